// kanlinear.cuh -- packed-parameter layout and the per-(row, input) feature evaluation shared by
// the any-width KANLinear kernels (CUDA-core forward/backward, tcgen05 forward).
#pragma once
#include <algorithm>
#include "smallnet.cuh"

namespace fetode {

// Packed image of one KANLinear layer ("wcat" in the C ABI):
//   wcat [I*F, O]   folded weights, row (i, f): f = 0 silu | 1..S spline s | S+1.. branch j
//   bias [O]        branch_bias + the input-independent part of the folded FERRO2D branch
//   tab  [I, TAB]   per input: padded knots kp[NKP], 1/(kp[j+p]-kp[j]) p=1..3, gi[4], bp[NB][4]
struct KLLayout {
    int I, O, G, NB, S, F, NK, NKP, TAB, tab_rows;
    int t_kn, t_rd1, t_rd2, t_rd3, t_gi, t_bp;
    long long wcat, bias, tab, total;
    // tensor-core weight image (kanlinear_tc.cu); tc_ok == 0 when the shape is served by the CUDA-core kernels only
    int tc_ok, tc_BN, tc_ntile, tc_units, tc_stages;
    long long tc;
    // input-gradient weight image (kanlinear_tc_dx.cu): [I/4 chunks][ceil(O/32) K slabs][hi|lo][4*td_FP x 32]
    int td_groups, td_KS, td_FP;
    long long td;
};

__host__ __device__ inline KLLayout kl_layout(int I, int O, int G, int NB) {
    KLLayout L{};
    L.I = I; L.O = O; L.G = G; L.NB = NB;
    L.S = G + 3;
    L.F = 1 + L.S + NB;
    L.NK = G + 7;
    L.NKP = round_up(L.NK + 6, 4);
    int t = 0;
    L.t_kn = t; t += L.NKP;
    L.t_rd1 = t; t += L.NKP;
    L.t_rd2 = t; t += L.NKP;
    L.t_rd3 = t; t += L.NKP;
    L.t_gi = t; t += 4;
    L.t_bp = t; t += 4 * NB;
    L.TAB = t;
    long long o = 0;
    L.wcat = o; o += ((long long)I * L.F * O + 3) / 4 * 4;
    L.bias = o; o += (O + 3) / 4 * 4;
    L.tc_ok = (G == 5 && (NB == 0 || NB == 8 || NB == 10) && (long long)I * L.F >= 1024 && O >= 8 && I % 4 == 0) ? 1 : 0;
    L.tab_rows = L.tc_ok ? round_up(I, 32) : I;          // the tcgen05 kernel streams whole 32-input slabs of the table
    L.tab = o; o += (long long)L.tab_rows * L.TAB;
    L.tc = 0;
    if (L.tc_ok) {
        L.tc_BN = O <= 16 ? 16 : (O <= 32 ? 32 : (O <= 64 ? 64 : 128));
        L.tc_ntile = (O + L.tc_BN - 1) / L.tc_BN;
        L.tc_units = (I + 31) / 32;
        L.tc_stages = L.tc_units * 2 * (5 + NB / 2);       // 16 input pairs x (5 + NB/2) 16-byte chunks, 8 chunks per stage
        o = (o + 255) / 256 * 256;                       // 1 KB alignment for the bulk copies
        L.tc = o;
        o += (long long)L.tc_ntile * L.tc_stages * 2 * L.tc_BN * 32;
        L.td_groups = I / 4;
        L.td_KS = (O + 31) / 32;
        L.td_FP = NB > 0 ? 20 : 12;
        o = (o + 255) / 256 * 256;
        L.td = o;
        o += (long long)L.td_groups * L.td_KS * 2 * (4 * L.td_FP) * 32;
    }
    L.total = o;
    return L;
}

// silu and the single input-dependent factor of each branch basis, see smallnet.cu header:
//   LOGISTIC: r_j = 1/(1+2^(-a(x-b) log2e))                 (phi = 2 r)
//   FERRO2D : r_j = 1/(1+2^(2 log2e k (x + Ec*momentum)))   (tanh = 1 - 2 r)
template <int BR>
__device__ __forceinline__ float kl_branch_r(float x, float omu, float ex, const float4 p) {
    if constexpr (BR == FETODE_BRANCH_FERRO2D) {
        // ex = ferro_ex(x, gl), p.x = 2^(c1_j): 1 + min(2^(x gl + c1_j), 2^30)   (smallnet.cuh, bq_c1)
        const float cn = rcp_approx(1.0f + fminf(ex * p.x, 1073741824.0f));
        const float sl = omu * cn;
        const float t = fmaf(p.w, sl, fmaf(p.y, x, p.z));
        return rcp_approx(1.0f + ex2_approx(fminf(t, FETODE_EXP_CLAMP)));
    } else {
        return rcp_approx(1.0f + ex2_approx(fminf(fmaf(x, p.x, p.y), FETODE_EXP_CLAMP)));
    }
}

// All F feature values of one (row, input) pair, written to col[f * stride].
template <int BR>
__device__ __forceinline__ void kl_features(float x, float prev, float gl, const float* __restrict__ tab,
                                            const KLLayout& L, float* col, int stride) {
    col[0] = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));                 // efficientkan.py:166
    const int S = L.S;
    for (int s = 0; s < S; ++s) col[(1 + s) * stride] = 0.f;
    const float4 gi = *reinterpret_cast<const float4*>(tab + L.t_gi);
    if (x >= gi.x && x < gi.z) {                                                     // :122 half-open test
        const float* kp = tab + L.t_kn;
        const int m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
        float v[4];
        bspline3_local(x, kp, tab + L.t_rd1, tab + L.t_rd2, tab + L.t_rd3, m + 3, v);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int s = m - 3 + c;
            if (s >= 0 && s < S) col[(1 + s) * stride] = v[c];
        }
    }
    if constexpr (BR != FETODE_BRANCH_NONE) {
        float omu = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D)
            omu = rcp_approx(1.0f + ex2_approx(gl * (x - prev)));                    // 1 - is_moving_up  ferro_class.py:560
        const float4* bp = reinterpret_cast<const float4*>(tab + L.t_bp);
        const float ex = BR == FETODE_BRANCH_FERRO2D ? ferro_ex(x, gl) : 0.f;
        for (int j = 0; j < L.NB; ++j) col[(1 + S + j) * stride] = kl_branch_r<BR>(x, omu, ex, bp[j]);
    }
}

int kl_check(const fetode_kanlayer_t* P);
int kl_tc_pack(const KLLayout& L, float* blob, cudaStream_t st);
int kl_tc_dw(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
             int prev_is_zero, const float* gy, float* ws, long long B, int accumulate, cudaStream_t st);
int kl_tc_dx(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
             int prev_is_zero, const float* gy, float* gx, float* gbp0, float* gbp1, long long B, cudaStream_t st);
int kl_td_pack(const KLLayout& L, float* blob, cudaStream_t st);
long long kl_tc_min_rows();
int kl_tc_fwd(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
              int prev_is_zero, float* y, long long B, cudaStream_t st);

}  // namespace fetode
