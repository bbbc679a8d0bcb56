// kanlinear.cu -- KANLinear of any width on the FP32 CUDA-core pipe (sm_100a): forward.
//
// y[b,o] = sum_i  Wb[o,i] silu(x_bi) + sum_s Ws[o,i,s] sc[o,i] N_s(x_bi) + sum_j Wl'[o,i,j] phi_ij(x_bi)
// evaluated as ONE contraction  [B, I*F] x [I*F, O]  (F = 1 + (G+k) + nb) whose left operand is
// never written to HBM: each CTA builds the feature columns of a [64 rows x 4 inputs] slab in
// shared memory (the "prologue") and multiplies it against the matching slab of the folded weight
// matrix `wcat`.  This is the general path (any I, O, nb; O = 2 ... 672); the wide
// ECG / ETT shapes are routed to the tcgen05 kernel in kanlinear_tc.cu by fetode_kanlinear_fwd.
//
// Reference arithmetic being replaced (file:line relative to the reference checkout):
//   KANLinear.forward                       efficient_kan/efficientkan.py:160-182
//   b_splines                               efficient_kan/efficientkan.py:117-131
//   scaled_spline_weight / _logistic_weight efficient_kan/efficientkan.py:145-158
//   LogisticBasis.forward                   efficient_kan/efficientkan.py:20-24
//   TwoDimensionFerroelectricBasis.forward  ferro_class.py:547-592 (KANFET branch slot)
//   MNIST copy with logistic_bias           mnist_kuramoto_kan.py:112-142
#include <cstdlib>
#include "kanlinear.cuh"

namespace fetode {

// ------------------------------------------------------------------------------------------
// pack
// ------------------------------------------------------------------------------------------
__global__ void kl_pack_kernel(fetode_kanlayer_t P, KLLayout L, float* blob) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nt = (long long)gridDim.x * blockDim.x;
    const int I = L.I, O = L.O, F = L.F, S = L.S, NB = L.NB;
    // folded weights, rows (i, f), O contiguous
    const long long nW = (long long)I * F * O;
    for (long long idx = tid; idx < nW; idx += nt) {
        const int o = (int)(idx % O);
        const long long r = idx / O;
        const int f = (int)(r % F), i = (int)(r / F);
        float w;
        if (f == 0) {
            w = P.base_weight[(size_t)o * I + i];
        } else if (f <= S) {
            const float sc = P.spline_scaler ? P.spline_scaler[(size_t)o * I + i] : 1.0f;
            w = P.spline_weight[((size_t)o * I + i) * S + (f - 1)] * sc;
        } else {
            double wr, wc;
            branch_fold(P, o, i, f - 1 - S, wr, wc);
            w = (float)wr;
        }
        blob[L.wcat + idx] = w;
    }
    // per-output bias: branch_bias + input-independent part of the folded branch
    for (long long o = tid; o < O; o += nt) {
        double b = P.branch_bias ? (double)P.branch_bias[o] : 0.0;
        if (P.branch == FETODE_BRANCH_FERRO2D)
            for (int i = 0; i < I; ++i)
                for (int j = 0; j < NB; ++j) { double wr, wc; branch_fold(P, (int)o, i, j, wr, wc); b += wc; }
        blob[L.bias + o] = (float)b;
    }
    // per-input tables
    for (long long i = tid; i < L.tab_rows; i += nt) {
        float* t = blob + L.tab + i * L.TAB;
        if (i >= I) {                                   // padding rows: finite features, never inside a grid
            for (int q = 0; q < L.TAB; ++q) t[q] = 0.f;
            continue;
        }
        pack_knots_row(P.grid + (size_t)i * L.NK, L.NK, L.NKP, t + L.t_kn, t + L.t_rd1, t + L.t_rd2, t + L.t_rd3,
                       t + L.t_gi);
        for (int j = 0; j < NB; ++j) {
            float4 q = branch_params(P, (int)i * NB + j);
            q.x = bq_c1(P, q.x);                            // FERRO2D: 2^(c1) (smallnet.cuh, bq_c1)
            reinterpret_cast<float4*>(t + L.t_bp)[j] = q;
        }
    }
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
constexpr int KL_BM = 64, KL_KI = 4, KL_THREADS = 256;

struct KLFwdArgs {
    const float* blob;
    const float* x;
    const float* prev;
    float* prev_out;
    float* y;
    long long B;
    KLLayout L;
    float gl;
    int prev_is_zero;
};

template <int BR, int BN>
__global__ void __launch_bounds__(KL_THREADS) kl_fwd_kernel(const KLFwdArgs a) {
    extern __shared__ float4 kl_smem4[];
    float* As = reinterpret_cast<float*>(kl_smem4);
    const KLLayout& L = a.L;
    const int F = L.F, I = L.I, O = L.O;
    const int KC = KL_KI * F;
    float* Ws = As + KC * KL_BM;
    constexpr int TN = BN / 16;
    const int tid = threadIdx.x;
    const long long row0 = (long long)blockIdx.x * KL_BM;
    const int n0 = blockIdx.y * BN;
    const int frow = tid % KL_BM, fii = tid / KL_BM;
    const int ty = tid / 16, tx = tid % 16;
    const float* wcat = a.blob + L.wcat;

    float acc[4][TN];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < TN; ++c) acc[r][c] = 0.f;

    for (int i0 = 0; i0 < I; i0 += KL_KI) {
        // ---- prologue: feature columns of (row frow, input i0+fii)
        {
            const int i = i0 + fii;
            const long long b = row0 + frow;
            float* col = As + (fii * F) * KL_BM + frow;
            if (i < I && b < a.B) {
                const float x = a.x[b * I + i];
                float prev = 0.f;
                if constexpr (BR == FETODE_BRANCH_FERRO2D) {
                    if (!a.prev_is_zero) prev = a.prev[b * I + i];
                    if (a.prev_out && blockIdx.y == 0) a.prev_out[b * I + i] = x;
                }
                kl_features<BR>(x, prev, a.gl, a.blob + L.tab + (size_t)i * L.TAB, L, col, KL_BM);
            } else {
                for (int f = 0; f < F; ++f) col[f * KL_BM] = 0.f;
            }
        }
        // ---- weight slab [KC, BN]
        {
            const long long grow0 = (long long)i0 * F, grows = (long long)I * F;
            for (int idx = tid; idx < KC * BN; idx += KL_THREADS) {
                const int kk = idx / BN, n = idx % BN;
                float w = 0.f;
                if (grow0 + kk < grows && n0 + n < O) w = wcat[(grow0 + kk) * O + n0 + n];
                Ws[kk * BN + n] = w;
            }
        }
        __syncthreads();
#pragma unroll 4
        for (int kk = 0; kk < KC; ++kk) {
            const float4 av = *reinterpret_cast<const float4*>(As + kk * KL_BM + ty * 4);
            float bv[TN];
            if constexpr (TN == 4) {
                const float4 t = *reinterpret_cast<const float4*>(Ws + kk * BN + tx * 4);
                bv[0] = t.x; bv[1] = t.y; bv[2] = t.z; bv[3] = t.w;
            } else {
#pragma unroll
                for (int c = 0; c < TN; ++c) bv[c] = Ws[kk * BN + tx * TN + c];
            }
            const float ar[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < TN; ++c) acc[r][c] = fmaf(ar[r], bv[c], acc[r][c]);
        }
        __syncthreads();
    }
    const float* bias = a.blob + L.bias;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const long long b = row0 + ty * 4 + r;
        if (b >= a.B) continue;
#pragma unroll
        for (int c = 0; c < TN; ++c) {
            const int n = n0 + tx * TN + c;
            if (n < O) a.y[b * O + n] = acc[r][c] + bias[n];
        }
    }
}

template <int BR>
static int kl_fwd_launch(const KLFwdArgs& a, cudaStream_t st) {
    const KLLayout& L = a.L;
    const unsigned gx = (unsigned)((a.B + KL_BM - 1) / KL_BM);
    if (L.O <= 16) {
        const size_t smem = (size_t)KL_KI * L.F * (KL_BM + 16) * 4;
        auto k = kl_fwd_kernel<BR, 16>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<dim3(gx, 1), KL_THREADS, smem, st>>>(a);
    } else {
        const size_t smem = (size_t)KL_KI * L.F * (KL_BM + 64) * 4;
        auto k = kl_fwd_kernel<BR, 64>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<dim3(gx, (L.O + 63) / 64), KL_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

// rows from which the tensor-core kernel is used.  Mode: -1 auto (>= 64 rows), 0 never, 1 always; set with
// fetode_set_tc_mode(), initialised ONCE from FETODE_TC (no getenv on the per-call path).
static int g_tc_mode = -2;
long long kl_tc_min_rows() {
    if (g_tc_mode == -2) {
        const char* e = getenv("FETODE_TC");
        g_tc_mode = (e && e[0] == '0') ? 0 : ((e && e[0] == '1') ? 1 : -1);
    }
    if (g_tc_mode == 0) return (1LL << 62);
    if (g_tc_mode == 1) return 1;
    return 64;
}
int kl_set_tc_mode(int mode) {
    if (mode < -1 || mode > 1) return FETODE_EINVAL;
    g_tc_mode = mode;
    return FETODE_OK;
}

int kl_check(const fetode_kanlayer_t* P) {
    if (!P) return FETODE_EINVAL;
    if (P->in_features < 1 || P->out_features < 1) return FETODE_EINVAL;
    if (P->spline_order != 3) return FETODE_EUNSUPPORTED;
    if (P->grid_size < 1 || P->grid_size > 32) return FETODE_EUNSUPPORTED;
    switch (P->branch) {
        case FETODE_BRANCH_NONE: if (P->num_basis != 0) return FETODE_EINVAL; break;
        case FETODE_BRANCH_LOGISTIC:
        case FETODE_BRANCH_FERRO2D:
            if (P->num_basis < 1 || P->num_basis > 64) return FETODE_EUNSUPPORTED;
            break;
        default: return FETODE_EINVAL;
    }
    return FETODE_OK;
}

}  // namespace fetode

using namespace fetode;

extern "C" int64_t fetode_kanlinear_wcat_floats(const fetode_kanlayer_t* P) {
    if (kl_check(P) != FETODE_OK) return -1;
    return kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis).total;
}

extern "C" int fetode_kanlinear_pack(const fetode_kanlayer_t* P, float* wcat, void* stream) {
    int rc = kl_check(P);
    if (rc) return rc;
    FETODE_CHECK_ARG(wcat && P->grid && P->base_weight && P->spline_weight);
    if (P->branch != FETODE_BRANCH_NONE) {
        FETODE_CHECK_ARG(P->branch_weight && P->bp[0] && P->bp[1]);
        if (P->branch == FETODE_BRANCH_FERRO2D) FETODE_CHECK_ARG(P->bp[2] && P->bp[3] && P->bp[4]);
    }
    const KLLayout L = kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis);
    const long long work = (long long)L.I * L.F * L.O;
    const int blocks = (int)std::min<long long>((work + 255) / 256, 148 * 8);
    kl_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(*P, L, wcat);
    rc = FETODE_LAUNCH_RC();
    if (rc) return rc;
    rc = kl_tc_pack(L, wcat, (cudaStream_t)stream);
    if (rc) return rc;
    return kl_td_pack(L, wcat, (cudaStream_t)stream);
}

extern "C" int fetode_kanlinear_fwd(const fetode_kanlayer_t* P, const float* wcat, const float* x,
                                    const float* prev, int prev_is_zero, float* prev_out, float* y,
                                    int64_t B, void* stream) {
    int rc = kl_check(P);
    if (rc) return rc;
    FETODE_CHECK_ARG(wcat && x && y && B >= 0);
    if (B == 0) return FETODE_OK;
    if (P->branch == FETODE_BRANCH_FERRO2D) {
        if (!prev_is_zero) FETODE_CHECK_ARG(prev != nullptr);
        FETODE_CHECK_ARG(prev_out == nullptr || prev_out != prev);
    }
    KLFwdArgs a;
    a.blob = wcat; a.x = x; a.prev = prev; a.prev_out = prev_out; a.y = y; a.B = B;
    a.L = kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis);
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero;
    cudaStream_t st = (cudaStream_t)stream;
    if (a.L.tc_ok && B >= kl_tc_min_rows()) {           // wide layer: tcgen05 path (kanlinear_tc.cu)
        rc = kl_tc_fwd(P, a.L, wcat, x, prev, prev_is_zero, y, B, st);
        if (rc != FETODE_EUNSUPPORTED) {
            if (rc == FETODE_OK && prev_out && P->branch == FETODE_BRANCH_FERRO2D)
                rc = (int)cudaMemcpyAsync(prev_out, x, (size_t)B * P->in_features * sizeof(float), cudaMemcpyDeviceToDevice, st);
            return rc;
        }
    }
    switch (P->branch) {
        case FETODE_BRANCH_NONE: return kl_fwd_launch<FETODE_BRANCH_NONE>(a, st);
        case FETODE_BRANCH_LOGISTIC: return kl_fwd_launch<FETODE_BRANCH_LOGISTIC>(a, st);
        case FETODE_BRANCH_FERRO2D: return kl_fwd_launch<FETODE_BRANCH_FERRO2D>(a, st);
    }
    return FETODE_EUNSUPPORTED;
}

// ------------------------------------------------------------------------------------------
// b_splines alone (efficientkan.py:117-131): out[b,i,:] = the G+3 cubic basis values.  Works from
// the raw knot table (no pack): interval by scan, local de Boor with true divisions.
// ------------------------------------------------------------------------------------------
namespace fetode {
__global__ void __launch_bounds__(256) bspline_basis_kernel(const float* __restrict__ grid, const float* __restrict__ x,
                                                            float* __restrict__ out, long long n, int I, int G) {
    const int NK = G + 7, S = G + 3;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % I);
        const float* t = grid + (size_t)i * NK;
        const float xv = x[e];
        float* o = out + e * S;
        for (int s = 0; s < S; ++s) o[s] = 0.f;
        if (!(xv >= t[0] && xv < t[NK - 1])) continue;
        int m = 0;
        while (m < NK - 2 && xv >= t[m + 1]) ++m;                   // t_m <= x < t_{m+1}
        auto T = [&](int j) -> float {                                // knots beyond the table: extrapolated, weights 0
            if (j < 0) return t[0] + (float)j * (t[1] - t[0]);
            if (j >= NK) return t[NK - 1] + (float)(j - NK + 1) * (t[NK - 1] - t[NK - 2]);
            return t[j];
        };
        const float left1 = xv - T(m), right1 = T(m + 1) - xv;
        const float left2 = xv - T(m - 1), right2 = T(m + 2) - xv;
        const float left3 = xv - T(m - 2), right3 = T(m + 3) - xv;
        float q = 1.0f / (T(m + 1) - T(m));
        const float a0 = right1 * q, a1 = left1 * q;
        q = a0 / (T(m + 1) - T(m - 1));
        const float b0 = right1 * q;
        float saved = left2 * q;
        q = a1 / (T(m + 2) - T(m));
        const float b1 = saved + right2 * q, b2 = left1 * q;
        float v[4];
        q = b0 / (T(m + 1) - T(m - 2));
        v[0] = right1 * q; saved = left3 * q;
        q = b1 / (T(m + 2) - T(m - 1));
        v[1] = saved + right2 * q; saved = left2 * q;
        q = b2 / (T(m + 3) - T(m));
        v[2] = saved + right3 * q; v[3] = left1 * q;
        for (int c = 0; c < 4; ++c) {
            const int s = m - 3 + c;
            if (s >= 0 && s < S) o[s] = v[c];
        }
    }
}
}  // namespace fetode

extern "C" int fetode_bspline_basis(const float* grid, const float* x, float* out, int64_t B, int32_t I,
                                    int32_t grid_size, int32_t spline_order, void* stream) {
    FETODE_CHECK_ARG(grid && x && out && B >= 0 && I >= 1 && grid_size >= 1);
    FETODE_CHECK_SUPPORTED(spline_order == 3);
    if (B == 0) return FETODE_OK;
    const long long n = (long long)B * I;
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 16);
    fetode::bspline_basis_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(grid, x, out, n, I, grid_size);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_set_tc_mode(int mode) { return fetode::kl_set_tc_mode(mode); }
