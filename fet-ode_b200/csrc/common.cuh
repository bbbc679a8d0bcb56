// common.cuh -- device helpers shared by the fetode kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/fetode.h"

#define FETODE_LOG2E 1.4426950408889634f

#define FETODE_CHECK_ARG(cond)            do { if (!(cond)) return FETODE_EINVAL; } while (0)
#define FETODE_CHECK_SHAPE(cond)          do { if (!(cond)) return FETODE_ESHAPE; } while (0)
#define FETODE_CHECK_SUPPORTED(cond)      do { if (!(cond)) return FETODE_EUNSUPPORTED; } while (0)
#define FETODE_LAUNCH_RC()                ((int)cudaGetLastError())

namespace fetode {

__host__ __device__ constexpr int round_up(int x, int m) { return (x + m - 1) / m * m; }

// MUFU wrappers.  ex2/rcp approximations are within 2 ulp / 1 ulp, well inside the 1e-5
// max-norm tolerance of the per-RHS parity gate; the bit-exact hard gate does NOT use them.
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Exponent clamp for values that enter a batched reciprocal: 2^30 per factor keeps a product
// of four below 2^127, and 1/(1+2^30) = 9.3e-10 is indistinguishable from 0 at fp32 output scale.
#define FETODE_EXP_CLAMP 30.0f

// r[i] = 1 / a[i] for N values with ceil(N/4) MUFU.RCP instead of N (a[i] in [1, 2^30+1]).
template <int N>
__device__ __forceinline__ void rcp_batch(const float (&a)[N], float (&r)[N]) {
    constexpr int Q = N / 4;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        const float a0 = a[4 * q], a1 = a[4 * q + 1], a2 = a[4 * q + 2], a3 = a[4 * q + 3];
        const float p01 = a0 * a1, p23 = a2 * a3;
        const float inv = rcp_approx(p01 * p23);
        const float q01 = inv * p23, q23 = inv * p01;
        r[4 * q] = q01 * a1;
        r[4 * q + 1] = q01 * a0;
        r[4 * q + 2] = q23 * a3;
        r[4 * q + 3] = q23 * a2;
    }
    constexpr int R = N - 4 * Q;
    if constexpr (R == 1) {
        r[4 * Q] = rcp_approx(a[4 * Q]);
    } else if constexpr (R == 2) {
        const float inv = rcp_approx(a[4 * Q] * a[4 * Q + 1]);
        r[4 * Q] = inv * a[4 * Q + 1];
        r[4 * Q + 1] = inv * a[4 * Q];
    } else if constexpr (R == 3) {
        const float p01 = a[4 * Q] * a[4 * Q + 1];
        const float inv = rcp_approx(p01 * a[4 * Q + 2]);
        const float q01 = inv * a[4 * Q + 2];
        r[4 * Q] = q01 * a[4 * Q + 1];
        r[4 * Q + 1] = q01 * a[4 * Q];
        r[4 * Q + 2] = inv * p01;
    }
}

// ---- packed fp32 pairs (Blackwell FFMA2 / FMUL2 / FADD2: one issue slot for two lanes) -------------------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 rcp2(f32x2 a) { float l, h; upk2(a, l, h); return pk2(rcp_approx(l), rcp_approx(h)); }

// r[i] = 1 / a[i], both halves of N packed pairs, with 2*ceil(N/4) MUFU.RCP (a[i] in [1, 2^30+1]).
template <int N>
__device__ __forceinline__ void rcp_batch2(const f32x2 (&a)[N], f32x2 (&r)[N]) {
    constexpr int Q = N / 4;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        const f32x2 p01 = mul2(a[4 * q], a[4 * q + 1]), p23 = mul2(a[4 * q + 2], a[4 * q + 3]);
        const f32x2 inv = rcp2(mul2(p01, p23));
        const f32x2 q01 = mul2(inv, p23), q23 = mul2(inv, p01);
        r[4 * q] = mul2(q01, a[4 * q + 1]);
        r[4 * q + 1] = mul2(q01, a[4 * q]);
        r[4 * q + 2] = mul2(q23, a[4 * q + 3]);
        r[4 * q + 3] = mul2(q23, a[4 * q + 2]);
    }
    constexpr int R = N - 4 * Q;
    if constexpr (R == 1) {
        r[4 * Q] = rcp2(a[4 * Q]);
    } else if constexpr (R == 2) {
        const f32x2 inv = rcp2(mul2(a[4 * Q], a[4 * Q + 1]));
        r[4 * Q] = mul2(inv, a[4 * Q + 1]);
        r[4 * Q + 1] = mul2(inv, a[4 * Q]);
    } else if constexpr (R == 3) {
        const f32x2 p01 = mul2(a[4 * Q], a[4 * Q + 1]);
        const f32x2 inv = rcp2(mul2(p01, a[4 * Q + 2]));
        const f32x2 q01 = mul2(inv, a[4 * Q + 2]);
        r[4 * Q] = mul2(q01, a[4 * Q + 1]);
        r[4 * Q + 1] = mul2(q01, a[4 * Q]);
        r[4 * Q + 2] = mul2(inv, p01);
    }
}

// sigmoid(z) with z already multiplied by log2(e):  1 / (1 + 2^-zl)
__device__ __forceinline__ float sigmoid_l2(float zl) { return rcp_approx(1.0f + ex2_approx(-zl)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Local cubic B-spline evaluation on a padded knot table (3 extrapolated knots either side,
// so kp[m+3] is the reference's t_m).  Given t_m <= x < t_{m+1} it returns the four values
// N^3_{m-3..m}(x) that the Cox-de Boor recursion of efficientkan.py:122-128 leaves non-zero;
// every other entry of the reference's [.., G+k] basis row is exactly 0.  rdp[j] holds
// 1 / (kp[j+p] - kp[j]).
__device__ __forceinline__ void bspline3_local(float x, const float* kp, const float* rd1,
                                               const float* rd2, const float* rd3, int mp,
                                               float (&v)[4]) {
    const float left1 = x - kp[mp], right1 = kp[mp + 1] - x;
    const float left2 = x - kp[mp - 1], right2 = kp[mp + 2] - x;
    const float left3 = x - kp[mp - 2], right3 = kp[mp + 3] - x;
    float t = rd1[mp];
    const float a0 = right1 * t, a1 = left1 * t;
    t = a0 * rd2[mp - 1];
    const float b0 = right1 * t;
    float saved = left2 * t;
    t = a1 * rd2[mp];
    const float b1 = saved + right2 * t;
    const float b2 = left1 * t;
    t = b0 * rd3[mp - 2];
    v[0] = right1 * t;
    saved = left3 * t;
    t = b1 * rd3[mp - 1];
    v[1] = saved + right2 * t;
    saved = left2 * t;
    t = b2 * rd3[mp];
    v[2] = saved + right3 * t;
    v[3] = left1 * t;
}

// Same evaluation plus the derivatives d/dx of the four values:
//   N'^3_j = 3 (N^2_j / (t_{j+3}-t_j) - N^2_{j+1} / (t_{j+4}-t_{j+1}))   
__device__ __forceinline__ void bspline3_local_d(float x, const float* kp, const float* rd1, const float* rd2,
                                                 const float* rd3, int mp, float (&v)[4], float (&dv)[4]) {
    const float left1 = x - kp[mp], right1 = kp[mp + 1] - x;
    const float left2 = x - kp[mp - 1], right2 = kp[mp + 2] - x;
    const float left3 = x - kp[mp - 2], right3 = kp[mp + 3] - x;
    float t = rd1[mp];
    const float a0 = right1 * t, a1 = left1 * t;
    t = a0 * rd2[mp - 1];
    const float b0 = right1 * t;
    float saved = left2 * t;
    t = a1 * rd2[mp];
    const float b1 = saved + right2 * t;
    const float b2 = left1 * t;
    const float t0 = b0 * rd3[mp - 2], t1 = b1 * rd3[mp - 1], t2 = b2 * rd3[mp];
    v[0] = right1 * t0;
    v[1] = left3 * t0 + right2 * t1;
    v[2] = left2 * t1 + right3 * t2;
    v[3] = left1 * t2;
    dv[0] = -3.0f * t0;
    dv[1] = 3.0f * (t0 - t1);
    dv[2] = 3.0f * (t1 - t2);
    dv[3] = 3.0f * t2;
}

// Column sums across a warp: every lane holds 32 values v[0..31]; returns in lane l the sum over all lanes of
// v[l].  31 shuffles for 32 columns (recursive halving) instead of 5 per column.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool hi = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const float send = hi ? v[i] : v[i + off];
            const float keep = hi ? v[i + off] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

// Interval search: largest m in [0, NK-2] with t_m <= x, for x known to lie in [t_0, t_{NK-1}).
// gi = {t_0, (NK-1)/(t_last - t_0), t_last, -}; the guess is exact for the default uniform
// grid and the two loops repair it for the non-uniform grids update_grid() produces
// (efficientkan.py:184-221).  Comparisons are against the stored fp32 knots, so membership is
// bit-identical to the reference's (x >= t_j) & (x < t_{j+1}) test (:122).
__device__ __forceinline__ int bspline_interval(float x, const float* kp, float t0, float inv_h,
                                                int NK) {
    int g = (int)((x - t0) * inv_h);
    g = min(max(g, 0), NK - 2);
    while (g > 0 && x < kp[g + 3]) --g;
    while (g < NK - 2 && x >= kp[g + 4]) ++g;
    return g;
}

// ---- parameter folding shared by the pack kernels (smallnet.cu, kanlinear.cu) -------------
// padded knots: kp[j+3] = t_j ; 3 knots extrapolated either side with the end spacings so that every
// read of the local evaluation is finite (those basis values meet zero coefficients)
__device__ inline void pack_knots_row(const float* t, int NK, int NKP, float* kp, float* rd1,
                                      float* rd2, float* rd3, float* gi) {
    for (int j = 0; j < NKP; ++j) {
        const int jj = j - 3;
        float v;
        if (jj < 0) v = t[0] + (float)jj * (t[1] - t[0]);
        else if (jj >= NK) v = t[NK - 1] + (float)(jj - NK + 1) * (t[NK - 1] - t[NK - 2]);
        else v = t[jj];
        kp[j] = v;
    }
    for (int j = 0; j < NKP; ++j) {
        rd1[j] = (j + 1 < NKP) ? 1.0f / (kp[j + 1] - kp[j]) : 0.f;
        rd2[j] = (j + 2 < NKP) ? 1.0f / (kp[j + 2] - kp[j]) : 0.f;
        rd3[j] = (j + 3 < NKP) ? 1.0f / (kp[j + 3] - kp[j]) : 0.f;
    }
    gi[0] = t[0];
    gi[1] = (float)(NK - 1) / (t[NK - 1] - t[0]);
    gi[2] = t[NK - 1];
    gi[3] = 0.f;
}

__device__ inline double branch_w(const fetode_kanlayer_t& P, int o, int i, int j) {
    double w = (double)P.branch_weight[(size_t)o * P.in_features * P.num_basis + i * P.num_basis + j] *
               (double)P.scale_logistic;
    if (P.branch_scaler) w *= (double)P.branch_scaler[o];
    return w;
}

// bp[j] = the four folded constants of branch basis (i, j)
__device__ inline float4 branch_params(const fetode_kanlayer_t& P, int idx) {
    const double l2e = 1.4426950408889634;
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (P.branch == FETODE_BRANCH_FERRO2D) {
        const double k = P.bp[0][idx], Ec = P.bp[1][idx];
        q.x = (float)((double)P.gate_slope * Ec * l2e);      // c1: cneg = 1/(1+2^(x*gl+c1))
        q.y = (float)(2.0 * l2e * k);                          // k'
        q.z = (float)(2.0 * l2e * k * Ec);                     // c2'
        q.w = (float)(-2.0 * (1.0 - (double)P.alpha) * 2.0 * l2e * k * Ec);   // c3
    } else if (P.branch == FETODE_BRANCH_LOGISTIC) {
        const double a = P.bp[0][idx], b = P.bp[1][idx];
        q.x = (float)(-a * l2e);
        q.y = (float)(a * b * l2e);
    }
    return q;
}

// multiplier of r = 1/(1+2^t) in the folded branch contraction, and its constant companion
__device__ inline void branch_fold(const fetode_kanlayer_t& P, int o, int i, int j, double& wr, double& wc) {
    const double w = branch_w(P, o, i, j);
    const int idx = i * P.num_basis + j;
    if (P.branch == FETODE_BRANCH_FERRO2D) {
        const double Ps = P.bp[2][idx], bias = P.bp[3][idx], coef = P.bp[4][idx];
        wr = -2.0 * w * coef * Ps;
        wc = w * coef * (Ps + bias);
    } else {
        wr = 2.0 * w;
        wc = 0.0;
    }
}

}  // namespace fetode
