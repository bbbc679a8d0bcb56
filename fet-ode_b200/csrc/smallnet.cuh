// smallnet.cuh -- shared-memory image layout and the per-trajectory right-hand side of the fused small-width
// KAN / KANFET kernels (forward: smallnet.cu, reverse sweep: smallnet_bwd.cu).  See smallnet.cu for the
// reference lines being replaced and the folding identities.
#pragma once
#include "common.cuh"

namespace fetode {

constexpr int SN_THREADS = 128;

// Shared-memory image.  Everything whose size depends only on (D, NB, G) sits at a compile-time
// offset; everything per hidden unit h is gathered into one record of REC floats so the fused
// hidden-unit loop walks a single pointer with immediate offsets.
//   (l1_bq / r_bq / r_wv1 repeat l1_bp / r_bp / r_wu1 with basis j paired with j + NB/2: the forward kernels
//    evaluate the branch two bases at a time with packed FFMA2 / FMUL2 / FADD2; the reverse sweep reads the plain copy)
//   hdr[4]      {gate_slope*log2e layer 1, layer 2, -, -}
//   layer-1 input tables, per input d < D: padded knots kp[NKP], 1/(kp[j+p]-kp[j]) p=1..3,
//               gi[4] = {t_0, (NK-1)/(t_last-t_0), t_last, -}, branch params bp[NB][4]
//   b2[4]       layer-2 output bias (folded branch constants)
//   rec[H]:     b1[4] (layer-1 bias of output h) | wu1[D][FU] ([base | branch j] weights of
//               output h) | ws1[D][SPP] (spline coefficients, s stored at s+3) |
//               layer-2 tables of input h: kp, rd1, rd2, rd3, gi, bp | wu2[FU][DV] | ws2[SPP][DV]
struct SNLayout {
    int D, H, NB, NK, NKP, S, SPP, FU, DV;
    int hdr, l1_kn, l1_rd1, l1_rd2, l1_rd3, l1_gi, l1_bp, l1_bq, b2, rec0, REC;
    int r_b1, r_wu1, r_wv1, r_ws1, r_kn, r_rd1, r_rd2, r_rd3, r_gi, r_bp, r_bq, r_wu2, r_ws2;
    int total;          // floats the forward kernels stage
    int raw1, raw2, total_ext;   // raw branch params (k,Ec | a,b) per (input, basis): reverse sweep only
};

__host__ __device__ constexpr SNLayout sn_layout(int D, int H, int NB, int G) {
    SNLayout L{};
    L.D = D; L.H = H; L.NB = NB;
    L.NK = G + 7;                       // knots per input (spline_order 3)
    L.NKP = round_up(L.NK + 6, 4);      // + 3 extrapolated knots either side
    L.S = G + 3;                        // spline coefficients per input
    L.SPP = round_up(L.S + 6, 4);       // + 3 zero coefficients either side
    L.FU = round_up(1 + NB, 4);         // [silu | branch j] weights, padded
    L.DV = D <= 1 ? 1 : (D == 2 ? 2 : 4);
    int o = 0;
    L.hdr = o; o += 4;
    L.l1_kn = o; o += D * L.NKP;
    L.l1_rd1 = o; o += D * L.NKP;
    L.l1_rd2 = o; o += D * L.NKP;
    L.l1_rd3 = o; o += D * L.NKP;
    L.l1_gi = o; o += D * 4;
    L.l1_bp = o; o += D * NB * 4;
    L.l1_bq = o; o += D * NB * 4;       // the same constants, basis j paired with j + NB/2 (forward: FFMA2 lanes)
    L.b2 = o; o += 4;
    L.rec0 = o;
    int r = 0;
    L.r_b1 = r; r += 4;
    L.r_wu1 = r; r += D * L.FU;
    L.r_wv1 = r; r += D * L.FU;         // [ (w_p, w_{p+NB/2}) p < NB/2 | base | 0.. ] : pair order of r_wu1
    L.r_ws1 = r; r += D * L.SPP;
    L.r_kn = r; r += L.NKP;
    L.r_rd1 = r; r += L.NKP;
    L.r_rd2 = r; r += L.NKP;
    L.r_rd3 = r; r += L.NKP;
    L.r_gi = r; r += 4;
    L.r_bp = r; r += NB * 4;
    L.r_bq = r; r += NB * 4;
    L.r_wu2 = r; r += L.FU * L.DV;
    L.r_ws2 = r; r += L.SPP * L.DV;
    r = round_up(r, 4);
    if ((r / 4) % 2 == 0) r += 4;       // record stride = an odd number of 16-byte units: lanes that read DIFFERENT records (the
                                        // hidden-units-on-lanes kernels) spread over the shared-memory banks (<= 2-way conflicts)
    L.REC = r;
    L.total = o + H * r;
    L.raw1 = L.total;
    L.raw2 = L.raw1 + 2 * D * NB;
    L.total_ext = round_up(L.raw2 + 2 * H * NB, 4);
    return L;
}

// ------------------------------------------------------------------------------------------
// device: per-input features
// ------------------------------------------------------------------------------------------
// FERRO2D: crossed_neg_j = 1 / (1 + 2^(x gl + c1_j)) for the NB thresholds of one input shares its x-dependent factor,
// 2^(x gl + c1_j) = 2^(x gl) * 2^(c1_j): the pair-ordered image stores C_j = 2^(c1_j) and the kernel spends ONE ex2 per
// input instead of NB (the MUFU pipe is this kernel's busiest).  Both exponents are clamped to +-FETODE_EXP_SPLIT so each
// factor is finite and non-zero: the product may overflow to inf (-> min(.., 2^30), exactly what the clamped sum gives)
// or underflow to 0 (2^t ~ 0 there), never NaN.  min(2^t, 2^30) keeps rcp_batch2's operand range.
#define FETODE_EXP_SPLIT 126.0f
__device__ __forceinline__ float ferro_ex(float x, float gl) {
    return ex2_approx(fminf(fmaxf(x * gl, -FETODE_EXP_SPLIT), FETODE_EXP_SPLIT));
}
__device__ inline float bq_c1(const fetode_kanlayer_t& P, float c1) {
    if (P.branch != FETODE_BRANCH_FERRO2D) return c1;
    return exp2f(fminf(fmaxf(c1, -FETODE_EXP_SPLIT), FETODE_EXP_SPLIT));
}

// r[j] = 1/(1+2^t_j): the only input-dependent factor of branch basis j (see header comment)
// cn_out (FERRO2D, optional): the crossed_neg factors, for a backward pass that would otherwise recompute them
template <int NB, int BR>
__device__ __forceinline__ void branch_eval(float x, float omu, float gl, const float4* __restrict__ bp,
                                            float (&r)[NB > 0 ? NB : 1], float* cn_out = nullptr) {
    if constexpr (NB > 0 && BR == FETODE_BRANCH_FERRO2D) {
        float a[NB], cn[NB];
        const float ex = ferro_ex(x, gl);
#pragma unroll
        for (int j = 0; j < NB; ++j)
            a[j] = 1.0f + fminf(ex * bp[j].x, 1073741824.0f);   // bp.x = 2^(c1_j): 1 + min(2^(x gl + c1_j), 2^30)
        rcp_batch<NB>(a, cn);                                   // crossed_neg_Ec  ferro_class.py:563
        if (cn_out) {
#pragma unroll
            for (int j = 0; j < NB; ++j) cn_out[j] = cn[j];
        }
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const float4 p = bp[j];
            const float sl = omu * cn[j];                       // switch_to_lower :567
            const float t = fmaf(p.w, sl, fmaf(p.y, x, p.z));   // 2*log2e*k*(x + Ec*momentum) :572-575
            a[j] = 1.0f + ex2_approx(fminf(t, FETODE_EXP_CLAMP));
        }
        rcp_batch<NB>(a, r);                                    // tanh = 1 - 2 r           :578
    } else if constexpr (NB > 0 && BR == FETODE_BRANCH_LOGISTIC) {
        float a[NB];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const float4 p = bp[j];
            a[j] = 1.0f + ex2_approx(fminf(fmaf(x, p.x, p.y), FETODE_EXP_CLAMP));
        }
        rcp_batch<NB>(a, r);                                    // phi = 2 r  efficientkan.py:24
    }
}

struct SNArgs {
    const float* packed;
    const float* y0;
    float* y_out;
    float* state;        // [B, D+H] in/out or null
    float* traj;         // [n_dump, B, D] or null
    float* ckpt;         // [n_steps, D+D+H, B] or null
    long long B;
    float dt, hdt, dt6;
    int H, tab, n_steps, dump_every, state_is_zero;
};

constexpr int TAB_RHS_ONLY = -1;

// One right-hand-side evaluation out = KAN(y).  prev1 (registers) / prev2 (shared-memory column,
// stride SN_THREADS) hold the previous input of each layer and are advanced to the current one,
// exactly like the reference's prev_x buffer is by every forward call (ferro_class.py:584-590).
template <int D, int NB, int G, int BR>
__device__ __forceinline__ void sn_rhs_scalar(const float* __restrict__ sm, int H, const float (&y)[D],
                                              float (&prev1)[D], float* prev2, float (&out)[D]) {
    constexpr SNLayout L = sn_layout(D, 1, NB, G);
    constexpr int NBR = NB > 0 ? NB : 1;
    constexpr int DV = L.DV, FU = L.FU, NKP = L.NKP, SPP = L.SPP, NK = L.NK;
    const float gl1 = sm[L.hdr], gl2 = sm[L.hdr + 1];
    float silu1[D], v1[D][4], r1[D][NBR];
    int m1[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float x = y[d];
        silu1[d] = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));       // efficientkan.py:166
        float omu = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D) {
            omu = rcp_approx(1.0f + ex2_approx(gl1 * (x - prev1[d])));           // 1 - is_moving_up :560
            prev1[d] = x;
        }
        const float4 gi = reinterpret_cast<const float4*>(sm + L.l1_gi)[d];
        m1[d] = -1;
        v1[d][0] = v1[d][1] = v1[d][2] = v1[d][3] = 0.f;
        if (x >= gi.x && x < gi.z) {
            const float* kp = sm + L.l1_kn + d * NKP;
            const int m = bspline_interval(x, kp, gi.x, gi.y, NK);
            bspline3_local(x, kp, sm + L.l1_rd1 + d * NKP, sm + L.l1_rd2 + d * NKP,
                           sm + L.l1_rd3 + d * NKP, m + 3, v1[d]);
            m1[d] = m;
        }
        branch_eval<NB, BR>(x, omu, gl1, reinterpret_cast<const float4*>(sm + L.l1_bp) + d * NB, r1[d]);
    }
    {
        const float4 b2 = *reinterpret_cast<const float4*>(sm + L.b2);
        out[0] = b2.x;
        if constexpr (D > 1) out[1] = b2.y;
        if constexpr (D > 2) out[2] = b2.z;
        if constexpr (D > 3) out[3] = b2.w;
    }
    const float* rec = sm + L.rec0;
#pragma unroll 1
    for (int h = 0; h < H; ++h, rec += L.REC) {
        // ---- layer 1, output h
        float acc = rec[L.r_b1];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float4* w4 = reinterpret_cast<const float4*>(rec + L.r_wu1 + d * FU);
            float feat[FU];
            feat[0] = silu1[d];
#pragma unroll
            for (int j = 0; j < NB; ++j) feat[1 + j] = r1[d][j];
#pragma unroll
            for (int j = 1 + NB; j < FU; ++j) feat[j] = 0.f;
#pragma unroll
            for (int q = 0; q < FU / 4; ++q) {
                const float4 w = w4[q];
                acc = fmaf(w.x, feat[4 * q], acc);
                acc = fmaf(w.y, feat[4 * q + 1], acc);
                acc = fmaf(w.z, feat[4 * q + 2], acc);
                acc = fmaf(w.w, feat[4 * q + 3], acc);
            }
            if (m1[d] >= 0) {
                const float* ws = rec + L.r_ws1 + d * SPP + m1[d];
                acc = fmaf(ws[0], v1[d][0], acc);
                acc = fmaf(ws[1], v1[d][1], acc);
                acc = fmaf(ws[2], v1[d][2], acc);
                acc = fmaf(ws[3], v1[d][3], acc);
            }
        }
        // ---- layer 2, input h
        const float x = acc;
        const float sl = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
        float omu = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D) {
            omu = rcp_approx(1.0f + ex2_approx(gl2 * (x - prev2[h * SN_THREADS])));
            prev2[h * SN_THREADS] = x;
        }
        float r2[NBR];
        branch_eval<NB, BR>(x, omu, gl2, reinterpret_cast<const float4*>(rec + L.r_bp), r2);
        {
            const float* wu = rec + L.r_wu2;
            float feat[1 + NB];
            feat[0] = sl;
#pragma unroll
            for (int j = 0; j < NB; ++j) feat[1 + j] = r2[j];
#pragma unroll
            for (int f = 0; f < 1 + NB; ++f) {
                if constexpr (DV == 1) {
                    out[0] = fmaf(wu[f], feat[f], out[0]);
                } else if constexpr (DV == 2) {
                    const float2 w = reinterpret_cast<const float2*>(wu)[f];
                    out[0] = fmaf(w.x, feat[f], out[0]);
                    out[1] = fmaf(w.y, feat[f], out[1]);
                } else {
                    const float4 w = reinterpret_cast<const float4*>(wu)[f];
                    out[0] = fmaf(w.x, feat[f], out[0]);
                    out[1] = fmaf(w.y, feat[f], out[1]);
                    if constexpr (D > 2) out[2] = fmaf(w.z, feat[f], out[2]);
                    if constexpr (D > 3) out[3] = fmaf(w.w, feat[f], out[3]);
                }
            }
        }
        const float4 gi = *reinterpret_cast<const float4*>(rec + L.r_gi);
        if (x >= gi.x && x < gi.z) {
            const float* kp = rec + L.r_kn;
            const int m = bspline_interval(x, kp, gi.x, gi.y, NK);
            float v[4];
            bspline3_local(x, kp, rec + L.r_rd1, rec + L.r_rd2, rec + L.r_rd3, m + 3, v);
            const float* ws = rec + L.r_ws2 + m * DV;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if constexpr (DV == 1) {
                    out[0] = fmaf(ws[c], v[c], out[0]);
                } else if constexpr (DV == 2) {
                    const float2 w = reinterpret_cast<const float2*>(ws)[c];
                    out[0] = fmaf(w.x, v[c], out[0]);
                    out[1] = fmaf(w.y, v[c], out[1]);
                } else {
                    const float4 w = reinterpret_cast<const float4*>(ws)[c];
                    out[0] = fmaf(w.x, v[c], out[0]);
                    out[1] = fmaf(w.y, v[c], out[1]);
                    if constexpr (D > 2) out[2] = fmaf(w.z, v[c], out[2]);
                    if constexpr (D > 3) out[3] = fmaf(w.w, v[c], out[3]);
                }
            }
        }
    }
}

// The NB branch factors of one input, two bases per lane pair: R[p] = (r_p, r_{p + NB/2}).  bq: the pair-ordered
// constants (l1_bq / r_bq), two float4 per pair: (P_lo.x, P_hi.x, P_lo.y, P_hi.y), (P_lo.z, P_hi.z, P_lo.w, P_hi.w).
template <int NB, int BR>
__device__ __forceinline__ void branch_eval2(float x, float omu, float gl, const float4* __restrict__ bq,
                                             f32x2 (&R)[NB / 2 > 0 ? NB / 2 : 1]) {
    constexpr int NP = NB / 2;
    const f32x2 one2 = pk2(1.0f, 1.0f), x2 = pk2(x, x);
    f32x2 A[NP > 0 ? NP : 1];
    if constexpr (NB > 0 && BR == FETODE_BRANCH_FERRO2D) {
        f32x2 CN[NP];
        const float ex = ferro_ex(x, gl);
        const f32x2 ex2v = pk2(ex, ex), omu2 = pk2(omu, omu);
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            const float4 q0 = bq[2 * p];
            float t0, t1;
            upk2(mul2(ex2v, pk2(q0.x, q0.y)), t0, t1);          // 2^(x gl) * 2^(c1_j)
            A[p] = add2(pk2(fminf(t0, 1073741824.0f), fminf(t1, 1073741824.0f)), one2);
        }
        rcp_batch2<NP>(A, CN);                                  // crossed_neg_Ec  ferro_class.py:563
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            const float4 q0 = bq[2 * p], q1 = bq[2 * p + 1];
            const f32x2 sl = mul2(omu2, CN[p]);                 // switch_to_lower :567
            f32x2 t = fma2(pk2(q0.z, q0.w), x2, pk2(q1.x, q1.y));
            t = fma2(pk2(q1.z, q1.w), sl, t);                   // 2*log2e*k*(x + Ec*momentum) :572-575
            float t0, t1;
            upk2(t, t0, t1);
            A[p] = add2(pk2(ex2_approx(fminf(t0, FETODE_EXP_CLAMP)), ex2_approx(fminf(t1, FETODE_EXP_CLAMP))), one2);
        }
        rcp_batch2<NP>(A, R);                                   // tanh = 1 - 2 r           :578
    } else if constexpr (NB > 0 && BR == FETODE_BRANCH_LOGISTIC) {
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            const float4 q0 = bq[2 * p];
            float t0, t1;
            upk2(fma2(x2, pk2(q0.x, q0.y), pk2(q0.z, q0.w)), t0, t1);
            A[p] = add2(pk2(ex2_approx(fminf(t0, FETODE_EXP_CLAMP)), ex2_approx(fminf(t1, FETODE_EXP_CLAMP))), one2);
        }
        rcp_batch2<NP>(A, R);                                   // phi = 2 r  efficientkan.py:24
    }
}

// sn_rhs with the branch evaluation and the contractions on packed fp32 pairs (FFMA2): same arithmetic as
// sn_rhs_scalar up to the order of the partial sums.  D = 2..4 (DV = 2 or 4), even NB > 0.
template <int D, int NB, int G, int BR>
__device__ __forceinline__ void sn_rhs_packed(const float* __restrict__ sm, int H, const float (&y)[D],
                                              float (&prev1)[D], float* prev2, float (&out)[D]) {
    constexpr SNLayout L = sn_layout(D, 1, NB, G);
    constexpr int NP = NB / 2;
    constexpr int DV = L.DV, FU = L.FU, NKP = L.NKP, SPP = L.SPP, NK = L.NK;
    constexpr int NO = DV / 2;                                   // packed output pairs
    static_assert(NB > 0 && NB % 2 == 0 && (DV == 2 || DV == 4), "packed path: even NB, D in 2..4");
    const float gl1 = sm[L.hdr], gl2 = sm[L.hdr + 1];
    float silu1[D], v1[D][4];
    f32x2 r1[D][NP];
    int m1[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float x = y[d];
        silu1[d] = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));       // efficientkan.py:166
        float omu = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D) {
            omu = rcp_approx(1.0f + ex2_approx(gl1 * (x - prev1[d])));           // 1 - is_moving_up :560
            prev1[d] = x;
        }
        const float4 gi = reinterpret_cast<const float4*>(sm + L.l1_gi)[d];
        m1[d] = -1;
        v1[d][0] = v1[d][1] = v1[d][2] = v1[d][3] = 0.f;
        if (x >= gi.x && x < gi.z) {
            const float* kp = sm + L.l1_kn + d * NKP;
            const int m = bspline_interval(x, kp, gi.x, gi.y, NK);
            bspline3_local(x, kp, sm + L.l1_rd1 + d * NKP, sm + L.l1_rd2 + d * NKP,
                           sm + L.l1_rd3 + d * NKP, m + 3, v1[d]);
            m1[d] = m;
        }
        branch_eval2<NB, BR>(x, omu, gl1, reinterpret_cast<const float4*>(sm + L.l1_bq) + d * NB, r1[d]);
    }
    f32x2 out2[NO];
    {
        const float4 b2 = *reinterpret_cast<const float4*>(sm + L.b2);
        out2[0] = pk2(b2.x, b2.y);
        if constexpr (NO > 1) out2[1] = pk2(b2.z, b2.w);
    }
    const float* rec = sm + L.rec0;
#pragma unroll 1
    for (int h = 0; h < H; ++h, rec += L.REC) {
        // ---- layer 1, output h: two partial sums (bases p | p + NB/2) per input, joined at the end
        f32x2 acc2 = pk2(rec[L.r_b1], 0.f);
        float accs = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float4* w4 = reinterpret_cast<const float4*>(rec + L.r_wv1 + d * FU);
#pragma unroll
            for (int q = 0; q < (NP + 1) / 2; ++q) {
                const float4 w = w4[q];
                acc2 = fma2(pk2(w.x, w.y), r1[d][2 * q], acc2);
                if (2 * q + 1 < NP) acc2 = fma2(pk2(w.z, w.w), r1[d][2 * q + 1], acc2);
            }
            accs = fmaf(rec[L.r_wv1 + d * FU + 2 * NP], silu1[d], accs);
            if (m1[d] >= 0) {
                const float* ws = rec + L.r_ws1 + d * SPP + m1[d];
                accs = fmaf(ws[0], v1[d][0], accs);
                accs = fmaf(ws[1], v1[d][1], accs);
                accs = fmaf(ws[2], v1[d][2], accs);
                accs = fmaf(ws[3], v1[d][3], accs);
            }
        }
        float alo, ahi;
        upk2(acc2, alo, ahi);
        // ---- layer 2, input h
        const float x = (alo + ahi) + accs;
        const float sl = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
        float omu = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D) {
            omu = rcp_approx(1.0f + ex2_approx(gl2 * (x - prev2[h * SN_THREADS])));
            prev2[h * SN_THREADS] = x;
        }
        f32x2 r2[NP];
        branch_eval2<NB, BR>(x, omu, gl2, reinterpret_cast<const float4*>(rec + L.r_bq), r2);
        {
            const float* wu = rec + L.r_wu2;                      // [1 + NB][DV]: silu | basis j
            auto axpy = [&](int f, float feat) {
                const f32x2 ff = pk2(feat, feat);
                if constexpr (DV == 2) {
                    const float2 w = reinterpret_cast<const float2*>(wu)[f];
                    out2[0] = fma2(pk2(w.x, w.y), ff, out2[0]);
                } else {
                    const float4 w = reinterpret_cast<const float4*>(wu)[f];
                    out2[0] = fma2(pk2(w.x, w.y), ff, out2[0]);
                    out2[1] = fma2(pk2(w.z, w.w), ff, out2[1]);
                }
            };
            axpy(0, sl);
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                float rlo, rhi;
                upk2(r2[p], rlo, rhi);
                axpy(1 + p, rlo);
                axpy(1 + p + NP, rhi);
            }
        }
        const float4 gi = *reinterpret_cast<const float4*>(rec + L.r_gi);
        if (x >= gi.x && x < gi.z) {
            const float* kp = rec + L.r_kn;
            const int m = bspline_interval(x, kp, gi.x, gi.y, NK);
            float v[4];
            bspline3_local(x, kp, rec + L.r_rd1, rec + L.r_rd2, rec + L.r_rd3, m + 3, v);
            const float* ws = rec + L.r_ws2 + m * DV;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const f32x2 vv = pk2(v[c], v[c]);
                if constexpr (DV == 2) {
                    const float2 w = reinterpret_cast<const float2*>(ws)[c];
                    out2[0] = fma2(pk2(w.x, w.y), vv, out2[0]);
                } else {
                    const float4 w = reinterpret_cast<const float4*>(ws)[c];
                    out2[0] = fma2(pk2(w.x, w.y), vv, out2[0]);
                    out2[1] = fma2(pk2(w.z, w.w), vv, out2[1]);
                }
            }
        }
    }
    {
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        upk2(out2[0], o[0], o[1]);
        if constexpr (NO > 1) upk2(out2[1], o[2], o[3]);
#pragma unroll
        for (int d = 0; d < D; ++d) out[d] = o[d];
    }
}

template <int D, int NB, int G, int BR>
__device__ __forceinline__ void sn_rhs(const float* __restrict__ sm, int H, const float (&y)[D],
                                       float (&prev1)[D], float* prev2, float (&out)[D]) {
#ifndef FETODE_SN_SCALAR
    if constexpr (NB > 0 && NB % 2 == 0 && D >= 2) sn_rhs_packed<D, NB, G, BR>(sm, H, y, prev1, prev2, out);
    else
#endif
        sn_rhs_scalar<D, NB, G, BR>(sm, H, y, prev1, prev2, out);
}

// d(loss)/d(branch inputs) for one (input, basis j): given G = d(loss)/d r_j returns d(loss)/dx and the two
// raw-parameter contributions.  FERRO2D: SURVEY appendix A.4 with branch_sign == 1; LOGISTIC: A.3.
// HAVE_CN: cn_in holds crossed_neg from the forward evaluation of the same point (branch_eval's cn_out); otherwise it is
// recomputed from et = 2^(x gl) * 2^(c1_j), the caller's ferro_ex(x, gl) times the packed constant (one rcp).
template <int BR, bool HAVE_CN = false>
__device__ __forceinline__ void branch_bwd(float x, float omu, float gl, float al2, float et, float r, float G,
                                           float p0, float p1, float& gx, float& g0, float& g1, float cn_in = 0.f) {
    if constexpr (BR == FETODE_BRANCH_FERRO2D) {
        const float g = gl * (1.0f / FETODE_LOG2E);
        const float k = p0, Ec = p1;
        const float cn = HAVE_CN ? cn_in : rcp_approx(1.0f + fminf(et, 1073741824.0f));
        const float dcn = -g * cn * (1.0f - cn);                  // d/dx == d/dEc
        const float domu = -g * omu * (1.0f - omu);
        const float m = 1.0f - al2 * omu * cn;
        const float dm_dx = -al2 * (domu * cn + omu * dcn);
        const float dm_dE = -al2 * omu * dcn;
        const float e = G * (-2.0f) * r * (1.0f - r);             // d(loss)/dz,  z = k (x + Ec m)
        gx = e * k * fmaf(Ec, dm_dx, 1.0f);
        g0 = e * fmaf(Ec, m, x);
        g1 = e * k * fmaf(Ec, dm_dE, m);
    } else {
        const float a = p0, b = p1;
        const float e = G * r * (1.0f - r);
        gx = e * a;
        g0 = e * (x - b);
        g1 = -e * a;
    }
}

int sn_check(const fetode_smallnet_t* net);

}  // namespace fetode
