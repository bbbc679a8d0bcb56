// smallnet_wide.cu -- the fused KAN / KANFET([2,10,2]) Runge-Kutta loop and its reverse sweep with the HIDDEN UNITS on
// the lanes (sm_100a): the training path.
//
// The lane-per-trajectory kernels (smallnet.cu / smallnet_v2.cu / smallnet_bwd.cu) are built for throughput over
// 10^5..10^6 rows.  Training batches are small (the reference's own Lotka-Volterra run is ONE trajectory,
// train_kanfet_node_predprey.py:149-155) and the reverse sweep must also reduce ~1000 parameter-gradient columns over
// the lanes after every right-hand side (two thirds of its instructions).  Here a warp holds three trajectories, ten
// lanes each: lane (g, h) owns hidden unit h of trajectory g -- its layer-1 output, its layer-2 input features, its
// rows of both weight matrices.  Consequences:
//   * the dependency chain of a right-hand side is 3 feature evaluations instead of 12 (B = 1: 8x shorter);
//   * every parameter-gradient column of unit h is produced by ONE lane: it accumulates privately (shared memory,
//     [column][lane]) over all stages, steps and trajectories the warp processes, and the cross-lane / cross-CTA
//     reduction happens once, at the end of the launch -- no per-right-hand-side reduction at all;
//   * the forward pass checkpoints every stage's input and hidden pre-activations ([step][slot][row][12] floats, one
//     coalesced 48-byte record per trajectory and slot), so the reverse sweep recomputes nothing but features.
// Lane exchanges are warp shuffles / a warp-private shared-memory transpose; there is no block-level synchronisation.
//
// Reference arithmetic replaced (reference checkout 704a0df4): as smallnet.cu (KAN.forward efficientkan.py:160-182,
// 274-279; TwoDimensionFerroelectricBasis ferro_class.py:547-592; integrate_rk4 train_ecg_kan_fet_nn_ode.py:693-705)
// and loss.backward() through the solver (train_kanfet_node_predprey.py:254-256).  Image: the first-generation layout
// (smallnet.cuh); folded-gradient layout: sn_gfold_layout (smallnet_bwd.cu), unfolded by fetode_smallnet_unfold.
#include <algorithm>
#include <cstdlib>
#include "smallnet.cuh"

namespace fetode {

constexpr int SW_D = 2, SW_H = 10, SW_NB = 10, SW_G = 5, SW_GRP = 3;      // 3 trajectories x 10 lanes per warp
constexpr int SW_WARPS = 4, SW_THREADS = SW_WARPS * 32;
constexpr int SW_REC = SW_D + SW_H;                                        // checkpoint record: y[2] | x2[10]
constexpr int SW_SLOTS = 5;                                                // slot 0: (prev1, prev2) at step start; 1+s: stage s

struct SNGfoldW { int ws1, gsum1, ws2, gsum2, bp1, bp2, total; };
__host__ __device__ inline SNGfoldW snw_gfold(int D, int H, int F, int NB) {
    SNGfoldW g{};
    int o = 0;
    g.ws1 = o; o += D * F * H;
    g.gsum1 = o; o += H;
    g.ws2 = o; o += H * F * D;
    g.gsum2 = o; o += D;
    g.bp1 = o; o += 2 * D * NB;
    g.bp2 = o; o += 2 * H * NB;
    g.total = (o + 3) / 4 * 4;
    return g;
}

struct SWArgs {
    const float* packed;
    const float* y0;          // [B, D]
    float* y_out;             // [B, D]
    float* state;             // [B, D + H] in/out or null
    float* traj;              // [n_dump, B, D] or null
    float* ckpt;              // [n_steps, SW_SLOTS, B, SW_REC] or null (forward: written; backward: read)
    const float* g_yout;      // backward
    const float* g_traj;
    float* g_y0;
    float* gfold;
    long long B;
    float dt;
    float A[4][4], bw[4];     // stage s input = y + dt sum_r A[s][r] k_r ; y_next = y + dt sum_s bw[s] k_s
    int nst, n_steps, dump_every, state_is_zero;
};

// sum over the ten lanes of a group in four shuffles; every lane of the group gets the total, BIT-IDENTICALLY: the tree
// ((0+1)+(2+3)) + ((4+5)+(6+7)) and (8+9) is built from exchanges between partners (a + b == b + a), so the ten replicas
// of a trajectory never drift apart and a solve resumed from saved state reproduces an uninterrupted one exactly
__device__ __forceinline__ float grp_sum(float v, int gbase, int h, bool lane_ok) {
    const bool low = h < 8;
    auto from = [&](float x, int hh) { return __shfl_sync(0xffffffffu, x, lane_ok ? gbase + hh : gbase + h); };
    const float s1 = v + from(v, h ^ 1);
    const float t2 = from(s1, low ? (h ^ 2) : h);
    const float s2 = low ? s1 + t2 : s1;
    const float t4 = from(s2, low ? (h ^ 4) : h);
    const float s4 = low ? s2 + t4 : s2;
    return s4 + from(s4, low ? 8 : 0);
}

// per-lane layer-1 work of one input d: silu, 1 - moving_up, and the branch factor r of basis j = h
template <int BR>
__device__ __forceinline__ void sw_l1_own(const float* __restrict__ sm, const SNLayout& L, int d, int h, float x, float prev,
                                          float gl1, float& silu, float& sg, float& omu, float& ex, float& r) {
    sg = rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
    silu = x * sg;
    const float4 bp = reinterpret_cast<const float4*>(sm + L.l1_bp)[d * SW_NB + h];
    omu = 0.f;
    ex = 0.f;
    if constexpr (BR == FETODE_BRANCH_FERRO2D) {
        omu = rcp_approx(1.0f + ex2_approx(gl1 * (x - prev)));
        ex = ferro_ex(x, gl1);
        const float cn = rcp_approx(1.0f + fminf(ex * bp.x, 1073741824.0f));
        const float t = fmaf(bp.w, omu * cn, fmaf(bp.y, x, bp.z));
        r = rcp_approx(1.0f + ex2_approx(fminf(t, FETODE_EXP_CLAMP)));
    } else {
        r = rcp_approx(1.0f + ex2_approx(fminf(fmaf(x, bp.x, bp.y), FETODE_EXP_CLAMP)));
    }
}

// One right-hand side for lane (g, h).  y / prev1 are replicated over the group; prev2 is this lane's.  Returns out[2]
// (replicated) and leaves this lane's hidden pre-activation in x2.
template <int BR>
__device__ __forceinline__ void sw_rhs(const float* __restrict__ sm, const SNLayout& L, int gbase, int h, bool lane_ok,
                                       const float (&y)[SW_D], float (&prev1)[SW_D], float& prev2, float (&out)[SW_D],
                                       float& x2_out) {
    constexpr int NB = SW_NB;
    const int FU = L.FU;
    const float gl1 = sm[L.hdr], gl2 = sm[L.hdr + 1];
    const float* rec = sm + L.rec0 + h * L.REC;
    float acc = rec[L.r_b1];
#pragma unroll
    for (int d = 0; d < SW_D; ++d) {
        const float x = y[d];
        float silu, sg, omu, ex, r;
        sw_l1_own<BR>(sm, L, d, h, x, prev1[d], gl1, silu, sg, omu, ex, r);
        prev1[d] = x;
        const float* wu = rec + L.r_wu1 + d * FU;
        acc = fmaf(wu[0], silu, acc);
#pragma unroll
        for (int j = 0; j < NB; ++j) acc = fmaf(wu[1 + j], __shfl_sync(0xffffffffu, r, lane_ok ? gbase + j : gbase + h), acc);
        const float4 gi = reinterpret_cast<const float4*>(sm + L.l1_gi)[d];
        if (x >= gi.x && x < gi.z) {                      // uniform over the group (x is replicated)
            const float* kp = sm + L.l1_kn + d * L.NKP;
            const int m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
            float v[4];
            bspline3_local(x, kp, sm + L.l1_rd1 + d * L.NKP, sm + L.l1_rd2 + d * L.NKP, sm + L.l1_rd3 + d * L.NKP, m + 3, v);
            const float* ws = rec + L.r_ws1 + d * L.SPP + m;
#pragma unroll
            for (int c = 0; c < 4; ++c) acc = fmaf(ws[c], v[c], acc);
        }
    }
    // ---- layer 2, input h
    const float x = acc;
    x2_out = x;
    const float sl = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
    float omu = 0.f;
    if constexpr (BR == FETODE_BRANCH_FERRO2D) omu = rcp_approx(1.0f + ex2_approx(gl2 * (x - prev2)));
    prev2 = x;
    float r2[NB];
    branch_eval<NB, BR>(x, omu, gl2, reinterpret_cast<const float4*>(rec + L.r_bp), r2);
    const float2* wu2 = reinterpret_cast<const float2*>(rec + L.r_wu2);
    float p0 = wu2[0].x * sl, p1 = wu2[0].y * sl;
#pragma unroll
    for (int j = 0; j < NB; ++j) { const float2 w = wu2[1 + j]; p0 = fmaf(w.x, r2[j], p0); p1 = fmaf(w.y, r2[j], p1); }
    const float4 gi = *reinterpret_cast<const float4*>(rec + L.r_gi);
    if (x >= gi.x && x < gi.z) {
        const float* kp = rec + L.r_kn;
        const int m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
        float v[4];
        bspline3_local(x, kp, rec + L.r_rd1, rec + L.r_rd2, rec + L.r_rd3, m + 3, v);
        const float2* ws = reinterpret_cast<const float2*>(rec + L.r_ws2) + m;
#pragma unroll
        for (int c = 0; c < 4; ++c) { p0 = fmaf(ws[c].x, v[c], p0); p1 = fmaf(ws[c].y, v[c], p1); }
    }
    const float4 b2 = *reinterpret_cast<const float4*>(sm + L.b2);
    out[0] = grp_sum(p0, gbase, h, lane_ok) + b2.x;
    out[1] = grp_sum(p1, gbase, h, lane_ok) + b2.y;
}

template <int BR>
__global__ void __launch_bounds__(SW_THREADS) smallnet_wide_fwd_kernel(const SWArgs a) {
    extern __shared__ float4 sw_smem4[];
    float* sm = reinterpret_cast<float*>(sw_smem4);
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    constexpr int SF = FERRO ? SW_D + SW_H : 0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < L.total / 4; i += SW_THREADS) sw_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    __syncthreads();
    const bool lane_ok = lane < SW_GRP * SW_H;
    const int g = lane_ok ? lane / SW_H : 0, h = lane_ok ? lane % SW_H : 0, gbase = lane_ok ? g * SW_H : lane;
    const long long n_tri = (a.B + SW_GRP - 1) / SW_GRP;
    for (long long tri = (long long)blockIdx.x * SW_WARPS + warp; tri < n_tri; tri += (long long)gridDim.x * SW_WARPS) {
        const long long t = tri * SW_GRP + g;
        const bool live = lane_ok && t < a.B;
        const long long tt = live ? t : 0;
        float y[SW_D], prev1[SW_D], prev2 = 0.f;
#pragma unroll
        for (int d = 0; d < SW_D; ++d) { y[d] = a.y0[tt * SW_D + d]; prev1[d] = 0.f; }
        if (FERRO && !a.state_is_zero) {
#pragma unroll
            for (int d = 0; d < SW_D; ++d) prev1[d] = a.state[tt * SF + d];
            prev2 = a.state[tt * SF + SW_D + h];
        }
        for (int step = 0; step < a.n_steps; ++step) {
            float* ck = a.ckpt ? a.ckpt + (((size_t)step * SW_SLOTS) * a.B + tt) * SW_REC : nullptr;
            const size_t slot_stride = (size_t)a.B * SW_REC;
            if (ck && live) {
                if (h < SW_D) ck[h] = prev1[h];
                ck[SW_D + h] = prev2;
            }
            float ks[4][SW_D], ys[SW_D];
#pragma unroll
            for (int d = 0; d < SW_D; ++d) ys[d] = y[d];
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                if (s < a.nst) {
                    float x2;
                    sw_rhs<BR>(sm, L, gbase, h, lane_ok, ys, prev1, prev2, ks[s], x2);
                    if (ck && live) {
                        float* c = ck + (size_t)(1 + s) * slot_stride;
                        if (h < SW_D) c[h] = ys[h];
                        c[SW_D + h] = x2;
                    }
                    if (s + 1 < a.nst) {
#pragma unroll
                        for (int d = 0; d < SW_D; ++d) {
                            float acc = 0.f;
#pragma unroll
                            for (int r = 0; r <= s; ++r) acc = fmaf(a.A[s + 1][r], ks[r][d], acc);
                            ys[d] = fmaf(a.dt, acc, y[d]);
                        }
                    }
                }
            }
#pragma unroll
            for (int d = 0; d < SW_D; ++d) {
                float acc = 0.f;
#pragma unroll
                for (int s = 0; s < 4; ++s)
                    if (s < a.nst) acc = fmaf(a.bw[s], ks[s][d], acc);
                y[d] = fmaf(a.dt, acc, y[d]);
            }
            if (a.traj && (step + 1) % a.dump_every == 0 && live && h < SW_D)
                a.traj[((size_t)((step + 1) / a.dump_every - 1) * a.B + tt) * SW_D + h] = y[h];
        }
        if (live) {
            if (h < SW_D) a.y_out[tt * SW_D + h] = y[h];
            if (FERRO && a.state) {
                if (h < SW_D) a.state[tt * SF + h] = prev1[h];
                a.state[tt * SF + SW_D + h] = prev2;
            }
        }
    }
}


// ------------------------------------------------------------------------------------------
// reverse sweep
// ------------------------------------------------------------------------------------------
// private accumulator columns of lane (g, h), F = 1 + 8 + NB features (0 = silu, 1 + s = spline s, 9 + j = branch j)
constexpr int SW_F = 1 + 8 + SW_NB;
constexpr int SW_C_W2 = 0;                          // [f][o]
constexpr int SW_C_BP2 = SW_C_W2 + SW_F * SW_D;      // c0[j] | c1[j]
constexpr int SW_C_W1 = SW_C_BP2 + 2 * SW_NB;        // [d][f]
constexpr int SW_C_GS1 = SW_C_W1 + SW_D * SW_F;
constexpr int SW_C_BP1 = SW_C_GS1 + 1;               // c0[d] | c1[d]  of basis j = h
constexpr int SW_C_GS2 = SW_C_BP1 + 2 * SW_D;        // lanes h < D
constexpr int SW_NACC = SW_C_GS2 + 1;
constexpr int SW_TSTRIDE = SW_NB + 1;                // transpose tile: [d][lane][j], padded

// Adjoint of one right-hand side for lane (g, h): gk (replicated) -> gin (replicated); parameter contributions go to
// this lane's accumulator columns acc[c * 32].  x2 is the saved hidden pre-activation of this stage.
template <int BR>
__device__ __forceinline__ void sw_rhs_bwd(const float* __restrict__ sm, const SNLayout& L, int gbase, int h, int lane,
                                           bool lane_ok, const float (&y)[SW_D], const float (&pin1)[SW_D], float pin2,
                                           float x2, const float (&gk)[SW_D], float (&gin)[SW_D], float* acc, float* tile) {
    constexpr int NB = SW_NB, D = SW_D, H = SW_H;
    const int FU = L.FU;
    const float gl1 = sm[L.hdr], gl2 = sm[L.hdr + 1], al1 = sm[L.hdr + 2], al2 = sm[L.hdr + 3];
    const float* rec = sm + L.rec0 + h * L.REC;
    const float* raw1 = sm + L.raw1;
    const float* raw2 = sm + L.raw2;
    auto ACC = [&](int c) -> float& { return acc[c * 32]; };
    // ---- layer-2 features of the saved pre-activation, d(loss)/d(x2)
    const float x = x2;
    const float sg = rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
    const float sl = x * sg;
    float omu = 0.f;
    if constexpr (BR == FETODE_BRANCH_FERRO2D) omu = rcp_approx(1.0f + ex2_approx(gl2 * (x - pin2)));
    float r2[NB], cn2[NB];
    branch_eval<NB, BR>(x, omu, gl2, reinterpret_cast<const float4*>(rec + L.r_bp), r2,
                        BR == FETODE_BRANCH_FERRO2D ? cn2 : nullptr);
    const float2* wu2 = reinterpret_cast<const float2*>(rec + L.r_wu2);
    const float g0 = gk[0], g1 = gk[1];
    float gx2 = fmaf(g0, wu2[0].x, g1 * wu2[0].y) * sg * fmaf(x, 1.0f - sg, 1.0f);
    ACC(SW_C_W2 + 0) += sl * g0;
    ACC(SW_C_W2 + 1) += sl * g1;
    {
        const float4 gi = *reinterpret_cast<const float4*>(rec + L.r_gi);
        if (x >= gi.x && x < gi.z) {
            const float* kp = rec + L.r_kn;
            const int m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
            float v[4], dv[4];
            bspline3_local_d(x, kp, rec + L.r_rd1, rec + L.r_rd2, rec + L.r_rd3, m + 3, v, dv);
            const float2* ws = reinterpret_cast<const float2*>(rec + L.r_ws2) + m;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                gx2 = fmaf(dv[c], fmaf(g0, ws[c].x, g1 * ws[c].y), gx2);
                const int sidx = m - 3 + c;                   // spline coefficient index (rows outside [0, 8) are padding)
                if (sidx >= 0 && sidx < 8) {
                    ACC(SW_C_W2 + (1 + sidx) * D) += v[c] * g0;
                    ACC(SW_C_W2 + (1 + sidx) * D + 1) += v[c] * g1;
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NB; ++j) {
        const float2 w = wu2[1 + j];
        const float Gj = fmaf(g0, w.x, g1 * w.y);
        const float4 bp = reinterpret_cast<const float4*>(rec + L.r_bp)[j];
        float gx, c0, c1;
        branch_bwd<BR, true>(x, omu, gl2, al2, bp.x, r2[j], Gj, raw2[h * NB + j], raw2[(H + h) * NB + j], gx, c0, c1,
                             BR == FETODE_BRANCH_FERRO2D ? cn2[j] : 0.f);
        gx2 += gx;
        ACC(SW_C_W2 + (9 + j) * D) += r2[j] * g0;
        ACC(SW_C_W2 + (9 + j) * D + 1) += r2[j] * g1;
        ACC(SW_C_BP2 + j) += c0;
        ACC(SW_C_BP2 + NB + j) += c1;
    }
    ACC(SW_C_GS1) += gx2;
    if (h < D) ACC(SW_C_GS2) += (h == 0 ? g0 : g1);
    // ---- layer 1: features of the stage input, W1 columns of unit h, adjoints back to the inputs
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float xd = y[d];
        float silu, sg1, omu1, ex1, r;
        sw_l1_own<BR>(sm, L, d, h, xd, pin1[d], gl1, silu, sg1, omu1, ex1, r);
        const float* wu = rec + L.r_wu1 + d * FU;
        ACC(SW_C_W1 + d * SW_F) += silu * gx2;
        float wd = gx2 * wu[0] * sg1 * fmaf(xd, 1.0f - sg1, 1.0f);            // silu' part of d(loss)/d(x_d), this unit's share
        float* trow = tile + (d * 32 + lane) * SW_TSTRIDE;
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const float rj = __shfl_sync(0xffffffffu, r, lane_ok ? gbase + j : gbase + h);
            ACC(SW_C_W1 + d * SW_F + 9 + j) += rj * gx2;
            trow[j] = gx2 * wu[1 + j];                                      // this unit's share of d(loss)/d r1[d][j]
        }
        const float4 gi = reinterpret_cast<const float4*>(sm + L.l1_gi)[d];
        if (xd >= gi.x && xd < gi.z) {
            const float* kp = sm + L.l1_kn + d * L.NKP;
            const int m = bspline_interval(xd, kp, gi.x, gi.y, L.NK);
            float v[4], dv[4];
            bspline3_local_d(xd, kp, sm + L.l1_rd1 + d * L.NKP, sm + L.l1_rd2 + d * L.NKP, sm + L.l1_rd3 + d * L.NKP,
                             m + 3, v, dv);
            const float* ws = rec + L.r_ws1 + d * L.SPP + m;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                wd = fmaf(gx2 * ws[c], dv[c], wd);
                const int sidx = m - 3 + c;
                if (sidx >= 0 && sidx < 8) ACC(SW_C_W1 + d * SW_F + 1 + sidx) += v[c] * gx2;
            }
        }
        __syncwarp();
        // d(loss)/d r1[d][j = h]: the shares of the ten units of this trajectory
        float G1 = 0.f;
        if (lane_ok) {
#pragma unroll
            for (int hh = 0; hh < H; ++hh) G1 += tile[(d * 32 + gbase + hh) * SW_TSTRIDE + h];
        }
        __syncwarp();
        if constexpr (NB > 0) {
            const float4 bp = reinterpret_cast<const float4*>(sm + L.l1_bp)[d * NB + h];
            float gx, c0, c1;
            branch_bwd<BR>(xd, omu1, gl1, al1, ex1 * bp.x, r, G1, raw1[d * NB + h], raw1[(D + d) * NB + h], gx, c0, c1);
            wd += gx;
            ACC(SW_C_BP1 + d) += c0;
            ACC(SW_C_BP1 + D + d) += c1;
        }
        gin[d] = grp_sum(wd, gbase, h, lane_ok);
    }
}

template <int BR>
__global__ void __launch_bounds__(SW_THREADS) smallnet_wide_bwd_kernel(const SWArgs a) {
    extern __shared__ float4 sw_smem4[];
    float* sm = reinterpret_cast<float*>(sw_smem4);
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < L.total_ext / 4; i += SW_THREADS) sw_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    float* acc_all = sm + L.total_ext;                                   // [SW_WARPS][SW_NACC][32]
    float* tile_all = acc_all + SW_WARPS * SW_NACC * 32;                 // [SW_WARPS][D][32][SW_TSTRIDE]
    for (int i = tid; i < SW_WARPS * SW_NACC * 32; i += SW_THREADS) acc_all[i] = 0.f;
    __syncthreads();
    float* acc = acc_all + warp * SW_NACC * 32 + lane;
    float* tile = tile_all + warp * SW_D * 32 * SW_TSTRIDE;
    const bool lane_ok = lane < SW_GRP * SW_H;
    const int g = lane_ok ? lane / SW_H : 0, h = lane_ok ? lane % SW_H : 0, gbase = lane_ok ? g * SW_H : lane;
    const long long n_tri = (a.B + SW_GRP - 1) / SW_GRP;
    const size_t slot_stride = (size_t)a.B * SW_REC;
    for (long long tri = (long long)blockIdx.x * SW_WARPS + warp; tri < n_tri; tri += (long long)gridDim.x * SW_WARPS) {
        const long long t = tri * SW_GRP + g;
        const bool live = lane_ok && t < a.B;
        const long long tt = live ? t : 0;
        float gy[SW_D];
#pragma unroll
        for (int d = 0; d < SW_D; ++d) gy[d] = (live && a.g_yout) ? a.g_yout[tt * SW_D + d] : 0.f;
        for (int step = a.n_steps - 1; step >= 0; --step) {
            if (a.g_traj && (step + 1) % a.dump_every == 0 && live) {
                const float* gt = a.g_traj + ((size_t)((step + 1) / a.dump_every - 1) * a.B + tt) * SW_D;
#pragma unroll
                for (int d = 0; d < SW_D; ++d) gy[d] += gt[d];
            }
            const float* ck = a.ckpt + (((size_t)step * SW_SLOTS) * a.B + tt) * SW_REC;
            float ys[5][SW_D], x2s[5];                                   // slot 0 = the snapshots at the step start
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                if (q <= a.nst) {
                    const float* c = ck + (size_t)q * slot_stride;
                    ys[q][0] = c[0]; ys[q][1] = c[1]; x2s[q] = c[SW_D + h];
                } else { ys[q][0] = ys[q][1] = 0.f; x2s[q] = 0.f; }
            }
            float gys[4][SW_D], gnext[SW_D];
#pragma unroll
            for (int d = 0; d < SW_D; ++d) gnext[d] = gy[d];
#pragma unroll
            for (int s = 3; s >= 0; --s) {
                if (s < a.nst) {
                    float gk[SW_D], gin[SW_D];
#pragma unroll
                    for (int d = 0; d < SW_D; ++d) {
                        float v = a.bw[s] * gy[d];
#pragma unroll
                        for (int q = s + 1; q < 4; ++q)
                            if (q < a.nst) v = fmaf(a.A[q][s], gys[q][d], v);
                        gk[d] = live ? a.dt * v : 0.f;               // dead lanes contribute exact zeros
                    }
                    // stage s saw prev = the previous stage's input / pre-activation (slot s = stage s-1, slot 0 = step start)
                    sw_rhs_bwd<BR>(sm, L, gbase, h, lane, lane_ok, ys[1 + s], ys[s], FERRO ? x2s[s] : 0.f, x2s[1 + s], gk,
                                   gin, acc, tile);
#pragma unroll
                    for (int d = 0; d < SW_D; ++d) { gys[s][d] = gin[d]; gnext[d] += gin[d]; }
                }
            }
#pragma unroll
            for (int d = 0; d < SW_D; ++d) gy[d] = gnext[d];
        }
        if (live && h < SW_D) a.g_y0[tt * SW_D + h] = gy[h];
    }
    // ---- flush: sum the lanes that own the same parameter (3 groups x SW_WARPS warps), one atomic per parameter and CTA
    __syncthreads();
    const SNGfoldW gf = snw_gfold(SW_D, SW_H, SW_F, SW_NB);
    for (int i = tid; i < SW_NACC * SW_H; i += SW_THREADS) {
        const int c = i / SW_H, hh = i % SW_H;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < SW_WARPS; ++w)
#pragma unroll
            for (int gg = 0; gg < SW_GRP; ++gg) v += acc_all[(w * SW_NACC + c) * 32 + gg * SW_H + hh];
        if (v == 0.f) continue;
        int dst = -1;
        if (c < SW_C_BP2) { const int f = c / SW_D, o = c % SW_D; dst = gf.ws2 + (hh * SW_F + f) * SW_D + o; }
        else if (c < SW_C_W1) { const int q = c - SW_C_BP2; dst = gf.bp2 + (q / SW_NB) * SW_H * SW_NB + hh * SW_NB + q % SW_NB; }
        else if (c < SW_C_GS1) { const int q = c - SW_C_W1; dst = gf.ws1 + q * SW_H + hh; }         // q = d * F + f
        else if (c == SW_C_GS1) dst = gf.gsum1 + hh;
        else if (c < SW_C_GS2) { const int q = c - SW_C_BP1; dst = gf.bp1 + (q / SW_D) * SW_D * SW_NB + (q % SW_D) * SW_NB + hh; }
        else if (hh < SW_D) dst = gf.gsum2 + hh;
        if (dst >= 0) atomicAdd(a.gfold + dst, v);
    }
}

// ------------------------------------------------------------------------------------------
// one explicit Runge-Kutta ATTEMPT with an arbitrary tableau (<= 7 stages): the building block of the adaptive solver
// ------------------------------------------------------------------------------------------
// torchdiffeq's dopri5 (call sites train_kanfet_node_predprey.py:252, the reference's default LV solver) evaluates the
// right-hand side six times per attempted step (stage 0 is the previous step's last stage: FSAL), forms the 5th-order
// solution, an embedded error estimate and ONE error norm over the whole batch.  Here the whole attempt is one launch:
//   k_0 given (first_eval = 1) or evaluated (first_eval = 0);  stage s input  ys_s = y + dt sum_{r<s} A[s][r] k_r ,
//   k_s = f(ys_s) for s >= first_eval;  y1 = ys_{ns-1} (y1_last != 0);  outputs ks [ns, B, D], y1 [B, D];
//   errsq += sum_{rows, d} ( dt sum_s E[s] k_s / (atol + rtol max(|y|, |y1|)) )^2     (one double, atomics per warp);
//   hysteresis state advanced by every evaluated stage (also on attempts the controller later rejects, as the
//   reference's modules are);  ckpt [1 + ns - first_eval][B][12] for the reverse sweep of an ACCEPTED attempt.
constexpr int SW_MAXS = 7;

struct SWStepArgs {
    const float* packed;
    const float* y;           // [B, D]
    const float* k0;          // [B, D] when first_eval == 1
    float* y1;                // [B, D] or null
    float* ks;                // [ns, B, D]
    float* state;             // [B, D + H] in/out (ferro) or null
    float* ckpt;              // [slots, B, SW_REC] or null
    double* errsq;            // or null
    const float *g_y1, *g_ks; // backward inputs ([B, D] or null; [ns, B, D])
    const float *x_y1, *x_klast, *x_y;   // optional [B, D] addends: to g_y1, to g_ks[ns-1], to the returned g_y
    float *g_y, *g_k0, *gfold;
    long long B;
    float dt, rtol, atol;
    float A[SW_MAXS][SW_MAXS], E[SW_MAXS];
    int ns, first_eval, y1_last, state_is_zero;
};

template <int BR>
__global__ void __launch_bounds__(SW_THREADS) smallnet_wide_step_fwd_kernel(const SWStepArgs a) {
    extern __shared__ float4 sw_smem4[];
    float* sm = reinterpret_cast<float*>(sw_smem4);
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    constexpr int SF = FERRO ? SW_D + SW_H : 0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < L.total / 4; i += SW_THREADS) sw_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    __syncthreads();
    const bool lane_ok = lane < SW_GRP * SW_H;
    const int g = lane_ok ? lane / SW_H : 0, h = lane_ok ? lane % SW_H : 0, gbase = lane_ok ? g * SW_H : lane;
    const long long n_tri = (a.B + SW_GRP - 1) / SW_GRP;
    const size_t slot_stride = (size_t)a.B * SW_REC;
    double esum = 0.0;
    for (long long tri = (long long)blockIdx.x * SW_WARPS + warp; tri < n_tri; tri += (long long)gridDim.x * SW_WARPS) {
        const long long t = tri * SW_GRP + g;
        const bool live = lane_ok && t < a.B;
        const long long tt = live ? t : 0;
        float y[SW_D], prev1[SW_D] = {0.f, 0.f}, prev2 = 0.f, k[SW_MAXS][SW_D], ys[SW_D];
#pragma unroll
        for (int d = 0; d < SW_D; ++d) y[d] = a.y[tt * SW_D + d];
        if (FERRO && !a.state_is_zero) {
#pragma unroll
            for (int d = 0; d < SW_D; ++d) prev1[d] = a.state[tt * SF + d];
            prev2 = a.state[tt * SF + SW_D + h];
        }
        float* ck = a.ckpt ? a.ckpt + tt * SW_REC : nullptr;
        if (ck && live) {
            if (h < SW_D) ck[h] = prev1[h];
            ck[SW_D + h] = prev2;
        }
#pragma unroll
        for (int s = 0; s < SW_MAXS; ++s) { k[s][0] = 0.f; k[s][1] = 0.f; }
        if (a.first_eval) {
#pragma unroll
            for (int d = 0; d < SW_D; ++d) k[0][d] = a.k0[tt * SW_D + d];
        }
#pragma unroll
        for (int s = 0; s < SW_MAXS; ++s) {
            if (s < a.ns) {
#pragma unroll
                for (int d = 0; d < SW_D; ++d) {
                    float acc = 0.f;
#pragma unroll
                    for (int r = 0; r < s; ++r) acc = fmaf(a.A[s][r], k[r][d], acc);
                    ys[d] = fmaf(a.dt, acc, y[d]);
                }
                if (s >= a.first_eval) {
                    float x2;
                    sw_rhs<BR>(sm, L, gbase, h, lane_ok, ys, prev1, prev2, k[s], x2);
                    if (ck && live) {
                        float* c = ck + (size_t)(1 + s - a.first_eval) * slot_stride;
                        if (h < SW_D) c[h] = ys[h];
                        c[SW_D + h] = x2;
                    }
                }
            }
        }
        if (live && h < SW_D) {
#pragma unroll
            for (int s = 0; s < SW_MAXS; ++s)
                if (s < a.ns) a.ks[((size_t)s * a.B + tt) * SW_D + h] = k[s][h];
            if (a.y1) a.y1[tt * SW_D + h] = a.y1_last ? ys[h] : y[h];
            if (a.errsq) {
                float e = 0.f;
#pragma unroll
                for (int s = 0; s < SW_MAXS; ++s)
                    if (s < a.ns) e = fmaf(a.E[s], k[s][h], e);
                e *= a.dt;
                const float tol = fmaf(a.rtol, fmaxf(fabsf(y[h]), fabsf(ys[h])), a.atol);
                const float q = e / tol;
                esum += (double)q * (double)q;
            }
        }
        if (live && FERRO && a.state) {
            if (h < SW_D) a.state[tt * SF + h] = prev1[h];
            a.state[tt * SF + SW_D + h] = prev2;
        }
    }
    if (a.errsq) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) esum += __shfl_xor_sync(0xffffffffu, esum, o);
        if (lane == 0 && esum != 0.0) atomicAdd(a.errsq, esum);
    }
}

template <int BR>
__global__ void __launch_bounds__(SW_THREADS) smallnet_wide_step_bwd_kernel(const SWStepArgs a) {
    extern __shared__ float4 sw_smem4[];
    float* sm = reinterpret_cast<float*>(sw_smem4);
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < L.total_ext / 4; i += SW_THREADS) sw_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    float* acc_all = sm + L.total_ext;
    float* tile_all = acc_all + SW_WARPS * SW_NACC * 32;
    for (int i = tid; i < SW_WARPS * SW_NACC * 32; i += SW_THREADS) acc_all[i] = 0.f;
    __syncthreads();
    float* acc = acc_all + warp * SW_NACC * 32 + lane;
    float* tile = tile_all + warp * SW_D * 32 * SW_TSTRIDE;
    const bool lane_ok = lane < SW_GRP * SW_H;
    const int g = lane_ok ? lane / SW_H : 0, h = lane_ok ? lane % SW_H : 0, gbase = lane_ok ? g * SW_H : lane;
    const long long n_tri = (a.B + SW_GRP - 1) / SW_GRP;
    const size_t slot_stride = (size_t)a.B * SW_REC;
    for (long long tri = (long long)blockIdx.x * SW_WARPS + warp; tri < n_tri; tri += (long long)gridDim.x * SW_WARPS) {
        const long long t = tri * SW_GRP + g;
        const bool live = lane_ok && t < a.B;
        const long long tt = live ? t : 0;
        const float* ck = a.ckpt + tt * SW_REC;
        float gY[SW_MAXS][SW_D];                       // adjoint of every stage INPUT
#pragma unroll
        for (int s = 0; s < SW_MAXS; ++s) { gY[s][0] = 0.f; gY[s][1] = 0.f; }
        float gy[SW_D] = {0.f, 0.f};
#pragma unroll
        for (int s = SW_MAXS - 1; s >= 0; --s) {
            if (s < a.ns && s >= a.first_eval) {
                float gk[SW_D], gin[SW_D], yin[SW_D], pin[SW_D];
                const int slot = 1 + s - a.first_eval;
                const float* c = ck + (size_t)slot * slot_stride;
                const float* cp = ck + (size_t)(slot - 1) * slot_stride;      // what this stage saw as prev_x
#pragma unroll
                for (int d = 0; d < SW_D; ++d) {
                    float v = a.g_ks ? a.g_ks[((size_t)s * a.B + tt) * SW_D + d] : 0.f;
                    if (s == a.ns - 1 && a.x_klast) v += a.x_klast[tt * SW_D + d];
#pragma unroll
                    for (int q = s + 1; q < SW_MAXS; ++q)
                        if (q < a.ns) v = fmaf(a.dt * a.A[q][s], gY[q][d], v);
                    gk[d] = live ? v : 0.f;
                    yin[d] = c[d];
                    pin[d] = cp[d];
                }
                sw_rhs_bwd<BR>(sm, L, gbase, h, lane, lane_ok, yin, pin, FERRO ? cp[SW_D + h] : 0.f, c[SW_D + h], gk, gin,
                               acc, tile);
#pragma unroll
                for (int d = 0; d < SW_D; ++d) {
                    float top = 0.f;
                    if (a.y1_last && s == a.ns - 1 && live) {
                        if (a.g_y1) top += a.g_y1[tt * SW_D + d];
                        if (a.x_y1) top += a.x_y1[tt * SW_D + d];
                    }
                    gY[s][d] = gin[d] + top;
                    gy[d] += gY[s][d];
                }
            }
        }
        if (live && h < SW_D) {
            float v = gy[h];
            if (!a.y1_last && a.g_y1) v += a.g_y1[tt * SW_D + h];
            if (!a.y1_last && a.x_y1) v += a.x_y1[tt * SW_D + h];
            if (a.x_y) v += a.x_y[tt * SW_D + h];
            a.g_y[tt * SW_D + h] = v;
            if (a.first_eval && a.g_k0) {
                float w = a.g_ks ? a.g_ks[tt * SW_D + h] : 0.f;
                if (a.ns == 1 && a.x_klast) w += a.x_klast[tt * SW_D + h];
#pragma unroll
                for (int q = 1; q < SW_MAXS; ++q)
                    if (q < a.ns) w = fmaf(a.dt * a.A[q][0], gY[q][h], w);
                a.g_k0[tt * SW_D + h] = w;
            }
        }
    }
    __syncthreads();
    const SNGfoldW gf = snw_gfold(SW_D, SW_H, SW_F, SW_NB);
    for (int i = tid; i < SW_NACC * SW_H; i += SW_THREADS) {
        const int c = i / SW_H, hh = i % SW_H;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < SW_WARPS; ++w)
#pragma unroll
            for (int gg = 0; gg < SW_GRP; ++gg) v += acc_all[(w * SW_NACC + c) * 32 + gg * SW_H + hh];
        if (v == 0.f) continue;
        int dst = -1;
        if (c < SW_C_BP2) { const int f = c / SW_D, o = c % SW_D; dst = gf.ws2 + (hh * SW_F + f) * SW_D + o; }
        else if (c < SW_C_W1) { const int q = c - SW_C_BP2; dst = gf.bp2 + (q / SW_NB) * SW_H * SW_NB + hh * SW_NB + q % SW_NB; }
        else if (c < SW_C_GS1) { const int q = c - SW_C_W1; dst = gf.ws1 + q * SW_H + hh; }
        else if (c == SW_C_GS1) dst = gf.gsum1 + hh;
        else if (c < SW_C_GS2) { const int q = c - SW_C_BP1; dst = gf.bp1 + (q / SW_D) * SW_D * SW_NB + (q % SW_D) * SW_NB + hh; }
        else if (hh < SW_D) dst = gf.gsum2 + hh;
        if (dst >= 0) atomicAdd(a.gfold + dst, v);
    }
}

// ------------------------------------------------------------------------------------------
// the whole adaptive solve on the device (small batches: one CTA holds every trajectory)
// ------------------------------------------------------------------------------------------
// For B <= SW_GRP * SW_WARPS trajectories -- the reference's Lotka-Volterra run is ONE -- the dopri5 controller itself
// runs in the kernel: attempt, error norm (block reduction), accept / reject, step-size update (double precision, the
// host controller's formulas), dense output at every requested time the accepted step covers.  The host launches it
// once per `cap` accepted steps (checkpoint chunk) and reads eight doubles back per launch instead of one per attempt.
//   ctrl[0] tb  [1] ta  [2] dt  [3] next output index  [4] accepted  [5] rejected  [6] status  [7] accepted in this chunk
//   status: 0 done, 1 launch again (chunk full / attempt budget spent), 2 dt underflow, 3 non-finite error ratio.
constexpr int SW_SOLVE_MAXB = SW_GRP * SW_WARPS;

struct SWSolveArgs {
    const float* packed;
    float *y, *fy;             // [B, D] in/out: solution and FSAL slope at tb
    float* state;              // [B, D + H] in/out or null
    float* out;                // [T, B, D]; rows 1.. written here
    const double* times;       // [T]
    double* ctrl;              // [8]
    float* ckpt;               // [cap][7][B][SW_REC] or null (no reverse sweep wanted)
    float* dts;                // [cap]
    int* rec_step;             // [T]  absolute index of the accepted step that produced output oi
    float* rec_w;              // [T][9] weights of (ya, yb, k_0..k_6)
    double cmid[SW_MAXS];
    float A[SW_MAXS][SW_MAXS], E[SW_MAXS];
    float rtol, atol;
    int B, T, cap, max_attempts, state_is_zero;
    // reverse sweep
    const float* g_out;        // [T, B, D]
    float *carry_y, *carry_k;  // [B, D] in/out
    float* gfold;
    int n_steps, step_base;    // chunk length, absolute index of its first step
    int* cursor;               // in/out: highest output index whose adjoint has not been consumed yet
};

__device__ __forceinline__ void sw_dense_weights(const double (&cmid)[SW_MAXS], double ta, double tb, double tq, double h,
                                                 float (&w)[9]) {
    if (tq == tb) {
#pragma unroll
        for (int i = 0; i < 9; ++i) w[i] = i == 1 ? 1.f : 0.f;
        return;
    }
    const double x = (tq - ta) / (tb - ta), x2 = x * x, x3 = x2 * x, x4 = x2 * x2;
    const double w_mid = 16 * x2 - 32 * x3 + 16 * x4;
    w[0] = (float)(1.0 - 11 * x2 + 18 * x3 - 8 * x4 + w_mid);
    w[1] = (float)(-5 * x2 + 14 * x3 - 8 * x4);
#pragma unroll
    for (int s = 0; s < SW_MAXS; ++s) {
        double v = w_mid * h * cmid[s];
        if (s == 0) v += h * (x - 4 * x2 + 5 * x3 - 2 * x4);
        if (s == 6) v += h * (x2 - 3 * x3 + 2 * x4);
        w[2 + s] = (float)v;
    }
}

template <int BR>
__global__ void __launch_bounds__(SW_THREADS) smallnet_wide_solve_fwd_kernel(const SWSolveArgs a) {
    extern __shared__ float4 sw_smem4[];
    float* sm = reinterpret_cast<float*>(sw_smem4);
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    constexpr int SF = FERRO ? SW_D + SW_H : 0;
    __shared__ double red[SW_WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < L.total / 4; i += SW_THREADS) sw_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    __syncthreads();
    const bool lane_ok = lane < SW_GRP * SW_H;
    const int g = lane_ok ? lane / SW_H : 0, h = lane_ok ? lane % SW_H : 0, gbase = lane_ok ? g * SW_H : lane;
    const int t = warp * SW_GRP + g;
    const bool live = lane_ok && t < a.B;
    const int tt = live ? t : 0;
    const size_t slot_stride = (size_t)a.B * SW_REC;
    float y[SW_D], prev1[SW_D] = {0.f, 0.f}, prev2 = 0.f, k[SW_MAXS][SW_D], ys[SW_D];
#pragma unroll
    for (int d = 0; d < SW_D; ++d) { y[d] = a.y[tt * SW_D + d]; k[0][d] = a.fy[tt * SW_D + d]; }
    if (FERRO && !a.state_is_zero) {
#pragma unroll
        for (int d = 0; d < SW_D; ++d) prev1[d] = a.state[tt * SF + d];
        prev2 = a.state[tt * SF + SW_D + h];
    }
    double tb = a.ctrl[0], ta = a.ctrl[1], dt = a.ctrl[2];
    int oi = (int)a.ctrl[3];
    long long n_acc = (long long)a.ctrl[4], n_rej = (long long)a.ctrl[5];
    int chunk = 0, attempts = 0, status = 0;
    const double n_el = (double)a.B * SW_D;
    while (true) {
        if (oi >= a.T) { status = 0; break; }
        if (attempts >= a.max_attempts || (a.ckpt && chunk >= a.cap)) { status = 1; break; }
        const double t_new = tb + dt;
        if (!(isfinite(dt) && t_new > tb)) { status = 2; break; }
        const float dtf = (float)dt;
        float* ck = a.ckpt ? a.ckpt + ((size_t)chunk * SW_MAXS) * slot_stride + (size_t)tt * SW_REC : nullptr;
        if (ck && live) {
            if (h < SW_D) ck[h] = prev1[h];
            ck[SW_D + h] = prev2;
        }
#pragma unroll
        for (int s = 1; s < SW_MAXS; ++s) {
#pragma unroll
            for (int d = 0; d < SW_D; ++d) {
                float acc = 0.f;
#pragma unroll
                for (int r = 0; r < s; ++r) acc = fmaf(a.A[s][r], k[r][d], acc);
                ys[d] = fmaf(dtf, acc, y[d]);
            }
            float x2;
            sw_rhs<BR>(sm, L, gbase, h, lane_ok, ys, prev1, prev2, k[s], x2);
            if (ck && live) {
                float* c = ck + (size_t)s * slot_stride;
                if (h < SW_D) c[h] = ys[h];
                c[SW_D + h] = x2;
            }
        }
        ++attempts;
        // ---- error norm over the whole batch (the controller's one global RMS)
        double esum = 0.0;
        if (live && h < SW_D) {
            float e = 0.f;
#pragma unroll
            for (int s = 0; s < SW_MAXS; ++s) e = fmaf(a.E[s], k[s][h], e);
            e *= dtf;
            const float q = e / fmaf(a.rtol, fmaxf(fabsf(y[h]), fabsf(ys[h])), a.atol);
            esum = (double)q * (double)q;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) esum += __shfl_xor_sync(0xffffffffu, esum, o);
        if (lane == 0) red[warp] = esum;
        __syncthreads();
        double tot = 0.0;
#pragma unroll
        for (int w = 0; w < SW_WARPS; ++w) tot += red[w];
        __syncthreads();
        const double ratio = sqrt(tot / n_el);
        if (!isfinite(ratio)) { status = 3; break; }
        if (ratio <= 1.0) {
            while (oi < a.T && a.times[oi] <= t_new) {
                float w[9];
                sw_dense_weights(a.cmid, tb, t_new, a.times[oi], dt, w);
                if (live && h < SW_D) {
                    float v;
                    if (w[1] == 1.f && w[0] == 0.f) v = ys[h];              // tq == tb: the step's end point itself
                    else {
                        v = w[0] * y[h];
                        v = fmaf(w[1], ys[h], v);
#pragma unroll
                        for (int s = 0; s < SW_MAXS; ++s) v = fmaf(w[2 + s], k[s][h], v);
                    }
                    a.out[((size_t)oi * a.B + tt) * SW_D + h] = v;
                }
                if (tid == 0 && a.rec_step) {
                    a.rec_step[oi] = (int)n_acc;
#pragma unroll
                    for (int i = 0; i < 9; ++i) a.rec_w[oi * 9 + i] = w[i];
                }
                ++oi;
            }
            if (tid == 0 && a.dts) a.dts[chunk] = dtf;
#pragma unroll
            for (int d = 0; d < SW_D; ++d) { y[d] = ys[d]; k[0][d] = k[SW_MAXS - 1][d]; }
            ta = tb; tb = t_new;
            ++n_acc; ++chunk;
        } else {
            ++n_rej;
        }
        if (ratio == 0.0) dt *= 10.0;
        else dt *= fmin(10.0, fmax(0.9 / pow(ratio, 0.2), ratio < 1.0 ? 1.0 : 0.2));
    }
    if (live) {
        if (h < SW_D) { a.y[tt * SW_D + h] = y[h]; a.fy[tt * SW_D + h] = k[0][h]; }
        if (FERRO && a.state) {
            if (h < SW_D) a.state[tt * SF + h] = prev1[h];
            a.state[tt * SF + SW_D + h] = prev2;
        }
    }
    if (tid == 0) {
        a.ctrl[0] = tb; a.ctrl[1] = ta; a.ctrl[2] = dt; a.ctrl[3] = (double)oi; a.ctrl[4] = (double)n_acc;
        a.ctrl[5] = (double)n_rej; a.ctrl[6] = (double)status; a.ctrl[7] = (double)chunk;
    }
}

template <int BR>
__global__ void __launch_bounds__(SW_THREADS) smallnet_wide_solve_bwd_kernel(const SWSolveArgs a) {
    extern __shared__ float4 sw_smem4[];
    float* sm = reinterpret_cast<float*>(sw_smem4);
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < L.total_ext / 4; i += SW_THREADS) sw_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    float* acc_all = sm + L.total_ext;
    float* tile_all = acc_all + SW_WARPS * SW_NACC * 32;
    for (int i = tid; i < SW_WARPS * SW_NACC * 32; i += SW_THREADS) acc_all[i] = 0.f;
    int oi = *a.cursor;                                  // read by everyone before the barrier, written after the sweep
    __syncthreads();
    float* acc = acc_all + warp * SW_NACC * 32 + lane;
    float* tile = tile_all + warp * SW_D * 32 * SW_TSTRIDE;
    const bool lane_ok = lane < SW_GRP * SW_H;
    const int g = lane_ok ? lane / SW_H : 0, h = lane_ok ? lane % SW_H : 0, gbase = lane_ok ? g * SW_H : lane;
    const int t = warp * SW_GRP + g;
    const bool live = lane_ok && t < a.B;
    const int tt = live ? t : 0;
    const size_t slot_stride = (size_t)a.B * SW_REC;
    float cy[SW_D], ckk[SW_D];
#pragma unroll
    for (int d = 0; d < SW_D; ++d) { cy[d] = live ? a.carry_y[tt * SW_D + d] : 0.f; ckk[d] = live ? a.carry_k[tt * SW_D + d] : 0.f; }
    for (int n = a.n_steps - 1; n >= 0; --n) {
        const float dtf = a.dts[n];
        float G[9][SW_D];
#pragma unroll
        for (int i = 0; i < 9; ++i) { G[i][0] = 0.f; G[i][1] = 0.f; }
        while (oi >= 1 && a.rec_step[oi] == a.step_base + n) {
#pragma unroll
            for (int d = 0; d < SW_D; ++d) {
                const float go = live ? a.g_out[((size_t)oi * a.B + tt) * SW_D + d] : 0.f;
#pragma unroll
                for (int i = 0; i < 9; ++i) G[i][d] = fmaf(a.rec_w[oi * 9 + i], go, G[i][d]);
            }
            --oi;
        }
        const float* ck = a.ckpt + ((size_t)n * SW_MAXS) * slot_stride + (size_t)tt * SW_REC;
        float gY[SW_MAXS][SW_D], gy[SW_D] = {0.f, 0.f};
#pragma unroll
        for (int s = 0; s < SW_MAXS; ++s) { gY[s][0] = 0.f; gY[s][1] = 0.f; }
#pragma unroll
        for (int s = SW_MAXS - 1; s >= 1; --s) {
            float gk[SW_D], gin[SW_D], yin[SW_D], pin[SW_D];
            const float* c = ck + (size_t)s * slot_stride;
            const float* cp = ck + (size_t)(s - 1) * slot_stride;
#pragma unroll
            for (int d = 0; d < SW_D; ++d) {
                float v = G[2 + s][d] + (s == SW_MAXS - 1 ? ckk[d] : 0.f);
#pragma unroll
                for (int q = s + 1; q < SW_MAXS; ++q) v = fmaf(dtf * a.A[q][s], gY[q][d], v);
                gk[d] = live ? v : 0.f;
                yin[d] = c[d];
                pin[d] = cp[d];
            }
            sw_rhs_bwd<BR>(sm, L, gbase, h, lane, lane_ok, yin, pin, FERRO ? cp[SW_D + h] : 0.f, c[SW_D + h], gk, gin, acc,
                           tile);
#pragma unroll
            for (int d = 0; d < SW_D; ++d) {
                gY[s][d] = gin[d] + (s == SW_MAXS - 1 ? G[1][d] + cy[d] : 0.f);
                gy[d] += gY[s][d];
            }
        }
#pragma unroll
        for (int d = 0; d < SW_D; ++d) {
            float w = G[2][d];
#pragma unroll
            for (int q = 1; q < SW_MAXS; ++q) w = fmaf(dtf * a.A[q][0], gY[q][d], w);
            ckk[d] = w;
            cy[d] = gy[d] + G[0][d];
        }
    }
    if (live && h < SW_D) { a.carry_y[tt * SW_D + h] = cy[h]; a.carry_k[tt * SW_D + h] = ckk[h]; }
    __syncthreads();
    if (tid == 0) *a.cursor = oi;
    const SNGfoldW gf = snw_gfold(SW_D, SW_H, SW_F, SW_NB);
    for (int i = tid; i < SW_NACC * SW_H; i += SW_THREADS) {
        const int c = i / SW_H, hh = i % SW_H;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < SW_WARPS; ++w)
#pragma unroll
            for (int gg = 0; gg < SW_GRP; ++gg) v += acc_all[(w * SW_NACC + c) * 32 + gg * SW_H + hh];
        if (v == 0.f) continue;
        int dst = -1;
        if (c < SW_C_BP2) { const int f = c / SW_D, o = c % SW_D; dst = gf.ws2 + (hh * SW_F + f) * SW_D + o; }
        else if (c < SW_C_W1) { const int q = c - SW_C_BP2; dst = gf.bp2 + (q / SW_NB) * SW_H * SW_NB + hh * SW_NB + q % SW_NB; }
        else if (c < SW_C_GS1) { const int q = c - SW_C_W1; dst = gf.ws1 + q * SW_H + hh; }
        else if (c == SW_C_GS1) dst = gf.gsum1 + hh;
        else if (c < SW_C_GS2) { const int q = c - SW_C_BP1; dst = gf.bp1 + (q / SW_D) * SW_D * SW_NB + (q % SW_D) * SW_NB + hh; }
        else if (hh < SW_D) dst = gf.gsum2 + hh;
        if (dst >= 0) atomicAdd(a.gfold + dst, v);
    }
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_smallnet_wide_fwd(const fetode_smallnet_t* net, const float* packed, int tableau, const float* y0,
                                        float* y_out, float* state, int state_is_zero, int64_t B, double dt,
                                        int32_t n_steps, float* traj, int32_t dump_every, float* ckpt, void* stream);

static bool snw_has(const fetode_smallnet_t* net) {
    return net && net->n_layers == 2 && net->widths[0] == SW_D && net->widths[1] == SW_H && net->widths[2] == SW_D &&
           net->grid_size == SW_G && net->spline_order == 3 && net->num_basis == SW_NB &&
           (net->branch == FETODE_BRANCH_FERRO2D || net->branch == FETODE_BRANCH_LOGISTIC);
}

static int snw_tableau(int tableau, SWArgs& a) {
    for (int i = 0; i < 4; ++i) { a.bw[i] = 0.f; for (int j = 0; j < 4; ++j) a.A[i][j] = 0.f; }
    switch (tableau) {
        case FETODE_TABLEAU_EULER: a.nst = 1; a.bw[0] = 1.f; break;
        case FETODE_TABLEAU_MIDPOINT: a.nst = 2; a.A[1][0] = 0.5f; a.bw[1] = 1.f; break;
        case FETODE_TABLEAU_RK4:
            a.nst = 4; a.A[1][0] = 0.5f; a.A[2][1] = 0.5f; a.A[3][2] = 1.f;
            a.bw[0] = a.bw[3] = (float)(1.0 / 6.0); a.bw[1] = a.bw[2] = (float)(1.0 / 3.0);
            break;
        case FETODE_TABLEAU_RK4_38:
            a.nst = 4; a.A[1][0] = (float)(1.0 / 3.0); a.A[2][0] = (float)(-1.0 / 3.0); a.A[2][1] = 1.f;
            a.A[3][0] = 1.f; a.A[3][1] = -1.f; a.A[3][2] = 1.f;
            a.bw[0] = a.bw[3] = 0.125f; a.bw[1] = a.bw[2] = 0.375f;
            break;
        default: return FETODE_EINVAL;
    }
    return FETODE_OK;
}

static unsigned snw_grid(long long B, int ctas_per_sm) {
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long n_tri = (B + SW_GRP - 1) / SW_GRP;
    const long long want = (n_tri + SW_WARPS - 1) / SW_WARPS;
    return (unsigned)std::max(1LL, std::min<long long>(want, (long long)sms * ctas_per_sm));
}

namespace fetode {
// entry for the generic forward dispatch (smallnet.cu): small batches without lane-per-trajectory checkpoints
bool snw_prefers(const fetode_smallnet_t* net, long long B, bool wants_v1_ckpt) {
    static int forced = -2;
    if (forced == -2) { const char* e = getenv("FETODE_SN_WIDE_FWD"); forced = e ? atoi(e) : -1; }
    if (!snw_has(net) || wants_v1_ckpt || forced == 0) return false;
    return forced == 1 || B <= 12288;     // measured (1000 RK4 steps): B = 1: 3.8 vs 27 ms, 4096: 7.4 vs 27.4; 16384: ~28 vs 27.6
}
int snw_forward(const fetode_smallnet_t* net, const SNArgs& s, cudaStream_t st) {
    return fetode_smallnet_wide_fwd(net, s.packed, s.tab, s.y0, s.y_out, s.state, s.state_is_zero, s.B, (double)s.dt,
                                    s.n_steps, s.traj, s.dump_every, nullptr, (void*)st);
}
}  // namespace fetode

extern "C" int fetode_smallnet_wide_supported(const fetode_smallnet_t* net) {
    return snw_has(net) ? FETODE_OK : FETODE_EUNSUPPORTED;
}

extern "C" int64_t fetode_smallnet_wide_ckpt_floats(const fetode_smallnet_t* net, int64_t B, int32_t n_steps) {
    if (!snw_has(net) || B < 0 || n_steps < 0) return -1;
    return (int64_t)n_steps * SW_SLOTS * B * SW_REC;
}

extern "C" int fetode_smallnet_wide_fwd(const fetode_smallnet_t* net, const float* packed, int tableau, const float* y0,
                                        float* y_out, float* state, int state_is_zero, int64_t B, double dt,
                                        int32_t n_steps, float* traj, int32_t dump_every, float* ckpt, void* stream) {
    if (!snw_has(net)) return FETODE_EUNSUPPORTED;
    FETODE_CHECK_ARG(packed && y0 && y_out && B >= 0 && n_steps >= 0);
    if (B == 0) return FETODE_OK;
    if (net->branch == FETODE_BRANCH_FERRO2D && !state_is_zero) FETODE_CHECK_ARG(state != nullptr);
    if (traj) FETODE_CHECK_ARG(dump_every >= 1);
    SWArgs a{};
    int rc = snw_tableau(tableau, a);
    if (rc) return rc;
    a.packed = packed; a.y0 = y0; a.y_out = y_out; a.state = state; a.traj = traj; a.ckpt = ckpt;
    a.B = B; a.dt = (float)dt; a.n_steps = n_steps; a.dump_every = dump_every > 0 ? dump_every : 1;
    a.state_is_zero = state_is_zero;
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    const size_t smem = (size_t)L.total * 4;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = snw_grid(B, 6);
    if (net->branch == FETODE_BRANCH_FERRO2D) {
        auto k = smallnet_wide_fwd_kernel<FETODE_BRANCH_FERRO2D>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    } else {
        auto k = smallnet_wide_fwd_kernel<FETODE_BRANCH_LOGISTIC>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_smallnet_wide_bwd(const fetode_smallnet_t* net, const float* packed, int tableau, const float* ckpt,
                                        const float* g_yout, const float* g_traj, int32_t dump_every, float* g_y0,
                                        float* gfold, int64_t B, double dt, int32_t n_steps, void* stream) {
    if (!snw_has(net)) return FETODE_EUNSUPPORTED;
    FETODE_CHECK_ARG(packed && g_y0 && gfold && B >= 0 && n_steps >= 0 && (n_steps == 0 || ckpt));
    if (g_traj) FETODE_CHECK_ARG(dump_every >= 1);
    if (B == 0) return FETODE_OK;
    SWArgs a{};
    int rc = snw_tableau(tableau, a);
    if (rc) return rc;
    a.packed = packed; a.ckpt = const_cast<float*>(ckpt); a.g_yout = g_yout; a.g_traj = g_traj; a.g_y0 = g_y0; a.gfold = gfold;
    a.B = B; a.dt = (float)dt; a.n_steps = n_steps; a.dump_every = dump_every > 0 ? dump_every : 1;
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    const size_t smem = (size_t)(L.total_ext + SW_WARPS * SW_NACC * 32 + SW_WARPS * SW_D * 32 * SW_TSTRIDE) * 4;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = snw_grid(B, 3);
    if (net->branch == FETODE_BRANCH_FERRO2D) {
        auto k = smallnet_wide_bwd_kernel<FETODE_BRANCH_FERRO2D>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    } else {
        auto k = smallnet_wide_bwd_kernel<FETODE_BRANCH_LOGISTIC>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

// ---- one Runge-Kutta attempt with a caller-supplied tableau (adaptive solvers) ---------------------------------------
static int snw_attempt_args(const fetode_smallnet_t* net, const fetode_rk_attempt_t* rk, SWStepArgs& a) {
    if (!snw_has(net)) return FETODE_EUNSUPPORTED;
    if (!rk || rk->n_stages < 1 || rk->n_stages > SW_MAXS || rk->first_eval < 0 || rk->first_eval > 1 ||
        rk->first_eval >= rk->n_stages)
        return FETODE_EINVAL;
    a.ns = rk->n_stages; a.first_eval = rk->first_eval; a.y1_last = rk->y1_is_last_input ? 1 : 0;
    a.dt = rk->dt; a.rtol = rk->rtol; a.atol = rk->atol;
    for (int i = 0; i < SW_MAXS; ++i) {
        a.E[i] = rk->e[i];
        for (int j = 0; j < SW_MAXS; ++j) a.A[i][j] = j < i ? rk->a[i][j] : 0.f;
    }
    return FETODE_OK;
}

extern "C" int64_t fetode_smallnet_attempt_ckpt_floats(const fetode_smallnet_t* net, const fetode_rk_attempt_t* rk,
                                                       int64_t B) {
    SWStepArgs a{};
    if (snw_attempt_args(net, rk, a) != FETODE_OK || B < 0) return -1;
    return (int64_t)(1 + a.ns - a.first_eval) * B * SW_REC;
}

extern "C" int fetode_smallnet_attempt_fwd(const fetode_smallnet_t* net, const float* packed,
                                           const fetode_rk_attempt_t* rk, const float* y, const float* k0, float* y1,
                                           float* ks, float* state, int state_is_zero, int64_t B, float* ckpt,
                                           double* errsq, void* stream) {
    SWStepArgs a{};
    int rc = snw_attempt_args(net, rk, a);
    if (rc) return rc;
    FETODE_CHECK_ARG(packed && y && ks && B >= 0 && (!a.first_eval || k0));
    if (B == 0) return FETODE_OK;
    if (net->branch == FETODE_BRANCH_FERRO2D && !state_is_zero) FETODE_CHECK_ARG(state != nullptr);
    a.packed = packed; a.y = y; a.k0 = k0; a.y1 = y1; a.ks = ks; a.state = state; a.ckpt = ckpt; a.errsq = errsq;
    a.B = B; a.state_is_zero = state_is_zero;
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    const size_t smem = (size_t)L.total * 4;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = snw_grid(B, 4);
    if (net->branch == FETODE_BRANCH_FERRO2D) {
        auto k = smallnet_wide_step_fwd_kernel<FETODE_BRANCH_FERRO2D>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    } else {
        auto k = smallnet_wide_step_fwd_kernel<FETODE_BRANCH_LOGISTIC>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_smallnet_attempt_bwd(const fetode_smallnet_t* net, const float* packed,
                                           const fetode_rk_attempt_t* rk, const float* ckpt, const float* g_y1,
                                           const float* g_ks, const float* x_y1, const float* x_klast, const float* x_y,
                                           float* g_y, float* g_k0, float* gfold, int64_t B, void* stream) {
    SWStepArgs a{};
    int rc = snw_attempt_args(net, rk, a);
    if (rc) return rc;
    FETODE_CHECK_ARG(packed && ckpt && g_y && gfold && B >= 0);
    if (B == 0) return FETODE_OK;
    a.packed = packed; a.ckpt = const_cast<float*>(ckpt); a.g_y1 = g_y1; a.g_ks = g_ks; a.x_y1 = x_y1;
    a.x_klast = x_klast; a.x_y = x_y; a.g_y = g_y; a.g_k0 = g_k0; a.gfold = gfold; a.B = B;
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    const size_t smem = (size_t)(L.total_ext + SW_WARPS * SW_NACC * 32 + SW_WARPS * SW_D * 32 * SW_TSTRIDE) * 4;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = snw_grid(B, 3);
    if (net->branch == FETODE_BRANCH_FERRO2D) {
        auto k = smallnet_wide_step_bwd_kernel<FETODE_BRANCH_FERRO2D>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    } else {
        auto k = smallnet_wide_step_bwd_kernel<FETODE_BRANCH_LOGISTIC>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<grid, SW_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

// ---- the whole dopri5 solve on the device (B <= 12) -------------------------------------------------------------------
static void snw_dp5_tableau(SWSolveArgs& a) {
    static const double A[7][6] = {
        {0, 0, 0, 0, 0, 0},
        {1.0 / 5, 0, 0, 0, 0, 0},
        {3.0 / 40, 9.0 / 40, 0, 0, 0, 0},
        {44.0 / 45, -56.0 / 15, 32.0 / 9, 0, 0, 0},
        {19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729, 0, 0},
        {9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656, 0},
        {35.0 / 384, 0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84}};
    static const double E[7] = {35.0 / 384 - 1951.0 / 21600, 0, 500.0 / 1113 - 22642.0 / 50085, 125.0 / 192 - 451.0 / 720,
                                -2187.0 / 6784 + 12231.0 / 42400, 11.0 / 84 - 649.0 / 6300, -1.0 / 60};
    static const double M[7] = {6025192743.0 / 30085553152.0 / 2, 0, 51252292925.0 / 65400821598.0 / 2,
                                -2691868925.0 / 45128329728.0 / 2, 187940372067.0 / 1594534317056.0 / 2,
                                -1776094331.0 / 19743644256.0 / 2, 11237099.0 / 235043384.0 / 2};
    for (int i = 0; i < 7; ++i) {
        a.E[i] = (float)E[i];
        a.cmid[i] = M[i];
        for (int j = 0; j < 7; ++j) a.A[i][j] = j < 6 ? (float)A[i][j] : 0.f;
    }
}

extern "C" int32_t fetode_smallnet_dp5_max_rows(void) { return SW_SOLVE_MAXB; }

extern "C" int64_t fetode_smallnet_dp5_ckpt_floats(const fetode_smallnet_t* net, int64_t B, int32_t cap) {
    if (!snw_has(net) || B < 0 || B > SW_SOLVE_MAXB || cap < 0) return -1;
    return (int64_t)cap * SW_MAXS * B * SW_REC;
}

extern "C" int fetode_smallnet_dp5_solve_fwd(const fetode_smallnet_t* net, const float* packed, float* y, float* fy,
                                             float* state, int state_is_zero, int64_t B, const double* times, int32_t T,
                                             float* out, double* ctrl, double rtol, double atol, float* ckpt, float* dts,
                                             int32_t cap, int32_t* rec_step, float* rec_w, int32_t max_attempts,
                                             void* stream) {
    if (!snw_has(net)) return FETODE_EUNSUPPORTED;
    FETODE_CHECK_ARG(packed && y && fy && times && out && ctrl && T >= 1 && B >= 1 && max_attempts >= 1);
    if (B > SW_SOLVE_MAXB) return FETODE_EUNSUPPORTED;
    if (ckpt) FETODE_CHECK_ARG(dts && rec_step && rec_w && cap >= 1);
    if (net->branch == FETODE_BRANCH_FERRO2D && !state_is_zero) FETODE_CHECK_ARG(state != nullptr);
    SWSolveArgs a{};
    snw_dp5_tableau(a);
    a.packed = packed; a.y = y; a.fy = fy; a.state = state; a.out = out; a.times = times; a.ctrl = ctrl; a.ckpt = ckpt;
    a.dts = dts; a.rec_step = rec_step; a.rec_w = rec_w; a.rtol = (float)rtol; a.atol = (float)atol; a.B = (int)B; a.T = T;
    a.cap = cap; a.max_attempts = max_attempts; a.state_is_zero = state_is_zero;
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    const size_t smem = (size_t)L.total * 4;
    cudaStream_t st = (cudaStream_t)stream;
    if (net->branch == FETODE_BRANCH_FERRO2D) {
        auto k = smallnet_wide_solve_fwd_kernel<FETODE_BRANCH_FERRO2D>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<1, SW_THREADS, smem, st>>>(a);
    } else {
        auto k = smallnet_wide_solve_fwd_kernel<FETODE_BRANCH_LOGISTIC>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<1, SW_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_smallnet_dp5_solve_bwd(const fetode_smallnet_t* net, const float* packed, const float* ckpt,
                                             const float* dts, int32_t n_steps, int32_t step_base, const int32_t* rec_step,
                                             const float* rec_w, const float* g_out, int32_t* cursor, float* carry_y,
                                             float* carry_k, float* gfold, int64_t B, void* stream) {
    if (!snw_has(net)) return FETODE_EUNSUPPORTED;
    FETODE_CHECK_ARG(packed && ckpt && dts && rec_step && rec_w && g_out && cursor && carry_y && carry_k && gfold &&
                     n_steps >= 0 && B >= 1);
    if (B > SW_SOLVE_MAXB) return FETODE_EUNSUPPORTED;
    if (n_steps == 0) return FETODE_OK;
    SWSolveArgs a{};
    snw_dp5_tableau(a);
    a.packed = packed; a.ckpt = const_cast<float*>(ckpt); a.dts = const_cast<float*>(dts); a.n_steps = n_steps;
    a.step_base = step_base; a.rec_step = const_cast<int*>(rec_step); a.rec_w = const_cast<float*>(rec_w); a.g_out = g_out;
    a.cursor = cursor; a.carry_y = carry_y; a.carry_k = carry_k; a.gfold = gfold; a.B = (int)B;
    constexpr SNLayout L = sn_layout(SW_D, SW_H, SW_NB, SW_G);
    const size_t smem = (size_t)(L.total_ext + SW_WARPS * SW_NACC * 32 + SW_WARPS * SW_D * 32 * SW_TSTRIDE) * 4;
    cudaStream_t st = (cudaStream_t)stream;
    if (net->branch == FETODE_BRANCH_FERRO2D) {
        auto k = smallnet_wide_solve_bwd_kernel<FETODE_BRANCH_FERRO2D>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<1, SW_THREADS, smem, st>>>(a);
    } else {
        auto k = smallnet_wide_solve_bwd_kernel<FETODE_BRANCH_LOGISTIC>;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        k<<<1, SW_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}
