// smallnet_bwd.cu -- reverse sweep of the fused small-width KAN / KANFET Runge-Kutta loop (sm_100a).
//
// Replaces what loss.backward() does in the Lotka-Volterra scripts (train_kanfet_node_predprey.py:254-256,
// train_kanfet_mlp_node_predprey.py:270-272): plain autograd through every solver stage of
// odeint(calDeriv, X0, t) -- discretise-then-optimise, no adjoint -- with the hysteresis snapshots prev_x
// treated as constants (ferro_class.py:553-554).
//
// One thread owns one trajectory.  For every step, last to first: reload (y, hysteresis state) from the
// checkpoint the forward kernel wrote, recompute the stage inputs, then pull the adjoint back through the
// stages.  Parameter gradients are the expensive part: each thread produces ~1000 per-parameter
// contributions per stage.  They are reduced across the warp 32 columns at a time by recursive halving
// (warp_colsum32: 31 shuffles per 32 columns, lane l ends up owning column l) and accumulated with
// shared-memory atomics into one per-CTA buffer that is flushed to HBM once, at the end of the launch.
// Gradients are produced for the FOLDED weights (same [I*F, O] + gsum layout as kanlinear_bwd.cu) and
// for the raw branch parameters (k, Ec | a, b); fetode_smallnet_unfold applies the chain rule back to
// the reference's parameters with the same kernel the any-width path uses.
#include "smallnet.cuh"

namespace fetode {

struct SNBwdArgs {
    const float* packed;
    const float* ckpt;       // [n_steps, D + SF, B]
    const float* g_yout;     // [B, D] or null
    const float* g_traj;     // [n_steps / dump_every, B, D] or null
    float* g_y0;             // [B, D]
    float* gfold;            // folded gradient accumulators (see sn_gfold_layout)
    long long B;
    float dt;
    float A[4][4];           // stage s input = y + dt sum_r A[s][r] k_r
    float bw[4];             // y_next = y + dt sum_s bw[s] k_s
    int H, nst, n_steps, dump_every;
};

struct SNGfold { int ws1, gsum1, ws2, gsum2, bp1, bp2, total; };
__host__ __device__ inline SNGfold sn_gfold_layout(int D, int H, int F, int NB) {
    SNGfold g{};
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

template <int NCOL, class Fn>
__device__ __forceinline__ void reduce_columns(Fn&& colfn, float* dst, int lane) {
    constexpr int NG = (NCOL + 31) / 32;
#pragma unroll
    for (int g = 0; g < NG; ++g) {
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = (g * 32 + i < NCOL) ? colfn(g * 32 + i) : 0.f;
        const float t = warp_colsum32(v, lane);
        if (g * 32 + lane < NCOL) atomicAdd(dst + g * 32 + lane, t);
    }
}

// dense spline feature s from the local values: v[c] sits at s = m - 3 + c (m < 0: outside the grid)
__device__ __forceinline__ float spl_at(int s, int m, const float (&v)[4]) {
    float r = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) r = (m >= 0 && s == m - 3 + c) ? v[c] : r;
    return r;
}

// Pull the adjoint `gout` of one right-hand-side evaluation out = KAN(y) back to its input (`gy`, overwritten)
// and add this trajectory's parameter-gradient contributions to the per-CTA accumulators `sacc`.
// prev1 / prev2 are the hysteresis snapshots the forward evaluation of this stage saw (read only).
//   sacc layout: per hidden unit h a block of NCH columns
//                  [ W2: f*DV+o | bp2: p0 j, p1 j | W1: d*F+f ]            at sacc + h * NCHP
//                then per input d  [ bp1: p0 j, p1 j ]  at sacc + H*NCHP + d*32*NG1, then gsum2[DV], gsum1[H]
template <int D, int NB, int G, int BR>
struct SNAcc {
    static constexpr SNLayout L = sn_layout(D, 1, NB, G);
    static constexpr int F = 1 + L.S + NB;
    static constexpr int NCH = F * L.DV + 2 * NB + D * F;
    static constexpr int NCHP = round_up(NCH, 32);
    static constexpr int NT = 2 * NB > 0 ? round_up(2 * NB, 32) : 0;       // per input d
    __host__ __device__ static int tail(int H) { return H * NCHP; }
    __host__ __device__ static int gsum2(int H) { return H * NCHP + D * NT; }
    __host__ __device__ static int gsum1(int H) { return gsum2(H) + 4; }
    __host__ __device__ static int total(int H) { return round_up(gsum1(H) + H, 4); }
};

template <int D, int NB, int G, int BR>
__device__ __forceinline__ void sn_rhs_bwd(const float* __restrict__ sm, const float* __restrict__ raw1,
                                           const float* __restrict__ raw2, int H, const float (&y)[D],
                                           const float (&prev1)[D], const float* prev2, const float (&gout)[D],
                                           float (&gy)[D], float* sacc, int lane) {
    using A = SNAcc<D, NB, G, BR>;
    constexpr SNLayout L = A::L;
    constexpr int NBR = NB > 0 ? NB : 1;
    constexpr int DV = L.DV, FU = L.FU, NKP = L.NKP, SPP = L.SPP, NK = L.NK, S = L.S, F = A::F;
    const float gl1 = sm[L.hdr], gl2 = sm[L.hdr + 1], al1 = sm[L.hdr + 2], al2 = sm[L.hdr + 3];
    // ---- layer-1 features (same arithmetic as sn_rhs) and spline derivatives
    float silu1[D], sg1[D], omu1[D], v1[D][4], dv1[D][4], r1[D][NBR];
    int m1[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float x = y[d];
        sg1[d] = rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
        silu1[d] = x * sg1[d];
        omu1[d] = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D) omu1[d] = rcp_approx(1.0f + ex2_approx(gl1 * (x - prev1[d])));
        const float4 gi = reinterpret_cast<const float4*>(sm + L.l1_gi)[d];
        m1[d] = -1;
#pragma unroll
        for (int c = 0; c < 4; ++c) { v1[d][c] = 0.f; dv1[d][c] = 0.f; }
        if (x >= gi.x && x < gi.z) {
            const float* kp = sm + L.l1_kn + d * NKP;
            const int m = bspline_interval(x, kp, gi.x, gi.y, NK);
            bspline3_local_d(x, kp, sm + L.l1_rd1 + d * NKP, sm + L.l1_rd2 + d * NKP, sm + L.l1_rd3 + d * NKP,
                             m + 3, v1[d], dv1[d]);
            m1[d] = m;
        }
        branch_eval<NB, BR>(x, omu1[d], gl1, reinterpret_cast<const float4*>(sm + L.l1_bp) + d * NB, r1[d]);
    }
    float U1u[D][1 + NBR], U1s[D][4];          // sum_h g_h * W1 row: d(loss)/d(feature) of layer 1
#pragma unroll
    for (int d = 0; d < D; ++d) {
#pragma unroll
        for (int j = 0; j < 1 + NBR; ++j) U1u[d][j] = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) U1s[d][c] = 0.f;
    }
    float go[DV];
#pragma unroll
    for (int o = 0; o < DV; ++o) go[o] = o < D ? gout[o] : 0.f;

    const float* rec = sm + L.rec0;
#pragma unroll 1
    for (int h = 0; h < H; ++h, rec += L.REC) {
        // ---- layer 1, output h (forward recompute)
        float acc = rec[L.r_b1];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float* wu = rec + L.r_wu1 + d * FU;
            acc = fmaf(wu[0], silu1[d], acc);
#pragma unroll
            for (int j = 0; j < NB; ++j) acc = fmaf(wu[1 + j], r1[d][j], acc);
            if (m1[d] >= 0) {
                const float* ws = rec + L.r_ws1 + d * SPP + m1[d];
#pragma unroll
                for (int c = 0; c < 4; ++c) acc = fmaf(ws[c], v1[d][c], acc);
            }
        }
        // ---- layer 2, input h: features, derivatives, d(loss)/d(x2)
        const float x = acc;
        const float sg = rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
        const float sl = x * sg;
        float omu = 0.f;
        if constexpr (BR == FETODE_BRANCH_FERRO2D) omu = rcp_approx(1.0f + ex2_approx(gl2 * (x - prev2[h * SN_THREADS])));
        float r2[NBR], cn2[NBR];
        branch_eval<NB, BR>(x, omu, gl2, reinterpret_cast<const float4*>(rec + L.r_bp), r2,
                            BR == FETODE_BRANCH_FERRO2D ? cn2 : nullptr);
        const float* wu2 = rec + L.r_wu2;
        float G0 = 0.f;
#pragma unroll
        for (int o = 0; o < DV; ++o) G0 = fmaf(go[o], wu2[o], G0);
        float gx2 = G0 * sg * fmaf(x, 1.0f - sg, 1.0f);
        int m2 = -1;
        float v2[4] = {0.f, 0.f, 0.f, 0.f};
        {
            const float4 gi = *reinterpret_cast<const float4*>(rec + L.r_gi);
            if (x >= gi.x && x < gi.z) {
                const float* kp = rec + L.r_kn;
                m2 = bspline_interval(x, kp, gi.x, gi.y, NK);
                float dv2[4];
                bspline3_local_d(x, kp, rec + L.r_rd1, rec + L.r_rd2, rec + L.r_rd3, m2 + 3, v2, dv2);
                const float* ws = rec + L.r_ws2 + m2 * DV;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float Gs = 0.f;
#pragma unroll
                    for (int o = 0; o < DV; ++o) Gs = fmaf(go[o], ws[c * DV + o], Gs);
                    gx2 = fmaf(dv2[c], Gs, gx2);
                }
            }
        }
        float c0[NBR], c1[NBR];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            float Gj = 0.f;
#pragma unroll
            for (int o = 0; o < DV; ++o) Gj = fmaf(go[o], wu2[(1 + j) * DV + o], Gj);
            const float4 bp = reinterpret_cast<const float4*>(rec + L.r_bp)[j];
            float gx;
            branch_bwd<BR, true>(x, omu, gl2, al2, bp.x, r2[j], Gj, raw2[h * NB + j], raw2[(H + h) * NB + j], gx, c0[j],
                                 c1[j], cn2[j]);
            gx2 += gx;
        }
        // ---- layer 1, output h: d(loss)/d(features of layer 1)
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float* wu = rec + L.r_wu1 + d * FU;
#pragma unroll
            for (int j = 0; j < 1 + NB; ++j) U1u[d][j] = fmaf(gx2, wu[j], U1u[d][j]);
            if (m1[d] >= 0) {
                const float* ws = rec + L.r_ws1 + d * SPP + m1[d];
#pragma unroll
                for (int c = 0; c < 4; ++c) U1s[d][c] = fmaf(gx2, ws[c], U1s[d][c]);
            }
        }
        // ---- parameter-gradient columns of this hidden unit
        auto col = [&](int c) -> float {
            if (c < F * DV) {
                const int f = c / DV, o = c % DV;
                const float ft = f == 0 ? sl : (f <= S ? spl_at(f - 1, m2, v2) : r2[(f - 1 - S) < NBR ? (f - 1 - S) : 0]);
                return ft * go[o];
            }
            c -= F * DV;
            if (c < NB) return c0[c < NBR ? c : 0];
            c -= NB;
            if (c < NB) return c1[c < NBR ? c : 0];
            c -= NB;
            const int d = (c / F) < D ? (c / F) : 0, f = c % F;
            const float ft = f == 0 ? silu1[d] : (f <= S ? spl_at(f - 1, m1[d], v1[d]) : r1[d][(f - 1 - S) < NBR ? (f - 1 - S) : 0]);
            return ft * gx2;
        };
        reduce_columns<A::NCH>(col, sacc + h * A::NCHP, lane);
        const float gs = warp_sum(gx2);
        if (lane == 0) atomicAdd(sacc + A::gsum1(H) + h, gs);
    }
    // ---- layer-1 input gradient and branch-parameter columns
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float x = y[d];
        float g = U1u[d][0] * sg1[d] * fmaf(x, 1.0f - sg1[d], 1.0f);
#pragma unroll
        for (int c = 0; c < 4; ++c) g = fmaf(U1s[d][c], dv1[d][c], g);
        if constexpr (NB > 0) {
            float c0[NBR], c1[NBR];
            const float ex1 = BR == FETODE_BRANCH_FERRO2D ? ferro_ex(x, gl1) : 0.f;
#pragma unroll
            for (int j = 0; j < NB; ++j) {
                const float4 bp = reinterpret_cast<const float4*>(sm + L.l1_bp)[d * NB + j];
                float gx;
                branch_bwd<BR>(x, omu1[d], gl1, al1, ex1 * bp.x, r1[d][j], U1u[d][1 + j], raw1[d * NB + j],
                               raw1[(D + d) * NB + j], gx, c0[j], c1[j]);
                g += gx;
            }
            auto col = [&](int c) -> float { return c < NB ? c0[c < NBR ? c : 0] : c1[(c - NB) < NBR && c >= NB ? (c - NB) : 0]; };
            reduce_columns<2 * NB>(col, sacc + A::tail(H) + d * A::NT, lane);
        }
        gy[d] = g;
    }
#pragma unroll
    for (int o = 0; o < D; ++o) {
        const float gs = warp_sum(gout[o]);
        if (lane == 0) atomicAdd(sacc + A::gsum2(H) + o, gs);
    }
}

// MINB = 4 caps the kernel at 128 registers (a few spills) so that four CTAs fit an SM: B = 65 536 rows is 512 CTAs, which
// then run as ONE wave (3 CTAs per SM at 168 registers needed a second, nearly empty wave: 11.5 -> 10.1 ms per LV
// training step).  Small batches are a single serial lane chain per warp: they take the spill-free build (MINB = 3).
template <int D, int NB, int G, int BR, int MINB>
__global__ void __launch_bounds__(SN_THREADS, MINB) smallnet_rk_bwd_kernel(const SNBwdArgs a) {
    extern __shared__ float4 sn_smem4[];
    float* sm = reinterpret_cast<float*>(sn_smem4);
    using A = SNAcc<D, NB, G, BR>;
    constexpr SNLayout L0 = sn_layout(D, 0, NB, G);
    constexpr bool FERRO = BR == FETODE_BRANCH_FERRO2D;
    const int H = a.H;
    const int total = L0.total + H * L0.REC;
    const int total_ext = round_up(total + 2 * D * NB + 2 * H * NB, 4);
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < total_ext / 4; i += SN_THREADS)
        sn_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    const float* raw1 = sm + total;
    const float* raw2 = raw1 + 2 * D * NB;
    float* slots = sm + total_ext;                       // [4][H][SN_THREADS] previous layer-2 inputs per stage
    float* sacc = slots + 4 * H * SN_THREADS;
    const int nacc = A::total(H);
    for (int i = tid; i < nacc; i += SN_THREADS) sacc[i] = 0.f;
    __syncthreads();

    const long long b = (long long)blockIdx.x * SN_THREADS + tid;
    const bool live = b < a.B;
    const long long bb = live ? b : 0;
    const int SF = FERRO ? D + H : 0, NC = D + SF;
    float* myslot = slots + tid;
    const int SS = H * SN_THREADS;                      // stride between stage slots

    float gy[D];
#pragma unroll
    for (int d = 0; d < D; ++d) gy[d] = (live && a.g_yout) ? a.g_yout[bb * D + d] : 0.f;

#pragma unroll 1
    for (int step = a.n_steps - 1; step >= 0; --step) {
        if (a.g_traj && (step + 1) % a.dump_every == 0 && live) {
            const float* t = a.g_traj + ((size_t)((step + 1) / a.dump_every - 1) * a.B + bb) * D;
#pragma unroll
            for (int d = 0; d < D; ++d) gy[d] += t[d];
        }
        // ---- state at the start of the step
        const float* c = a.ckpt + ((size_t)step * NC) * a.B + bb;
        float ys[4][D], p1[4][D], ks[4][D];
#pragma unroll
        for (int d = 0; d < D; ++d) { ys[0][d] = c[(size_t)d * a.B]; p1[0][d] = 0.f; }
        if constexpr (FERRO) {
#pragma unroll
            for (int d = 0; d < D; ++d) p1[0][d] = c[(size_t)(D + d) * a.B];
#pragma unroll 1
            for (int h = 0; h < H; ++h) myslot[h * SN_THREADS] = c[(size_t)(2 * D + h) * a.B];
        }
        // ---- forward recompute of the stage inputs (and of every stage's hysteresis snapshot)
#pragma unroll 1
        for (int s = 0; s + 1 < a.nst; ++s) {
            float pv[D];
#pragma unroll
            for (int d = 0; d < D; ++d) pv[d] = p1[s][d];
            if constexpr (FERRO) {
#pragma unroll 1
                for (int h = 0; h < H; ++h) myslot[(s + 1) * SS + h * SN_THREADS] = myslot[s * SS + h * SN_THREADS];
            }
            sn_rhs<D, NB, G, BR>(sm, H, ys[s], pv, myslot + (s + 1) * SS, ks[s]);
#pragma unroll
            for (int d = 0; d < D; ++d) {
                float acc = 0.f;
                for (int r = 0; r <= s; ++r) acc = fmaf(a.A[s + 1][r], ks[r][d], acc);
                ys[s + 1][d] = fmaf(a.dt, acc, ys[0][d]);
                p1[s + 1][d] = ys[s][d];
            }
        }
        // ---- reverse sweep over the stages
        float gys[4][D], gnext[D];
#pragma unroll
        for (int d = 0; d < D; ++d) gnext[d] = gy[d];
#pragma unroll 1
        for (int s = a.nst - 1; s >= 0; --s) {
            float gk[D], yin[D], pin[D], gin[D];
#pragma unroll
            for (int d = 0; d < D; ++d) {
                float acc = a.bw[s] * gy[d];
                for (int q = s + 1; q < a.nst; ++q) acc = fmaf(a.A[q][s], gys[q][d], acc);
                gk[d] = a.dt * acc;
                yin[d] = ys[s][d];
                pin[d] = p1[s][d];
            }
            sn_rhs_bwd<D, NB, G, BR>(sm, raw1, raw2, H, yin, pin, myslot + s * SS, gk, gin, sacc, lane);
#pragma unroll
            for (int d = 0; d < D; ++d) { gys[s][d] = gin[d]; gnext[d] += gin[d]; }
        }
#pragma unroll
        for (int d = 0; d < D; ++d) gy[d] = gnext[d];
    }
    if (live) {
#pragma unroll
        for (int d = 0; d < D; ++d) a.g_y0[bb * D + d] = gy[d];
    }
    // ---- flush the CTA's accumulators into the folded-gradient buffer
    __syncthreads();
    constexpr int F = A::F, DV = A::L.DV;
    const SNGfold gf = sn_gfold_layout(D, H, F, NB);
    for (int i = tid; i < nacc; i += SN_THREADS) {
        const float v = sacc[i];
        if (v == 0.f) continue;
        int dst = -1;
        if (i < A::tail(H)) {
            const int h = i / A::NCHP;
            int col = i % A::NCHP;
            if (col < F * DV) {
                const int f = col / DV, o = col % DV;
                if (o < D) dst = gf.ws2 + (h * F + f) * D + o;
            } else if ((col -= F * DV) < 2 * NB) {
                dst = gf.bp2 + (col / (NB > 0 ? NB : 1)) * H * NB + h * NB + col % (NB > 0 ? NB : 1);
            } else if ((col -= 2 * NB) < D * F) {
                dst = gf.ws1 + col * H + h;              // col = d*F + f
            }
        } else if (i < A::gsum2(H)) {
            if constexpr (NB > 0) {
                const int r = i - A::tail(H), d = r / A::NT, col = r % A::NT;
                if (col < 2 * NB) dst = gf.bp1 + (col / NB) * D * NB + d * NB + col % NB;
            }
        } else if (i < A::gsum1(H)) {
            const int o = i - A::gsum2(H);
            if (o < D) dst = gf.gsum2 + o;
        } else if (i - A::gsum1(H) < H) {
            dst = gf.gsum1 + (i - A::gsum1(H));
        }
        if (dst >= 0) atomicAdd(a.gfold + dst, v);
    }
}

template <int D, int NB, int BR>
static int snb_launch_k(const SNBwdArgs& a, cudaStream_t st) {
    using A = SNAcc<D, NB, 5, BR>;
    constexpr SNLayout L0 = sn_layout(D, 0, NB, 5);
    const int total_ext = round_up(L0.total + a.H * L0.REC + 2 * D * NB + 2 * a.H * NB, 4);
    const size_t smem = (size_t)(total_ext + 4 * a.H * SN_THREADS + A::total(a.H)) * 4;
    if (smem > 227 * 1024) return FETODE_EUNSUPPORTED;
    const unsigned grid = (unsigned)((a.B + SN_THREADS - 1) / SN_THREADS);
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (grid > (unsigned)(3 * sms)) {
        auto kern = smallnet_rk_bwd_kernel<D, NB, 5, BR, 4>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, SN_THREADS, smem, st>>>(a);
    } else {
        auto kern = smallnet_rk_bwd_kernel<D, NB, 5, BR, 3>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        kern<<<grid, SN_THREADS, smem, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

template <int D>
static int snb_launch_d(int branch, const SNBwdArgs& a, cudaStream_t st) {
    switch (branch) {
        case FETODE_BRANCH_NONE: return snb_launch_k<D, 0, FETODE_BRANCH_NONE>(a, st);
        case FETODE_BRANCH_LOGISTIC: return snb_launch_k<D, 10, FETODE_BRANCH_LOGISTIC>(a, st);
        case FETODE_BRANCH_FERRO2D: return snb_launch_k<D, 10, FETODE_BRANCH_FERRO2D>(a, st);
    }
    return FETODE_EUNSUPPORTED;
}

// adds the raw branch-parameter gradients gathered by the sweep into the caller's buffers
__global__ void snb_bp_add_kernel(const float* src, float* g0, float* g1, int n) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    if (g0) g0[t] += src[t];
    if (g1) g1[t] += src[n + t];
}

int kl_unfold_launch(const fetode_kanlayer_t* P, const fetode_kanlayer_grads_t* grads, float* ws, cudaStream_t st);

}  // namespace fetode

using namespace fetode;

extern "C" int64_t fetode_smallnet_gfold_floats(const fetode_smallnet_t* net) {
    if (sn_check(net) != FETODE_OK) return -1;
    const int F = 1 + net->grid_size + 3 + net->num_basis;
    return sn_gfold_layout(net->widths[0], net->widths[1], F, net->num_basis).total;
}

extern "C" int fetode_smallnet_rk_bwd(const fetode_smallnet_t* net, const float* packed, int tableau,
                                      const float* ckpt, const float* g_yout, const float* g_traj,
                                      int32_t dump_every, float* g_y0, float* gfold, int64_t B, double dt,
                                      int32_t n_steps, void* stream) {
    int rc = sn_check(net);
    if (rc) return rc;
    FETODE_CHECK_ARG(packed && g_y0 && gfold && B >= 0 && n_steps >= 0);
    FETODE_CHECK_ARG(n_steps == 0 || ckpt != nullptr);
    FETODE_CHECK_ARG(tableau >= FETODE_TABLEAU_EULER && tableau <= FETODE_TABLEAU_RK4_38);
    if (g_traj) FETODE_CHECK_ARG(dump_every >= 1);
    if (B == 0) return FETODE_OK;
    SNBwdArgs a{};
    a.packed = packed; a.ckpt = ckpt; a.g_yout = g_yout; a.g_traj = g_traj; a.g_y0 = g_y0; a.gfold = gfold;
    a.B = B; a.dt = (float)dt; a.H = net->widths[1]; a.n_steps = n_steps; a.dump_every = dump_every > 0 ? dump_every : 1;
    switch (tableau) {
        case FETODE_TABLEAU_EULER: a.nst = 1; a.bw[0] = 1.f; break;
        case FETODE_TABLEAU_MIDPOINT: a.nst = 2; a.A[1][0] = 0.5f; a.bw[1] = 1.f; break;
        case FETODE_TABLEAU_RK4:
            a.nst = 4; a.A[1][0] = 0.5f; a.A[2][1] = 0.5f; a.A[3][2] = 1.f;
            a.bw[0] = a.bw[3] = (float)(1.0 / 6.0); a.bw[1] = a.bw[2] = (float)(1.0 / 3.0);
            break;
        default:
            a.nst = 4; a.A[1][0] = (float)(1.0 / 3.0); a.A[2][0] = (float)(-1.0 / 3.0); a.A[2][1] = 1.f;
            a.A[3][0] = 1.f; a.A[3][1] = -1.f; a.A[3][2] = 1.f;
            a.bw[0] = a.bw[3] = 0.125f; a.bw[1] = a.bw[2] = 0.375f;
            break;
    }
    cudaStream_t st = (cudaStream_t)stream;
    switch (net->widths[0]) {
        case 1: return snb_launch_d<1>(net->branch, a, st);
        case 2: return snb_launch_d<2>(net->branch, a, st);
        case 3: return snb_launch_d<3>(net->branch, a, st);
        case 4: return snb_launch_d<4>(net->branch, a, st);
    }
    return FETODE_EUNSUPPORTED;
}

extern "C" int fetode_smallnet_unfold(const fetode_smallnet_t* net, const fetode_kanlayer_t* layers,
                                      float* gfold, const fetode_kanlayer_grads_t* grads, void* stream) {
    int rc = sn_check(net);
    if (rc) return rc;
    FETODE_CHECK_ARG(layers && gfold && grads);
    const int D = net->widths[0], H = net->widths[1], NB = net->num_basis;
    const int F = 1 + net->grid_size + 3 + NB;
    const SNGfold gf = sn_gfold_layout(D, H, F, NB);
    cudaStream_t st = (cudaStream_t)stream;
    rc = kl_unfold_launch(&layers[0], &grads[0], gfold + gf.ws1, st);
    if (rc) return rc;
    rc = kl_unfold_launch(&layers[1], &grads[1], gfold + gf.ws2, st);
    if (rc) return rc;
    if (NB > 0) {
        if (grads[0].bp[0] || grads[0].bp[1])
            snb_bp_add_kernel<<<(D * NB + 127) / 128, 128, 0, st>>>(gfold + gf.bp1, grads[0].bp[0], grads[0].bp[1], D * NB);
        if (grads[1].bp[0] || grads[1].bp[1])
            snb_bp_add_kernel<<<(H * NB + 127) / 128, 128, 0, st>>>(gfold + gf.bp2, grads[1].bp[0], grads[1].bp[1], H * NB);
    }
    return FETODE_LAUNCH_RC();
}
