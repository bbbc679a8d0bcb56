// kanlinear_tc_dx.cu -- input gradient of the wide KANLinear layers on the tensor cores (sm_100a, tcgen05 + TMEM).
//
// Replaces, in the autograd graph PyTorch records for KANLinear.forward (efficient_kan/efficientkan.py:160-182;
// loss.backward() at train_ecg_kan_fet_nn_ode.py:1108 / train_kan_fet_ett.py), the contraction
//
//   G[b,(i,f)] = sum_o gy[b,o] wcat[(i,f),o]                (gradient w.r.t. the never-materialised feature matrix)
//
// and everything that consumes it:  gx[b,i] = sum_f G[b,(i,f)] dA_f/dx(x_bi)  (silu', cubic B-spline derivative,
// branch-basis derivative) and the gradients of the per-input branch parameters (LogisticBasis a,b efficientkan.py:17-24;
// 2-D ferro k,Ec ferro_class.py:537-578), which are reductions of G over the batch rows.  G lives only in TMEM.
//
// CTA = 128 rows x one 128-wide slice of the outputs x a contiguous range of 4-input chunks.
//   D[128 rows, NCOL = 4 inputs x FP features] = gy[128, O] . W^T : both operands K-major (K = outputs), SWIZZLE_128B.
//   gy is loaded once, split hi/lo (3xTF32: hi*hi + hi*lo + lo*hi, see kanlinear_tc.cu) and stays resident
//   (<= 128 KB); the weight chunks ([NCOL x 32 outputs] hi+lo tiles, pre-split / pre-swizzled by kl_td_pack_kernel)
//   stream through a ring via the TMA bulk-copy engine.  The contraction is only O <= 128 long (48 MMAs), so one
//   accumulator per chunk is enough.  Two consecutive chunks share every MMA (N = 2 NCOL: the kernel is bound by the
//   shared-memory operand reads, and this reads the 128-row gy operand once per 8 inputs instead of once per 4);
//   two such double-width accumulators alternate in TMEM.
//   warps 0-15  gy prologue, then epilogue: set (warp >> 2) takes every 4th chunk (the per-row derivative
//               evaluation is a latency-bound chain, so four sets run concurrently); thread = row (TMEM lane):
//               tcgen05.ld of the FP values of one input, derivative evaluation, gx store, branch-parameter partial
//               sums reduced across the warp with a recursive-halving shuffle (warp_colsum) and across the CTA's
//               four row quarters in shared memory, one global atomic per (input, parameter) per CTA
//   warp 16     lane 0: cp.async.bulk of the weight stages; whole warp: TMEM alloc / dealloc
//   warp 17     lane 0: tcgen05.mma issue
#include <cstdlib>
#include "tc_common.cuh"

namespace fetode {
namespace tcx {

using namespace tc;

constexpr int NSW = 2;                  // weight-stage ring depth (a stage = one K slab of a two-chunk group)
constexpr int NACC = 2;                 // TMEM accumulators in flight, each two chunks (8 inputs) wide
constexpr int NEPI = 4;                 // epilogue sets of 4 warps; set s owns chunks s, s + NEPI, .. (accumulator s)
static_assert(NACC % NEPI == 0 || NEPI % NACC == 0, "every accumulator must belong to one epilogue set");
constexpr uint32_t GY_SLAB = BM * 32 * 4;      // one [128 x 32] K-slab of gy (hi or lo): 16 KB

struct DXArgs {
    fetode_kanlayer_t P;
    const float* blob;
    const float* wtd;       // dX weight image
    const float* x;
    const float* prev;
    const float* gy;
    float* gx;              // [B, I] or null
    float* gbp0;            // gradient of bp[0] (a | k) or null
    float* gbp1;            // gradient of bp[1] (b | Ec) or null
    long long B;
    KLLayout L;
    float gl;
    int prev_is_zero, chunks_per_split;
};

template <int N>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, float* v);
template <>
__device__ __forceinline__ void tmem_ld<4>(uint32_t taddr, float* v) {
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] = __uint_as_float(r[e]);
}
template <>
__device__ __forceinline__ void tmem_ld<8>(uint32_t taddr, float* v) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[e]);
}
template <>
__device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
#pragma unroll
    for (int e = 0; e < 16; ++e) v[e] = __uint_as_float(r[e]);
}

// Column sums across a warp of N (power of two <= 32) per-lane values: lane l returns the total of v[l & (N-1)].
template <int N>
__device__ __forceinline__ float warp_colsum(float* v, int lane) {
#pragma unroll
    for (int off = N / 2; off >= 1; off >>= 1) {
        const bool hi = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const float send = hi ? v[i] : v[i + off];
            const float keep = hi ? v[i + off] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    float r = v[0];
#pragma unroll
    for (int off = N; off < 32; off <<= 1) r += __shfl_xor_sync(0xffffffffu, r, off);
    return r;
}

__device__ __forceinline__ void bar_sync_named(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// One (row, input) pair: g[0..F) = G[b,(i,.)] -> returns gx[b,i]; pv[0..NB) / pv[NB..2NB) receive this row's
// contribution to the gradient of bp[0][i,.] / bp[1][i,.].  Same formulas as kl_bwd_dx_kernel (kanlinear_bwd.cu).
template <int NB, int BR>
__device__ __forceinline__ float dx_input(const float* g, float x, float prev, const float* __restrict__ tab,
                                          const KLLayout& L, const DXArgs& a, int i, float* pv) {
    float gx;
    {   // silu'(x) = s (1 + x (1 - s))
        const float s = rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
        gx = g[0] * s * fmaf(x, 1.0f - s, 1.0f);
    }
    const float4 gi = __ldg(reinterpret_cast<const float4*>(tab + L.t_gi));
    if (x >= gi.x && x < gi.z) {
        const float* kp = tab + L.t_kn;
        const float* rd1 = tab + L.t_rd1; const float* rd2 = tab + L.t_rd2; const float* rd3 = tab + L.t_rd3;
        const int m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
        const int mp = m + 3;
        // order-2 values N^2_{m-2..m}, then N'^3_j = 3 (N^2_j rd3[j] - N^2_{j+1} rd3[j+1])
        const float left1 = x - kp[mp], right1 = kp[mp + 1] - x;
        const float left2 = x - kp[mp - 1], right2 = kp[mp + 2] - x;
        float t = rd1[mp];
        const float a0 = right1 * t, a1 = left1 * t;
        t = a0 * rd2[mp - 1];
        const float b0 = right1 * t;
        const float saved = left2 * t;
        t = a1 * rd2[mp];
        const float b1 = saved + right2 * t;
        const float b2 = left1 * t;
        const float e0 = 3.0f * b0 * rd3[mp - 2], e1 = 3.0f * b1 * rd3[mp - 1], e2 = 3.0f * b2 * rd3[mp];
        const float dv[4] = {-e0, e0 - e1, e1 - e2, e2};
#pragma unroll
        for (int s = 0; s < 8; ++s) {
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) d = (s == m - 3 + c) ? dv[c] : d;
            gx = fmaf(g[1 + s], d, gx);
        }
    }
    if constexpr (BR == FETODE_BRANCH_LOGISTIC) {
        const float* pa = a.P.bp[0] + (size_t)i * NB;
        const float* pb = a.P.bp[1] + (size_t)i * NB;
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const float aj = __ldg(pa + j), bj = __ldg(pb + j);
            const float r = rcp_approx(1.0f + ex2_approx(fminf(-FETODE_LOG2E * aj * (x - bj), FETODE_EXP_CLAMP)));
            const float e = g[9 + j] * r * (1.0f - r);
            gx = fmaf(e, aj, gx);
            pv[j] = e * (x - bj);
            pv[NB + j] = -e * aj;
        }
    } else if constexpr (BR == FETODE_BRANCH_FERRO2D) {
        const float gs = a.P.gate_slope, al2 = 2.0f * (1.0f - a.P.alpha);
        const float omu = rcp_approx(1.0f + ex2_approx(a.gl * (x - prev)));        // 1 - is_moving_up
        const float domu = -gs * omu * (1.0f - omu);
        const float* pk = a.P.bp[0] + (size_t)i * NB;
        const float* pE = a.P.bp[1] + (size_t)i * NB;
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const float k = __ldg(pk + j), Ec = __ldg(pE + j);
            const float cn = rcp_approx(1.0f + ex2_approx(fminf(a.gl * (x + Ec), FETODE_EXP_CLAMP)));
            const float dcn = -gs * cn * (1.0f - cn);                                // d/dx == d/dEc
            const float sl = omu * cn;
            const float mo = 1.0f - al2 * sl;
            const float dm_dx = -al2 * (domu * cn + omu * dcn);
            const float dm_dE = -al2 * omu * dcn;
            const float z2 = 2.0f * FETODE_LOG2E * k * fmaf(Ec, mo, x);
            const float r = rcp_approx(1.0f + ex2_approx(fminf(z2, FETODE_EXP_CLAMP)));
            const float e = g[9 + j] * (-2.0f) * r * (1.0f - r);                     // dL/dz
            gx = fmaf(e * k, fmaf(Ec, dm_dx, 1.0f), gx);
            pv[j] = e * fmaf(Ec, mo, x);
            pv[NB + j] = e * k * fmaf(Ec, dm_dE, mo);
        }
    }
    return gx;
}

template <int NB, int BR>
__global__ void __launch_bounds__(NEPI * 128 + 64, 1) kl_tc_dx_kernel(const DXArgs a) {
    constexpr int F = 9 + NB;
    constexpr int FP = NB > 0 ? 20 : 12;                // TMEM columns per input (tcgen05.ld x16+x4 / x8+x4)
    static_assert(F <= FP, "feature count exceeds the padded column count");
    constexpr int NCOL = 4 * FP;                        // MMA N: one chunk = 4 inputs
    constexpr int NV = 2 * NB;                          // branch-parameter partial sums per input
    static_assert(NV == 0 || NV == 16 || NV == 20, "warp reduction is laid out for NB = 8 or 10");
    constexpr int NCOL2 = 2 * NCOL;                     // two chunks share one MMA (N = NCOL2): the gy operand is read
                                                        // from shared memory once per 8 inputs instead of once per 4
    constexpr uint32_t W_TILE = NCOL * 32 * 4;          // one chunk, one of hi / lo
    constexpr uint32_t W_TILE2 = 2 * W_TILE;            // a stage holds [hi: chunk 0 | chunk 1][lo: chunk 0 | chunk 1]
    constexpr uint32_t TMEM_COLS = NACC * NCOL2 <= 256 ? 256u : 512u;
    constexpr int TMA_WARP = 4 * NEPI, MMA_WARP = TMA_WARP + 1;
    extern __shared__ __align__(1024) unsigned char tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    unsigned char* gbase = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const KLLayout& L = a.L;
    const int I = L.I, O = L.O;
    const int osplit = blockIdx.y;
    const int o0 = osplit * 128;
    const int KS = min(4, (O - o0 + 31) / 32);          // 32-output K slabs of this CTA
    const uint32_t offGy = 0, offW = (uint32_t)KS * 2 * GY_SLAB, offAcc = offW + NSW * 2 * W_TILE2,
                   offBar = offAcc + NEPI * 2 * 4 * (NV > 0 ? NV : 4) * 4;
    const uint32_t barWFull = base + offBar, barWEmpty = barWFull + 8 * NSW, barAccFull = barWEmpty + 8 * NSW,
                   barAccEmpty = barAccFull + 8 * NACC, barGy = barAccEmpty + 8 * NACC;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + offBar + 8 * (2 * NSW + 2 * NACC + 1));
    float* pacc = reinterpret_cast<float*>(gbase + offAcc);          // [set][parity][4 * NV]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long m0 = (long long)blockIdx.x * BM;
    const int chunk0 = blockIdx.z * a.chunks_per_split;
    const int nch = min(a.chunks_per_split, L.td_groups - chunk0);

    if (tid == 0) {
        for (int s = 0; s < NSW; ++s) { mbar_init(barWFull + 8 * s, 1); mbar_init(barWEmpty + 8 * s, 1); }
        for (int s = 0; s < NACC; ++s) { mbar_init(barAccFull + 8 * s, 1); mbar_init(barAccEmpty + 8 * s, 256); }
        mbar_init(barGy, NEPI * 128);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32((const void*)tmem_slot)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if constexpr (NV > 0)
        for (int q = tid; q < NEPI * 2 * 4 * NV; q += blockDim.x) pacc[q] = 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4 * NEPI) {
        // ===================== gy tile: [128 rows x 32*KS outputs] -> hi / lo K-major SWIZZLE_128B slabs =====================
        {
            const int c4 = KS * 8;                              // 16-byte chunks per row
            for (int q = tid; q < BM * c4; q += NEPI * 128) {
                const int r = q / c4, o4 = q % c4;
                const int o = o0 + o4 * 4;
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                if (m0 + r < a.B && o < O) {
                    const float* gp = a.gy + (m0 + r) * O + o;
                    if (o + 3 < O && (O & 3) == 0) g = __ldg(reinterpret_cast<const float4*>(gp));
                    else { g.x = gp[0]; if (o + 1 < O) g.y = gp[1]; if (o + 2 < O) g.z = gp[2]; if (o + 3 < O) g.w = gp[3]; }
                }
                float4 h;
                h.x = __uint_as_float(__float_as_uint(g.x) & 0xffffe000u); h.y = __uint_as_float(__float_as_uint(g.y) & 0xffffe000u);
                h.z = __uint_as_float(__float_as_uint(g.z) & 0xffffe000u); h.w = __uint_as_float(__float_as_uint(g.w) & 0xffffe000u);
                const uint32_t addr = base + offGy + (uint32_t)(o4 >> 3) * 2 * GY_SLAB +
                                      (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + (((o4 & 7) ^ (r & 7)) << 4));
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w) : "memory");
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr + GY_SLAB), "f"(g.x - h.x), "f"(g.y - h.y), "f"(g.z - h.z), "f"(g.w - h.w) : "memory");
            }
            fence_proxy_async();
            mbar_arrive(barGy);
        }
        // ===================== epilogue =====================
        const int set = warp >> 2;
        const int r = tid & 127;
        const long long b = m0 + r;
        const bool live = b < a.B;
        const bool has_prev = BR == FETODE_BRANCH_FERRO2D && !a.prev_is_zero;
        const bool need_bp = NV > 0 && (a.gbp0 || a.gbp1);
        auto ldx = [&](const float* p, int chunk) -> float4 {
            if (!live) return make_float4(0.f, 0.f, 0.f, 0.f);
            return __ldg(reinterpret_cast<const float4*>(p + b * I + 4 * chunk));
        };
        float4 xn = make_float4(0.f, 0.f, 0.f, 0.f), pn = xn;
        if (set < nch) {
            xn = ldx(a.x, chunk0 + set);
            if (has_prev) pn = ldx(a.prev, chunk0 + set);
        }
        const uint32_t tlane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        // set s takes chunks s, s + 4, ..: chunk lc is the (lc & 1) half of accumulator (lc >> 1) & 1, whose
        // (lc >> 2)-th use this is; the two sets sharing an accumulator both wait for it and both release it
        const int ncg = (nch + 1) / 2;
        for (int lc = set; lc < 2 * ncg; lc += NEPI) {
            const int chunk = chunk0 + lc;
            const int acc = (lc >> 1) & 1;
            if (lc >= nch) {                                    // odd tail: this half of the last group is empty
                mbar_wait(barAccFull + 8 * acc, (lc >> 2) & 1);
                tc_fence_after();
                tc_fence_before();
                mbar_arrive(barAccEmpty + 8 * acc);
                continue;
            }
            const float4 xc = xn, pc = pn;
            if (lc + NEPI < nch) {
                xn = ldx(a.x, chunk + NEPI);
                if (has_prev) pn = ldx(a.prev, chunk + NEPI);
            }
            mbar_wait(barAccFull + 8 * acc, (lc >> 2) & 1);
            tc_fence_after();
            float* pa = pacc + (set * 2 + ((lc / NEPI) & 1)) * 4 * (NV > 0 ? NV : 4);
#pragma unroll 1
            for (int ii = 0; ii < 4; ++ii) {
                float g[FP];
                const uint32_t taddr = tlane + (uint32_t)(acc * NCOL2 + (lc & 1) * NCOL + ii * FP);
                if constexpr (FP == 20) { tmem_ld<16>(taddr, g); tmem_ld<4>(taddr + 16, g + 16); }
                else { tmem_ld<8>(taddr, g); tmem_ld<4>(taddr + 8, g + 8); }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (ii == 3) { tc_fence_before(); mbar_arrive(barAccEmpty + 8 * acc); }
                const float x = ii == 0 ? xc.x : (ii == 1 ? xc.y : (ii == 2 ? xc.z : xc.w));
                const float pr = ii == 0 ? pc.x : (ii == 1 ? pc.y : (ii == 2 ? pc.z : pc.w));
                const int i = 4 * chunk + ii;
                float pv[NV > 0 ? NV : 1];
                const float gxv = dx_input<NB, BR>(g, x, pr, a.blob + L.tab + (size_t)i * L.TAB, L, a, i, pv);
                if (live && a.gx) {
                    if (gridDim.y == 1) a.gx[b * I + i] = gxv;
                    else atomicAdd(a.gx + b * I + i, gxv);
                }
                if constexpr (NV > 0) {
                    if (need_bp) {
                        float tot = warp_colsum<16>(pv, lane);
                        if constexpr (NV == 20) {
                            const float t4 = warp_colsum<4>(pv + 16, lane);
                            tot = lane < 16 ? tot : t4;
                        }
                        if (lane < NV) atomicAdd(pa + ii * NV + lane, tot);
                    }
                }
            }
            if constexpr (NV > 0) {
                if (need_bp) {
                    bar_sync_named(1 + set, 128);
                    if (r < 4 * NV) {
                        const float v = pa[r];
                        pa[r] = 0.f;
                        const int ii = r / NV, q = r % NV;
                        float* dst = q < NB ? a.gbp0 : a.gbp1;
                        if (dst) atomicAdd(dst + (size_t)(4 * chunk + ii) * NB + (q % NB), v);
                    }
                }
            }
        }
        tc_fence_before();
    } else if (warp == TMA_WARP) {
        // ===================== TMA: weight stages =====================
        if (lane == 0) {
            const size_t stage_floats = (size_t)2 * NCOL * 32;     // one chunk's [hi | lo] pair in the packed image
            const int ncg = (nch + 1) / 2;
            const int nit = ncg * KS;
            for (int it = 0; it < nit; ++it) {
                const int s = it % NSW;
                const int g = it / KS, ks = it % KS;
                const int nc = min(2, nch - 2 * g);
                mbar_wait(barWEmpty + 8 * s, ((it / NSW) & 1) ^ 1);
                mbar_arrive_expect_tx(barWFull + 8 * s, (uint32_t)nc * 2 * W_TILE);
                const uint32_t dst = base + offW + (uint32_t)s * 2 * W_TILE2;
                for (int j = 0; j < nc; ++j) {
                    const float* src = a.wtd + ((size_t)(chunk0 + 2 * g + j) * L.td_KS + (size_t)(osplit * 4 + ks)) * stage_floats;
                    bulk_g2s(dst + (uint32_t)j * W_TILE, src, W_TILE, barWFull + 8 * s);
                    bulk_g2s(dst + W_TILE2 + (uint32_t)j * W_TILE, src + NCOL * 32, W_TILE, barWFull + 8 * s);
                }
            }
        }
    } else if (warp == MMA_WARP) {
        // ===================== MMA issue =====================
        if (lane == 0) {
            const uint32_t idesc1 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NCOL >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            const uint32_t idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(NCOL2 >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            mbar_wait(barGy, 0);
            tc_fence_after();
            int it = 0;
            const int ncg = (nch + 1) / 2;
            for (int g = 0; g < ncg; ++g) {
                const int acc = g & 1;
                const uint32_t idesc = nch - 2 * g >= 2 ? idesc2 : idesc1;
                mbar_wait(barAccEmpty + 8 * acc, ((g >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)(acc * NCOL2);
                for (int ks = 0; ks < KS; ++ks, ++it) {
                    const int s = it % NSW;
                    mbar_wait(barWFull + 8 * s, (it / NSW) & 1);
                    tc_fence_after();
                    const uint32_t gHi = base + offGy + (uint32_t)ks * 2 * GY_SLAB, gLo = gHi + GY_SLAB;
                    const uint32_t wHi = base + offW + (uint32_t)s * 2 * W_TILE2, wLo = wHi + W_TILE2;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint64_t dAh = umma_desc(gHi + j * 32), dAl = umma_desc(gLo + j * 32);
                        const uint64_t dBh = umma_desc(wHi + j * 32), dBl = umma_desc(wLo + j * 32);
                        umma_tf32(d, dAh, dBh, idesc, (ks | j) != 0);
                        umma_tf32(d, dAh, dBl, idesc, 1);
                        umma_tf32(d, dAl, dBh, idesc, 1);
                    }
                    umma_commit(barWEmpty + 8 * s);
                }
                umma_commit(barAccFull + 8 * acc);
            }
        }
    }
    __syncthreads();
    if (warp == TMA_WARP) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// dX weight image: for chunk c (inputs 4c..4c+3) and K slab ks (outputs 32ks..32ks+31) one [hi | lo] pair of
// [NCOL x 32] K-major SWIZZLE_128B tiles; row p = ii * FP + f holds wcat[(4c+ii, f), 32ks + col], zero for f >= F.
__global__ void __launch_bounds__(256) kl_td_pack_kernel(KLLayout L, float* blob) {
    const int FP = L.td_FP, NCOL = 4 * FP;
    const long long per_tile = (long long)NCOL * 32;
    const long long n = (long long)L.td_groups * L.td_KS * per_tile;
    const float* wcat = blob + L.wcat;
    float* dst = blob + L.td;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long long)gridDim.x * blockDim.x) {
        const int col = (int)(idx % 32);
        const int row = (int)((idx / 32) % NCOL);
        const long long ts = idx / per_tile;                      // chunk * KS + ks
        const int ks = (int)(ts % L.td_KS);
        const long long c = ts / L.td_KS;
        const int ii = row / FP, f = row % FP;
        const int o = ks * 32 + col;
        float w = 0.f;
        if (f < L.F && o < L.O) w = wcat[((c * 4 + ii) * L.F + f) * L.O + o];
        const float hi = __uint_as_float(__float_as_uint(w) & 0xffffe000u);
        float* tile = dst + ts * 2 * per_tile;
        tile[swz(row, col)] = hi;
        tile[per_tile + swz(row, col)] = w - hi;
    }
}

template <int NB, int BR>
static int dx_launch_k(const DXArgs& a, dim3 grid, size_t smem, cudaStream_t st) {
    auto k = kl_tc_dx_kernel<NB, BR>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e) return (int)e;
    k<<<grid, NEPI * 128 + 64, smem, st>>>(a);
    return FETODE_LAUNCH_RC();
}

}  // namespace tcx

int kl_td_pack(const KLLayout& L, float* blob, cudaStream_t st) {
    if (!L.tc_ok) return FETODE_OK;
    const long long n = (long long)L.td_groups * L.td_KS * 4 * L.td_FP * 32;
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 8);
    tcx::kl_td_pack_kernel<<<blocks, 256, 0, st>>>(L, blob);
    return FETODE_LAUNCH_RC();
}

// gx = (gy W^T) . dA/dx and the branch-parameter gradients on the tensor cores.  gbp0 / gbp1 are accumulated into
// (caller zeroes them); gx is overwritten.  Returns FETODE_EUNSUPPORTED when the caller should use the CUDA-core kernel.
int kl_tc_dx(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
             int prev_is_zero, const float* gy, float* gx, float* gbp0, float* gbp1, long long B, cudaStream_t st) {
    if (!L.tc_ok) return FETODE_EUNSUPPORTED;
    if (((uintptr_t)x & 15) || (prev && ((uintptr_t)prev & 15)) || ((uintptr_t)gy & 15)) return FETODE_EUNSUPPORTED;
    tcx::DXArgs a{};
    a.P = *P; a.blob = blob; a.wtd = blob + L.td; a.x = x; a.prev = prev; a.gy = gy; a.gx = gx; a.gbp0 = gbp0; a.gbp1 = gbp1;
    a.B = B; a.L = L;
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero;
    const long long mt = (B + tc::BM - 1) / tc::BM;
    const int osplits = (L.O + 127) / 128;
    if (mt > 0x7fffffffLL || osplits > 65535) return FETODE_EUNSUPPORTED;
    long long splits = std::max(1LL, 148LL / (mt * osplits));
    splits = std::min<long long>(splits, L.td_groups);
    a.chunks_per_split = (int)((L.td_groups + splits - 1) / splits);
    splits = (L.td_groups + a.chunks_per_split - 1) / a.chunks_per_split;
    if (splits > 65535) return FETODE_EUNSUPPORTED;
    if (gx && osplits > 1) {
        cudaError_t e = cudaMemsetAsync(gx, 0, (size_t)B * L.I * sizeof(float), st);
        if (e) return (int)e;
    }
    const int KS = std::min(4, L.td_KS);
    const int NCOL = 4 * L.td_FP;
    size_t smem = (size_t)KS * 2 * tcx::GY_SLAB + (size_t)tcx::NSW * 2 * 2 * NCOL * 128 + tcx::NEPI * 2 * 4 * std::max(2 * L.NB, 4) * 4 + 512 + 1024;
    smem = std::max<size_t>(smem, 120 * 1024);              // one CTA per SM: the kernel allocates all of TMEM
    const dim3 grid((unsigned)mt, (unsigned)osplits, (unsigned)splits);
    switch (P->branch) {
        case FETODE_BRANCH_NONE: return tcx::dx_launch_k<0, FETODE_BRANCH_NONE>(a, grid, smem, st);
        case FETODE_BRANCH_LOGISTIC:
            if (L.NB == 10) return tcx::dx_launch_k<10, FETODE_BRANCH_LOGISTIC>(a, grid, smem, st);
            if (L.NB == 8) return tcx::dx_launch_k<8, FETODE_BRANCH_LOGISTIC>(a, grid, smem, st);
            break;
        case FETODE_BRANCH_FERRO2D:
            if (L.NB == 10) return tcx::dx_launch_k<10, FETODE_BRANCH_FERRO2D>(a, grid, smem, st);
            break;
    }
    return FETODE_EUNSUPPORTED;
}

}  // namespace fetode
