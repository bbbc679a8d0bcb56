// kanlinear_tc_bwd.cu -- backward of the wide KANLinear layers on the tensor cores (sm_100a, tcgen05 + TMEM).
//
// Replaces the two large contractions autograd records for KANLinear.forward (efficient_kan/efficientkan.py:160-182)
// in the ECG / ETT training configurations (train_ecg_kan_fet_nn_ode.py:1108 loss.backward(), train_kan_fet_ett.py):
//
//   kl_tc_dw_kernel   gW[(i,f), o] = sum_b A[b,(i,f)] gy[b,o]          (weight gradient, folded layout)
//       D[o, feature] += gy^T . A : both operands are "MN-major" for UMMA (contraction = batch rows), the feature
//       operand is built by the same prologue as the forward kernel and has the SAME shared-memory layout
//       (128B-swizzled [rows x 32 features] tiles); gy tiles are split hi/lo on the fly.
//
// Same 3xTF32 split as the forward (hi*hi + hi*lo + lo*hi) with the hi*hi products rotated over two TMEM
// accumulators and the corrections in a third (the tensor core truncates when it accumulates, see kanlinear_tc.cu).
// Per CTA: 8 inputs = 4 pairs, one 128-wide output tile, a contiguous range of 32-row chunks.  One chunk =
// one shared-memory buffer holding the gy tile [32 x 128] and the feature tile [32 x NCOLS] (NCOLS = 4 pairs x
// (20 spline-side + 2 NB branch-side columns)), both hi/lo; the MMA takes the whole feature tile as ONE N = NCOLS
// operand (M = 128 outputs, K = 32 rows: 12 instructions per chunk), so the 4 KB gy operand is re-read from shared
// memory once per 80-cycle instruction instead of once per 16-cycle N = 32 slice.  NPAR producer groups (spline
// role + branch role, thread = (row, pair)) own one buffer each and take every NPAR-th chunk.
#include <cstdlib>
#include "tc_common.cuh"

namespace fetode {
namespace tcb {

using namespace tc;

constexpr int RC = 32;                 // rows per chunk
constexpr uint32_t STAGE_BYTES = RC * 32 * 4;            // one [32 rows x 32 features] group, hi or lo
constexpr uint32_t GY_BYTES = RC * 128 * 4;              // one of hi / lo

struct DWArgs {
    const float* blob;
    const float* x;
    const float* prev;
    const float* gy;
    float* ws;              // [I*F, O] folded-weight gradient
    long long B;
    KLLayout L;
    float gl;
    int prev_is_zero, chunks_per_split, accumulate;
};

// shared with the forward kernel (kanlinear_tc.cu)
__device__ __forceinline__ void spline_features(float x, const float* __restrict__ tab, const KLLayout& L, float* f) {
    f[0] = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
    int m = -1;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    const float4 gi = *reinterpret_cast<const float4*>(tab + L.t_gi);
    if (x >= gi.x && x < gi.z) {
        const float* kp = tab + L.t_kn;
        m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
        bspline3_local(x, kp, tab + L.t_rd1, tab + L.t_rd2, tab + L.t_rd3, m + 3, v);
    }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        float r = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) r = (s == m - 3 + c) ? v[c] : r;
        f[1 + s] = (m >= 0) ? r : 0.f;
    }
}

template <int NB, int BR>
__device__ __forceinline__ void branch_features(float x, float prev, float gl, const float* __restrict__ tab,
                                                const KLLayout& L, float* f) {
    float omu = 0.f;
    if constexpr (BR == FETODE_BRANCH_FERRO2D) omu = rcp_approx(1.0f + ex2_approx(gl * (x - prev)));
    float r[NB > 0 ? NB : 1];
    branch_eval<NB, BR>(x, omu, gl, reinterpret_cast<const float4*>(tab + L.t_bp), r);
#pragma unroll
    for (int j = 0; j < NB; ++j) f[j] = r[j];
}



template <int NB> struct DWShape {
    static constexpr int NROLE = NB > 0 ? 2 : 1;             // spline | branch producer roles
    static constexpr int NPAR = 2;                           // producer groups = shared-memory buffers
    static constexpr int FPP = 20 + 2 * NB;                  // columns per pair of inputs
    static constexpr int NCOLS = 4 * FPP;                    // features per CTA (MMA N)
    static constexpr int NSTG = (NCOLS + 31) / 32;           // 32-feature groups per chunk
    static constexpr int NPROD = 128 * NROLE * NPAR;
    static constexpr int THREADS = NPROD + 64;
    static constexpr uint32_t BUF_BYTES = 2 * GY_BYTES + NSTG * 2 * STAGE_BYTES;
};

template <int NB, int BR>
__global__ void __launch_bounds__(DWShape<NB>::THREADS, 1) kl_tc_dw_kernel(const DWArgs a) {
    using SH = DWShape<NB>;
    constexpr int F = 9 + NB;
    constexpr int NROLE = SH::NROLE, NPAR = SH::NPAR, FPP = SH::FPP, NCOLS = SH::NCOLS, NSTG = SH::NSTG;
    constexpr int NSETS = NROLE * NPAR;
    constexpr int TMA_WARP = 4 * NSETS, MMA_WARP = TMA_WARP + 1;
    static_assert(NCOLS % 16 == 0 && 3 * NCOLS <= 512, "feature tile must be a legal MMA N and fit three accumulators");
    extern __shared__ __align__(1024) unsigned char tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    unsigned char* gbase = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const KLLayout& L = a.L;
    // buffer p: [gy hi | gy lo | NSTG x (features hi | lo)]
    const uint32_t offTab = NPAR * SH::BUF_BYTES, offBar = offTab + (uint32_t)(8 * L.TAB * 4 + 127) / 128 * 128;
    const uint32_t barFull = base + offBar, barEmpty = barFull + 8 * NPAR, barDone = barEmpty + 8 * NPAR;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + offBar + 8 * (2 * NPAR + 1));
    float* tabs = reinterpret_cast<float*>(gbase + offTab);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int I = L.I, O = L.O;
    const int i0 = blockIdx.x * 8;
    const int otile = blockIdx.y;
    const long long nchunks_total = (a.B + RC - 1) / RC;
    const long long chunk0 = (long long)blockIdx.z * a.chunks_per_split;
    const int nchunks = (int)min((long long)a.chunks_per_split, nchunks_total - chunk0);
    constexpr uint32_t tmem_cols = 512;                 // 2 rotating main accumulators + 1 correction, NCOLS each

    if (tid == 0) {
        for (int s = 0; s < NPAR; ++s) { mbar_init(barFull + 8 * s, 128 * NROLE); mbar_init(barEmpty + 8 * s, 1); }
        mbar_init(barDone, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32((const void*)tmem_slot)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // per-input tables of this CTA's 8 inputs (padding rows of the table are zero, see kl_pack_kernel)
    for (int q = tid; q < 8 * L.TAB; q += blockDim.x) tabs[q] = a.blob[L.tab + (size_t)i0 * L.TAB + q];
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4 * NSETS) {
        // ===================== producers: gy tile (A operand) and feature tile (B operand) of every NPAR-th chunk =====
        const int set = warp >> 2;
        const int role = set % NROLE, par = set / NROLE;
        const int t = tid & 127, row = t & 31, pg = t >> 5;          // pair pg: inputs i0 + 2 pg, + 1
        const int tg = (role * 128 + t);                             // index inside the parity group (gy tile work split)
        const bool has_prev = BR == FETODE_BRANCH_FERRO2D && !a.prev_is_zero;
        // MN-major tf32 tiles (SWIZZLE_128B_BASE32B): 4-row atoms of 512 B, 32-byte chunks XOR-ed with row & 3
        const uint32_t rowoff = (uint32_t)((row >> 2) * 512 + (row & 3) * 128);
        const uint32_t rx = (uint32_t)(row & 3);
        const int ig = i0 + 2 * pg;
        const uint32_t buf = base + (uint32_t)par * SH::BUF_BYTES;
        const uint32_t fbase = buf + 2 * GY_BYTES + rowoff;
        auto put = [&](int q, const float* v4) {                     // 16-byte chunk q of the chunk's feature tile
            float hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                hi[e] = __uint_as_float(__float_as_uint(v4[e]) & 0xffffe000u);
                lo[e] = v4[e] - hi[e];
            }
            const uint32_t cs = (uint32_t)q & 7u;
            const uint32_t addr = fbase + (uint32_t)(q >> 3) * 2 * STAGE_BYTES + ((((cs >> 1) ^ rx) << 5) | ((cs & 1u) << 4));
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(hi[0]), "f"(hi[1]), "f"(hi[2]), "f"(hi[3]) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr + STAGE_BYTES), "f"(lo[0]), "f"(lo[1]), "f"(lo[2]), "f"(lo[3]) : "memory");
        };
        auto ldx = [&](const float* p, long long b) -> float2 {
            if (b >= a.B || ig >= I) return make_float2(0.f, 0.f);
            return __ldg(reinterpret_cast<const float2*>(p + b * I + ig));
        };
        const float* tgp = tabs + (size_t)(2 * pg) * L.TAB;
        float2 xn = make_float2(0.f, 0.f), pn = xn;
        if (par < nchunks) {
            xn = ldx(a.x, (chunk0 + par) * RC + row);
            if (role == 1 && has_prev) pn = ldx(a.prev, (chunk0 + par) * RC + row);
        }
#pragma unroll 1
        for (int c = par; c < nchunks; c += NPAR) {
            const long long r0 = (chunk0 + c) * RC;
            const int use = c / NPAR;                                // how many times this buffer was filled before
            // ---- features of (row, pair) in registers first: the buffer may still be in use by the MMA
            const float2 xc = xn, pc = pn;
            if (c + NPAR < nchunks) {
                xn = ldx(a.x, r0 + NPAR * RC + row);
                if (role == 1 && has_prev) pn = ldx(a.prev, r0 + NPAR * RC + row);
            }
            float f[20];
            if (role == 0) {
                spline_features(xc.x, tgp, L, f);
                spline_features(xc.y, tgp + L.TAB, L, f + 9);
                f[18] = f[19] = 0.f;
            } else {
                if constexpr (NB > 0) {
                    branch_features<NB, BR>(xc.x, pc.x, a.gl, tgp, L, f);
                    branch_features<NB, BR>(xc.y, pc.y, a.gl, tgp + L.TAB, L, f + NB);
                }
            }
            // ---- this thread's share of the gy tile: [32 rows x 128 outputs] = 1024 float4
            float4 gv[1024 / (128 * NROLE)];
#pragma unroll
            for (int k = 0; k < 1024 / (128 * NROLE); ++k) {
                const int q = tg + k * 128 * NROLE;
                const int b = q >> 5, o4 = q & 31;
                const int o = otile * 128 + o4 * 4;
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r0 + b < a.B && o < O) {
                    const float* gp = a.gy + (r0 + b) * O + o;
                    if (o + 3 < O && (O & 3) == 0) g = __ldg(reinterpret_cast<const float4*>(gp));
                    else { g.x = gp[0]; if (o + 1 < O) g.y = gp[1]; if (o + 2 < O) g.z = gp[2]; if (o + 3 < O) g.w = gp[3]; }
                }
                gv[k] = g;
            }
            mbar_wait(barEmpty + 8 * par, (use & 1) ^ 1);
#pragma unroll
            for (int k = 0; k < 1024 / (128 * NROLE); ++k) {
                const int q = tg + k * 128 * NROLE;
                const int b = q >> 5, o4 = q & 31;
                const float4 g = gv[k];
                float4 h;
                h.x = __uint_as_float(__float_as_uint(g.x) & 0xffffe000u); h.y = __uint_as_float(__float_as_uint(g.y) & 0xffffe000u);
                h.z = __uint_as_float(__float_as_uint(g.z) & 0xffffe000u); h.w = __uint_as_float(__float_as_uint(g.w) & 0xffffe000u);
                const uint32_t addr = buf + (uint32_t)((o4 >> 3) * (RC * 128) + (b >> 2) * 512 + (b & 3) * 128 + (((((o4 & 7) >> 1) ^ (b & 3)) << 5) | ((o4 & 1) << 4)));
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w) : "memory");
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr + GY_BYTES), "f"(g.x - h.x), "f"(g.y - h.y), "f"(g.z - h.z), "f"(g.w - h.w) : "memory");
            }
            if (role == 0) {
#pragma unroll
                for (int k = 0; k < 5; ++k) put(pg * (FPP / 4) + k, f + 4 * k);
            } else {
#pragma unroll
                for (int k = 0; k < NB / 2; ++k) put(pg * (FPP / 4) + 5 + k, f + 4 * k);
            }
            fence_proxy_async();
            mbar_arrive(barFull + 8 * par);
        }
        // ===================== epilogue: TMEM [o lanes x features] -> folded-weight gradient =====================
        mbar_wait(barDone, 0);
        tc_fence_after();
        const int o = otile * 128 + (warp & 3) * 32 + lane;
        for (int c0 = 16 * set; c0 < NCOLS; c0 += 16 * NSETS) {
            float acc[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) acc[e] = 0.f;
            for (int part = 0; part < 3; ++part) {                     // main 0, main 1, correction
                if (part == 1 && nchunks < 2) continue;
                uint32_t v[16];
                const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(part * NCOLS + c0);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int e = 0; e < 16; ++e) acc[e] += __uint_as_float(v[e]);
            }
            if (o < O && nchunks > 0) {
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const int col = c0 + e;
                    const int g = col / FPP, p = col % FPP;
                    int ii, f;
                    if (p < 18) { ii = p / 9; f = p % 9; }
                    else if (p < 20) continue;
                    else { ii = (p - 20) / (NB > 0 ? NB : 1); f = 9 + (p - 20) % (NB > 0 ? NB : 1); }
                    const int i = i0 + 2 * g + ii;
                    if (i >= I) continue;
                    float* dstp = a.ws + ((size_t)i * F + f) * O + o;
                    if (gridDim.z == 1 && !a.accumulate) *dstp = acc[e]; else atomicAdd(dstp, acc[e]);
                }
            }
        }
        tc_fence_before();
    } else if (warp == MMA_WARP) {
        // ===================== MMA issue =====================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(NCOLS >> 3) << 17) |
                                   ((uint32_t)(128 >> 4) << 24);
            for (int c = 0; c < nchunks; ++c) {
                const int p = c % NPAR;
                mbar_wait(barFull + 8 * p, (c / NPAR) & 1);
                tc_fence_after();
                const uint32_t gHi = base + (uint32_t)p * SH::BUF_BYTES, gLo = gHi + GY_BYTES;
                const uint32_t fHi = gHi + 2 * GY_BYTES, fLo = fHi + STAGE_BYTES;
                const uint32_t dMain = tmem_base + (uint32_t)((c & 1) * NCOLS);
                const uint32_t dCorr = tmem_base + (uint32_t)(2 * NCOLS);
#pragma unroll
                for (int ks = 0; ks < RC / 8; ++ks) {
                    // one K = 8 slice = two 4-row atoms, 512 B apart (SBO); 32-element groups along M / N are LBO apart:
                    // RC*128 B for gy, 2*STAGE_BYTES for the features (hi and lo groups interleave)
                    const uint64_t dGh = umma_desc(gHi + ks * 1024, RC * 128, 512, 1), dGl = umma_desc(gLo + ks * 1024, RC * 128, 512, 1);
                    const uint64_t dFh = umma_desc(fHi + ks * 1024, 2 * STAGE_BYTES, 512, 1), dFl = umma_desc(fLo + ks * 1024, 2 * STAGE_BYTES, 512, 1);
                    umma_tf32(dMain, dGh, dFh, idesc, !(c < 2 && ks == 0));
                    umma_tf32(dCorr, dGh, dFl, idesc, !(c == 0 && ks == 0));
                    umma_tf32(dCorr, dGl, dFh, idesc, 1);
                }
                umma_commit(barEmpty + 8 * p);
            }
            umma_commit(barDone);
        }
    }
    __syncthreads();
    if (warp == TMA_WARP) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
    }
}

template <int NB, int BR>
static int dw_launch_k(const DWArgs& a, dim3 grid, cudaStream_t st) {
    using SH = DWShape<NB>;
    const size_t smem = (size_t)SH::NPAR * SH::BUF_BYTES + ((size_t)8 * a.L.TAB * 4 + 127) / 128 * 128 + 1024 + 256;
    auto k = kl_tc_dw_kernel<NB, BR>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e) return (int)e;
    k<<<grid, SH::THREADS, smem, st>>>(a);
    return FETODE_LAUNCH_RC();
}

}  // namespace tcb

// gW = A^T gy on the tensor cores; ws must be zeroed by the caller (row splits accumulate with atomics).
// Returns FETODE_EUNSUPPORTED when the caller should use the CUDA-core kernel.
int kl_tc_dw(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
             int prev_is_zero, const float* gy, float* ws, long long B, int accumulate, cudaStream_t st) {
    if (!L.tc_ok) return FETODE_EUNSUPPORTED;
    if (((uintptr_t)x & 15) || (prev && ((uintptr_t)prev & 15)) || ((uintptr_t)gy & 15)) return FETODE_EUNSUPPORTED;
    tcb::DWArgs a{};
    a.blob = blob; a.x = x; a.prev = prev; a.gy = gy; a.ws = ws; a.B = B; a.L = L; a.accumulate = accumulate;
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero;
    const long long nchunks = (B + tcb::RC - 1) / tcb::RC;
    const int blocks = (L.I + 7) / 8, otiles = (L.O + 127) / 128;
    // fill the 148 SMs with whole waves; <= 2048 rows per CTA keeps the truncating accumulation chains short
    const long long ctas = (long long)blocks * otiles;
    long long splits = 1;
    {
        double best = 1e30;
        for (long long s = 1; s <= 8; ++s) {
            const double t = (double)((ctas * s + 147) / 148) / (double)s;       // waves per unit of work
            if (t < best - 1e-9) { best = t; splits = s; }
        }
    }
    splits = std::max(splits, (nchunks + 63) / 64);
    splits = std::min(splits, nchunks);
    a.chunks_per_split = (int)((nchunks + splits - 1) / splits);
    splits = (nchunks + a.chunks_per_split - 1) / a.chunks_per_split;
    if (splits > 65535 || otiles > 65535) return FETODE_EUNSUPPORTED;
    const dim3 grid((unsigned)blocks, (unsigned)otiles, (unsigned)splits);
    switch (P->branch) {
        case FETODE_BRANCH_NONE: return tcb::dw_launch_k<0, FETODE_BRANCH_NONE>(a, grid, st);
        case FETODE_BRANCH_LOGISTIC:
            if (L.NB == 10) return tcb::dw_launch_k<10, FETODE_BRANCH_LOGISTIC>(a, grid, st);
            if (L.NB == 8) return tcb::dw_launch_k<8, FETODE_BRANCH_LOGISTIC>(a, grid, st);
            break;
        case FETODE_BRANCH_FERRO2D:
            if (L.NB == 10) return tcb::dw_launch_k<10, FETODE_BRANCH_FERRO2D>(a, grid, st);
            break;
    }
    return FETODE_EUNSUPPORTED;
}

}  // namespace fetode
