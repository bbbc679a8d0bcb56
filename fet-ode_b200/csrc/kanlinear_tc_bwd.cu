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
// Per CTA: 8 inputs (two groups of 4, K-stream padded to FP = 4*ceil(F/4) chunks per group so that a 64-row chunk
// is a whole number of 32-feature stages), all rows of its row range in chunks of 64, one 128-wide output tile.
#include <cstdlib>
#include "tc_common.cuh"

namespace fetode {
namespace tcb {

using namespace tc;

constexpr int RC = 64;                 // rows per chunk
constexpr int NSA = 5;                 // feature-stage ring depth ([64 x 32] hi+lo = 16 KB per stage)
constexpr uint32_t STAGE_BYTES = RC * 32 * 4;            // one of hi / lo
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
    int prev_is_zero, chunks_per_split;
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


template <int NB, int BR>
__global__ void __launch_bounds__(NB > 0 ? 320 : 192, 1) kl_tc_dw_kernel(const DWArgs a) {
    constexpr int F = 9 + NB;
    constexpr int FP = (F + 3) / 4 * 4;                 // 16-byte chunks per group of 4 inputs (padded)
    constexpr int NST = FP / 4;                         // 32-feature stages per 64-row chunk (2 groups)
    constexpr int NCOLS = 8 * FP;                       // features per CTA
    constexpr int NSETS = NB > 0 ? 2 : 1;
    constexpr int NPROD = 128 * NSETS;
    constexpr int TMA_WARP = 4 * NSETS, MMA_WARP = TMA_WARP + 1;
    extern __shared__ __align__(1024) unsigned char tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    unsigned char* gbase = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const KLLayout& L = a.L;
    const uint32_t offA = 0, offG = NSA * 2 * STAGE_BYTES, offTab = offG + 2 * 2 * GY_BYTES,
                   offBar = offTab + (uint32_t)(8 * L.TAB * 4 + 127) / 128 * 128;
    const uint32_t barFullA = base + offBar, barEmpty = barFullA + 8 * NSA, barGyFull = barEmpty + 8 * NSA,
                   barGyEmpty = barGyFull + 16, barDone = barGyEmpty + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + offBar + 8 * (2 * NSA + 5));
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
        for (int s = 0; s < NSA; ++s) { mbar_init(barFullA + 8 * s, NPROD); mbar_init(barEmpty + 8 * s, 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(barGyFull + 8 * s, NPROD); mbar_init(barGyEmpty + 8 * s, 1); }
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
        // ===================== producers: gy tiles (A operand) and feature stages (B operand) =====================
        const int set = warp >> 2;
        const int t = tid & 127, row = t & 63, grp = t >> 6;
        const bool has_prev = BR == FETODE_BRANCH_FERRO2D && !a.prev_is_zero;
        // MN-major tf32 tiles (SWIZZLE_128B_BASE32B): 4-row atoms of 512 B, 32-byte chunks XOR-ed with row & 3
        const uint32_t rowoff = (uint32_t)((row >> 2) * 512 + (row & 3) * 128);
        const uint32_t rx = (uint32_t)(row & 3);
        const int ig = i0 + 4 * grp;                     // first input of this thread's group
        int cur = -1;
        uint32_t dst = 0;
        auto advance = [&](int target) {
            while (cur < target) {
                if (cur >= 0) { fence_proxy_async(); mbar_arrive(barFullA + 8 * (cur % NSA)); }
                ++cur;
                mbar_wait(barEmpty + 8 * (cur % NSA), ((cur / NSA) & 1) ^ 1);
                dst = base + offA + (uint32_t)(cur % NSA) * 2 * STAGE_BYTES + rowoff;
            }
        };
        auto put = [&](int q, const float* v4) {
            const int st = q >> 3;
            if (st != cur) advance(st);
            float hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                hi[e] = __uint_as_float(__float_as_uint(v4[e]) & 0xffffe000u);
                lo[e] = v4[e] - hi[e];
            }
            const uint32_t cs = (uint32_t)q & 7u;
            const uint32_t addr = dst + ((((cs >> 1) ^ rx) << 5) | ((cs & 1u) << 4));
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(hi[0]), "f"(hi[1]), "f"(hi[2]), "f"(hi[3]) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr + STAGE_BYTES), "f"(lo[0]), "f"(lo[1]), "f"(lo[2]), "f"(lo[3]) : "memory");
        };
        auto ldx = [&](const float* p, long long b) -> float4 {
            if (b >= a.B || ig >= I) return make_float4(0.f, 0.f, 0.f, 0.f);
            return __ldg(reinterpret_cast<const float4*>(p + b * I + ig));
        };
        const float* tg = tabs + (size_t)(4 * grp) * L.TAB;
        const float zero4[4] = {0.f, 0.f, 0.f, 0.f};
        float4 xn = ldx(a.x, chunk0 * RC + row);
        float4 pn = has_prev ? ldx(a.prev, chunk0 * RC + row) : make_float4(0.f, 0.f, 0.f, 0.f);
        for (int c = 0; c < nchunks; ++c) {
            const long long r0 = (chunk0 + c) * RC;
            // ---- gy tile of this chunk: [64 rows x 128 outputs] -> hi / lo, MN-major SWIZZLE_128B_BASE32B atoms
            {
                mbar_wait(barGyEmpty + 8 * (c & 1), ((c >> 1) & 1) ^ 1);
                const uint32_t gdst = base + offG + (uint32_t)(c & 1) * 2 * GY_BYTES;
                for (int q = tid; q < RC * 32; q += NPROD) {
                    const int b = q >> 5, o4 = q & 31;
                    const int o = otile * 128 + o4 * 4;
                    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (r0 + b < a.B && o < O) {
                        const float* gp = a.gy + (r0 + b) * O + o;
                        if (o + 3 < O && (O & 3) == 0) g = __ldg(reinterpret_cast<const float4*>(gp));
                        else { g.x = gp[0]; if (o + 1 < O) g.y = gp[1]; if (o + 2 < O) g.z = gp[2]; if (o + 3 < O) g.w = gp[3]; }
                    }
                    float4 h;
                    h.x = __uint_as_float(__float_as_uint(g.x) & 0xffffe000u); h.y = __uint_as_float(__float_as_uint(g.y) & 0xffffe000u);
                    h.z = __uint_as_float(__float_as_uint(g.z) & 0xffffe000u); h.w = __uint_as_float(__float_as_uint(g.w) & 0xffffe000u);
                    const uint32_t addr = gdst + (uint32_t)((o4 >> 3) * 8192 + (b >> 2) * 512 + (b & 3) * 128 + (((((o4 & 7) >> 1) ^ (b & 3)) << 5) | ((o4 & 1) << 4)));
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w) : "memory");
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr + GY_BYTES), "f"(g.x - h.x), "f"(g.y - h.y), "f"(g.z - h.z), "f"(g.w - h.w) : "memory");
                }
                fence_proxy_async();
                mbar_arrive(barGyFull + 8 * (c & 1));
            }
            // ---- features of (row, 4 inputs)
            const float4 xc = xn, pc = pn;
            if (c + 1 < nchunks) {
                xn = ldx(a.x, r0 + RC + row);
                if (has_prev) pn = ldx(a.prev, r0 + RC + row);
            }
            const float xs[4] = {xc.x, xc.y, xc.z, xc.w}, ps[4] = {pc.x, pc.y, pc.z, pc.w};
            const int q0 = c * (2 * FP) + grp * FP;
            if (set == 0) {
                float f[36];
#pragma unroll
                for (int ii = 0; ii < 4; ++ii) spline_features(xs[ii], tg + ii * L.TAB, L, f + ii * 9);
#pragma unroll
                for (int k = 0; k < 9; ++k) put(q0 + k, f + 4 * k);
                if constexpr (NB == 0) {
#pragma unroll
                    for (int k = 9; k < FP; ++k) put(q0 + k, zero4);
                }
            } else {
                if constexpr (NB > 0) {
                    float f[4 * NB];
#pragma unroll
                    for (int ii = 0; ii < 4; ++ii) branch_features<NB, BR>(xs[ii], ps[ii], a.gl, tg + ii * L.TAB, L, f + ii * NB);
#pragma unroll
                    for (int k = 0; k < NB; ++k) put(q0 + 9 + k, f + 4 * k);
#pragma unroll
                    for (int k = 9 + NB; k < FP; ++k) put(q0 + k, zero4);
                }
            }
        }
        if (nchunks > 0) {
            advance(nchunks * NST - 1);
            fence_proxy_async();
            mbar_arrive(barFullA + 8 * (cur % NSA));
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
                    const int g = col / (4 * FP), p = col % (4 * FP);
                    int ii, f;
                    if (p < 36) { ii = p / 9; f = p % 9; }
                    else if (p < 36 + 4 * NB) { ii = (p - 36) / (NB > 0 ? NB : 1); f = 9 + (p - 36) % (NB > 0 ? NB : 1); }
                    else continue;
                    const int i = i0 + 4 * g + ii;
                    if (i >= I) continue;
                    float* dstp = a.ws + ((size_t)i * F + f) * O + o;
                    if (gridDim.z == 1) *dstp = acc[e]; else atomicAdd(dstp, acc[e]);
                }
            }
        }
        tc_fence_before();
    } else if (warp == MMA_WARP) {
        // ===================== MMA issue =====================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(32 >> 3) << 17) |
                                   ((uint32_t)(128 >> 4) << 24);
            for (int c = 0; c < nchunks; ++c) {
                mbar_wait(barGyFull + 8 * (c & 1), (c >> 1) & 1);
                const uint32_t gHi = base + offG + (uint32_t)(c & 1) * 2 * GY_BYTES, gLo = gHi + GY_BYTES;
                for (int s = 0; s < NST; ++s) {
                    const int it = c * NST + s;
                    mbar_wait(barFullA + 8 * (it % NSA), (it / NSA) & 1);
                    tc_fence_after();
                    const uint32_t fHi = base + offA + (uint32_t)(it % NSA) * 2 * STAGE_BYTES, fLo = fHi + STAGE_BYTES;
                    const uint32_t dMain = tmem_base + (uint32_t)((c & 1) * NCOLS + s * 32);
                    const uint32_t dCorr = tmem_base + (uint32_t)(2 * NCOLS + s * 32);
#pragma unroll
                    for (int ks = 0; ks < RC / 8; ++ks) {
                        // one K = 8 slice = two 4-row atoms, 512 B apart (SBO); output atoms of gy are 8 KB apart (LBO)
                        const uint64_t dGh = umma_desc(gHi + ks * 1024, 8192, 512, 1), dGl = umma_desc(gLo + ks * 1024, 8192, 512, 1);
                        const uint64_t dFh = umma_desc(fHi + ks * 1024, 4096, 512, 1), dFl = umma_desc(fLo + ks * 1024, 4096, 512, 1);
                        umma_tf32(dMain, dGh, dFh, idesc, !(c < 2 && ks == 0));
                        umma_tf32(dCorr, dGh, dFl, idesc, !(c == 0 && ks == 0));
                        umma_tf32(dCorr, dGl, dFh, idesc, 1);
                    }
                    umma_commit(barEmpty + 8 * (it % NSA));
                }
                umma_commit(barGyEmpty + 8 * (c & 1));
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
    const size_t smem = (size_t)NSA * 2 * STAGE_BYTES + 2 * 2 * GY_BYTES + ((size_t)8 * a.L.TAB * 4 + 127) / 128 * 128 + 1024 + 256;
    auto k = kl_tc_dw_kernel<NB, BR>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e) return (int)e;
    k<<<grid, NB > 0 ? 320 : 192, smem, st>>>(a);
    return FETODE_LAUNCH_RC();
}

}  // namespace tcb

// gW = A^T gy on the tensor cores; ws must be zeroed by the caller (row splits accumulate with atomics).
// Returns FETODE_EUNSUPPORTED when the caller should use the CUDA-core kernel.
int kl_tc_dw(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
             int prev_is_zero, const float* gy, float* ws, long long B, cudaStream_t st) {
    if (!L.tc_ok) return FETODE_EUNSUPPORTED;
    if (((uintptr_t)x & 15) || (prev && ((uintptr_t)prev & 15)) || ((uintptr_t)gy & 15)) return FETODE_EUNSUPPORTED;
    tcb::DWArgs a{};
    a.blob = blob; a.x = x; a.prev = prev; a.gy = gy; a.ws = ws; a.B = B; a.L = L;
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero;
    const long long nchunks = (B + tcb::RC - 1) / tcb::RC;
    const int blocks = (L.I + 7) / 8, otiles = (L.O + 127) / 128;
    long long splits = std::max(1LL, (148LL + (long long)blocks * otiles - 1) / ((long long)blocks * otiles));
    splits = std::max(splits, (nchunks + 31) / 32);          // <= 2048 rows per CTA: bounded accumulation chains
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
