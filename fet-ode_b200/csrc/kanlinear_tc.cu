// kanlinear_tc.cu -- wide KANLinear forward on the 5th-generation tensor cores (sm_100a): tcgen05.mma with the
// accumulator in TMEM, the weight operand streamed by the TMA bulk-copy engine and the feature operand
// (SiLU | cubic B-spline basis | logistic / ferroelectric branch basis) evaluated in the kernel's PROLOGUE and
// written straight into the 128B-swizzled UMMA shared-memory layout -- the [B, I*(1+G+k+nb)] feature matrix the
// reference materialises level by level in HBM (efficient_kan/efficientkan.py:117-131, 166-179) never exists.
//
// Reference arithmetic being replaced: KANLinear.forward efficientkan.py:160-182 for the wide layers of the
// ECG / ETT / diffusion configurations (kan.KAN([768,128]), ([128,128]) train_ecg_kan_fet_nn_ode.py:441-443;
// KAN([ctx*F,128]) train_kan_fet_ett.py:676-681; KANFET([in,256]).. kan_diffusion_ett.py:421-425).
//
// Precision: the 1e-5 per-evaluation bar rules out plain TF32 (10-bit mantissa, ~3e-4).  Both operands are split
// x = hi + lo (hi = top 19 bits, lo = x - hi exactly) and three MMAs are issued per K-slice,
// hi*hi + hi*lo + lo*hi, accumulating in fp32 in TMEM; the dropped lo*lo term is ~2^-22 relative.
//
// CTA = 128 rows x BN outputs x one contiguous range of 32-input units (split-K over the grid's z axis, at most 6
// units per CTA so that the truncating tensor-core accumulation chains stay short -- see the MMA loop).
//   producers   (thread r of every producer set owns row r)  The feature evaluation is a long dependent chain
//               (interval search, Cox-de Boor, MUFU ex2/rcp), so it is latency-bound per thread: 4 sets (NB > 0) or
//               2 sets (NB == 0) of 4 warps share the stream.  Inputs are taken in pairs; pair g2 goes to the sets
//               of parity g2 & 1:  "spline" set: silu + dense cubic B-spline row of the 2 inputs,
//               "branch" set (NB > 0): their NB branch-basis factors.  Values are split hi/lo and stored as 16-byte
//               chunks, conflict-free, into the 128B-swizzled A ring (the K axis is permuted per pair of inputs so
//               that every set owns whole chunks).
//               All sets double as the epilogue (TMEM lane quarter == warp & 3): tcgen05.ld of the 3 rotating
//               accumulators + the correction accumulator, summed in registers, (+bias) -> y / atomics
//   next warp   lane 0: cp.async.bulk (TMA bulk engine) of the pre-split, pre-swizzled weight stage (hi+lo) into the
//               B ring; whole warp: TMEM alloc / dealloc
//   next warp   lane 0: tcgen05.mma issue (12 per 32-wide K stage), tcgen05.commit releases ring slots
//   next warp   lane 0: cp.async.bulk of the next unit's per-input tables (knots, reciprocal differences, folded
//               branch constants: 32 x TAB floats) into a 2-slot shared-memory ring
#include <cstdlib>
#include "tc_common.cuh"

namespace fetode {
namespace tc {

struct TCArgs {
    const float* blob;      // standard layer image (bias, per-input tables)
    const float* wtc;       // tensor-core weight image
    const float* x;
    const float* prev;
    float* y;
    long long B;
    KLLayout L;
    float gl;
    int prev_is_zero, BN, units_total, units_per_split, stages_total;
};


// The K axis is ordered per pair of inputs as [ (input ii, silu | N_0..N_7) x 2, 0, 0 | (input ii, r_0..r_{NB-1}) x 2 ]
// (the weight image is permuted to match, kl_tc_pack_kernel): 5 spline-side chunks and NB/2 branch-side chunks of
// 4 values each, produced by two different sets of warps; CPG = 5 + NB/2 chunks per pair, 2*CPG stages per
// 32-input unit.

// silu and the dense cubic B-spline row of one (row, input) pair: f[0..8]
__device__ __forceinline__ void tc_spline_features(float x, const float* __restrict__ tab, const KLLayout& L, float* f) {
    f[0] = x * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));                     // efficientkan.py:166
    int m = -1;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    const float4 gi = *reinterpret_cast<const float4*>(tab + L.t_gi);
    if (x >= gi.x && x < gi.z) {                                                     // :122 half-open test
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

// the NB branch-basis factors of one (row, input) pair (LogisticBasis efficientkan.py:20-24 or the 2-D ferro neuron)
template <int NB, int BR>
__device__ __forceinline__ void tc_branch_features(float x, float prev, float gl, const float* __restrict__ tab,
                                                   const KLLayout& L, float* f) {
    float omu = 0.f;
    if constexpr (BR == FETODE_BRANCH_FERRO2D) omu = rcp_approx(1.0f + ex2_approx(gl * (x - prev)));
    float r[NB > 0 ? NB : 1];
    branch_eval<NB, BR>(x, omu, gl, reinterpret_cast<const float4*>(tab + L.t_bp), r);
#pragma unroll
    for (int j = 0; j < NB; ++j) f[j] = r[j];
}

// producer groups taking every NPAR-th input pair.  Measured on B200 (768->128, B=4096): 2 groups 0.129 ms, 3 groups
// 0.130 ms -- with a 3-stage ring (2.4 pairs of K) a third group has no free slot to fill, so 2 it is.
template <int NB, int BR> struct TCShape {
    static constexpr int NROLE = NB > 0 ? 2 : 1;
    static constexpr int NPAR = 2;
    static constexpr int THREADS = (4 * NROLE * NPAR + 3) * 32;
};

template <int NB, int BR>
__global__ void __launch_bounds__(TCShape<NB, BR>::THREADS, 1) kl_tc_fwd_kernel(const TCArgs a) {
    constexpr int CPG = 5 + NB / 2;                               // 16-byte chunks per pair of inputs
    constexpr int SPU = 2 * CPG;                                  // 32-wide K stages per 32-input unit
    constexpr int NROLE = NB > 0 ? 2 : 1;                         // spline | branch
    constexpr int NPAR = TCShape<NB, BR>::NPAR;                   // producer groups taking every NPAR-th input pair
    constexpr int NSETS = NPAR * NROLE;
    constexpr int NPROD = 128 * NSETS;
    extern __shared__ __align__(1024) unsigned char tc_smem_raw[];
    const int BN = a.BN;
    // carve: A ring (hi, lo per stage), B ring (hi, lo per stage), barriers
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    unsigned char* gbase = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    constexpr uint32_t A_TILE = BM * BK * 4;                     // 16 KB
    const uint32_t B_TILE = (uint32_t)BN * BK * 4;
    const uint32_t TAB_SLAB = (uint32_t)(32 * a.L.TAB * 4);        // per-input tables of one 32-input unit
    const uint32_t offA = 0, offB = NSTAGE * 2 * A_TILE, offTab = offB + NSTAGE * 2 * B_TILE, offBar = offTab + 2 * TAB_SLAB;
    const uint32_t barFullA = base + offBar, barFullB = barFullA + 8 * NSTAGE, barEmpty = barFullB + 8 * NSTAGE,
                   barDone = barEmpty + 8 * NSTAGE, barTabFull = barDone + 8, barTabEmpty = barTabFull + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + offBar + 8 * (3 * NSTAGE + 1 + 4));
    const float* tabs = reinterpret_cast<const float*>(gbase + offTab);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long m0 = (long long)blockIdx.x * BM;
    const int ntile = blockIdx.y, split = blockIdx.z;
    const int unit0 = split * a.units_per_split;
    const int nunits = min(a.units_per_split, a.units_total - unit0);
    const int nit = nunits * SPU;                                 // 32-wide K stages of this CTA
    const int stage0 = unit0 * SPU;
    const uint32_t tmem_cols = 4 * BN < 32 ? 32u : (uint32_t)(4 * BN);   // 3 rotating main accumulators + 1 correction
    constexpr int TMA_WARP = 4 * NSETS, MMA_WARP = TMA_WARP + 1, TAB_WARP = MMA_WARP + 1;

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(barFullA + 8 * s, NPROD);
            mbar_init(barFullB + 8 * s, 1);
            mbar_init(barEmpty + 8 * s, 1);
        }
        mbar_init(barDone, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(barTabFull + 8 * s, 1);
            mbar_init(barTabEmpty + 8 * s, NPROD);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32((const void*)tmem_slot)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4 * NSETS) {
        // ===================== producers: A operand =====================
        const int set = warp >> 2;
        const int role = set % NROLE, par = set / NROLE;
        const int r = tid & 127;
        const long long b = m0 + r;
        const bool live = b < a.B;
        const KLLayout& L = a.L;
        const int I = L.I;
        const float* xrow = a.x + (live ? b : 0) * I;
        const float* prow = a.prev + (live ? b : 0) * I;
        const bool has_prev = BR == FETODE_BRANCH_FERRO2D && !a.prev_is_zero;
        const uint32_t rowoff = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128);
        const uint32_t rx = (uint32_t)(r & 7);
        int cur = -1;                                            // stage this thread has acquired and not yet published
        uint32_t dstA = 0;
        // every thread visits every stage in order: acquire (empty barrier), write its own chunks, publish
        auto advance = [&](int target) {
            while (cur < target) {
                if (cur >= 0) { fence_proxy_async(); mbar_arrive(barFullA + 8 * (cur % NSTAGE)); }
                ++cur;
                mbar_wait(barEmpty + 8 * (cur % NSTAGE), ((cur / NSTAGE) & 1) ^ 1);
                dstA = base + offA + (uint32_t)(cur % NSTAGE) * 2 * A_TILE + rowoff;
            }
        };
        auto put = [&](int q, const float* v4) {                 // 16-byte chunk q of this CTA's K stream
            const int t = q >> 3;
            if (t != cur) advance(t);
            float hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                hi[e] = __uint_as_float(__float_as_uint(v4[e]) & 0xffffe000u);
                lo[e] = v4[e] - hi[e];
            }
            const uint32_t addr = dstA + ((((uint32_t)q & 7u) ^ rx) << 4);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(hi[0]), "f"(hi[1]), "f"(hi[2]), "f"(hi[3]) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr + A_TILE), "f"(lo[0]), "f"(lo[1]), "f"(lo[2]), "f"(lo[3]) : "memory");
        };
        const int npairs = nunits * 16;
        // 2 consecutive inputs of this row (8-byte aligned: I % 4 == 0), zeros for padding inputs / rows
        auto ldx = [&](const float* rowp, int i) -> float2 {
            if (!live || i >= I) return make_float2(0.f, 0.f);
            return __ldg(reinterpret_cast<const float2*>(rowp + i));
        };
        int tab_u = -1;                                           // unit whose table slab this thread currently holds
        float2 xn = ldx(xrow, unit0 * 32 + 2 * par);
        float2 pn = (role == 1 && has_prev) ? ldx(prow, unit0 * 32 + 2 * par) : make_float2(0.f, 0.f);
#pragma unroll 1
        for (int g2 = par; g2 < npairs; g2 += NPAR) {
            const int u = g2 >> 4, g = g2 & 15;
            if (u != tab_u) {
                if (tab_u >= 0) mbar_arrive(barTabEmpty + 8 * (tab_u & 1));
                mbar_wait(barTabFull + 8 * (u & 1), (u >> 1) & 1);
                tab_u = u;
            }
            const float2 xc = xn, pc = pn;
            if (g2 + NPAR < npairs) {
                xn = ldx(xrow, unit0 * 32 + 2 * (g2 + NPAR));
                if (role == 1 && has_prev) pn = ldx(prow, unit0 * 32 + 2 * (g2 + NPAR));
            }
            const float* tg = tabs + (size_t)((u & 1) * 32 + g * 2) * L.TAB;
            if (role == 0) {
                float f[20];
                tc_spline_features(xc.x, tg, L, f);
                tc_spline_features(xc.y, tg + L.TAB, L, f + 9);
                f[18] = f[19] = 0.f;
#pragma unroll
                for (int c = 0; c < 5; ++c) put(g2 * CPG + c, f + 4 * c);
            } else {
                if constexpr (NB > 0) {
                    float f[2 * NB];
                    tc_branch_features<NB, BR>(xc.x, pc.x, a.gl, tg, L, f);
                    tc_branch_features<NB, BR>(xc.y, pc.y, a.gl, tg + L.TAB, L, f + NB);
#pragma unroll
                    for (int c = 0; c < NB / 2; ++c) put(g2 * CPG + 5 + c, f + 4 * c);
                }
            }
        }
        if (tab_u >= 0) mbar_arrive(barTabEmpty + 8 * (tab_u & 1));
        advance(nit - 1);
        fence_proxy_async();
        mbar_arrive(barFullA + 8 * (cur % NSTAGE));
        // ===================== epilogue =====================
        mbar_wait(barDone, 0);
        tc_fence_after();
        const float* bias = a.blob + L.bias;
        const int O = L.O, n0 = ntile * BN;
        const int nacc = nit < 3 ? nit : 3;
        for (int c0 = 16 * set; c0 < BN; c0 += 16 * NSETS) {
            float acc[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) acc[e] = 0.f;
            for (int q = 0; q <= nacc; ++q) {                     // q == nacc: the correction accumulator
                uint32_t v[16];
                const uint32_t col = (uint32_t)((q == nacc ? 3 : q) * BN + c0);
                const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + col;
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int e = 0; e < 16; ++e) acc[e] += __uint_as_float(v[e]);
            }
            if (live && gridDim.z > 1 && (O & 3) == 0) {
                // split-K partial sums: 16-byte vector reductions (one L2 operation per four outputs)
#pragma unroll
                for (int e = 0; e < 16; e += 4) {
                    const int n = n0 + c0 + e;
                    if (n < O) {
                        float v0 = acc[e], v1 = acc[e + 1], v2 = acc[e + 2], v3 = acc[e + 3];
                        if (split == 0) { v0 += bias[n]; v1 += bias[n + 1]; v2 += bias[n + 2]; v3 += bias[n + 3]; }
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                                     ::"l"(a.y + b * O + n), "f"(v0), "f"(v1), "f"(v2), "f"(v3) : "memory");
                    }
                }
            } else if (live) {
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const int n = n0 + c0 + e;
                    if (n < O) {
                        const float val = acc[e];
                        if (gridDim.z == 1) a.y[b * O + n] = val + bias[n];
                        else atomicAdd(a.y + b * O + n, split == 0 ? val + bias[n] : val);
                    }
                }
            }
        }
        tc_fence_before();
    } else if (warp == TMA_WARP) {
        // ===================== TMA: B operand =====================
        if (lane == 0) {
            const uint32_t bytes = 2 * B_TILE;
            const float* src = a.wtc + ((size_t)ntile * a.stages_total + stage0) * (size_t)(2 * BN * BK);
            for (int it = 0; it < nit; ++it) {
                const int s = it % NSTAGE;
                mbar_wait(barEmpty + 8 * s, ((it / NSTAGE) & 1) ^ 1);
                mbar_arrive_expect_tx(barFullB + 8 * s, bytes);
                bulk_g2s(base + offB + (uint32_t)s * 2 * B_TILE, src + (size_t)it * (2 * BN * BK), bytes, barFullB + 8 * s);
            }
        }
    } else if (warp == TAB_WARP) {
        // ===================== TMA: per-input tables of the next unit =====================
        if (lane == 0) {
            const float* src = a.blob + a.L.tab + (size_t)unit0 * 32 * a.L.TAB;
            for (int u = 0; u < nunits; ++u) {
                mbar_wait(barTabEmpty + 8 * (u & 1), ((u >> 1) & 1) ^ 1);
                mbar_arrive_expect_tx(barTabFull + 8 * (u & 1), TAB_SLAB);
                bulk_g2s(base + offTab + (uint32_t)(u & 1) * TAB_SLAB, src + (size_t)u * 32 * a.L.TAB, TAB_SLAB, barTabFull + 8 * (u & 1));
            }
        }
    } else {
        // ===================== MMA issue =====================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            for (int it = 0; it < nit; ++it) {
                const int s = it % NSTAGE;
                const uint32_t ph = (it / NSTAGE) & 1;
                mbar_wait(barFullA + 8 * s, ph);
                mbar_wait(barFullB + 8 * s, ph);
                tc_fence_after();
                const uint32_t aHi = base + offA + (uint32_t)s * 2 * A_TILE, aLo = aHi + A_TILE;
                const uint32_t bHi = base + offB + (uint32_t)s * 2 * B_TILE, bLo = bHi + B_TILE;
#pragma unroll
                for (int j = 0; j < BK / 8; ++j) {
                    const uint64_t dAh = umma_desc(aHi + j * 32), dAl = umma_desc(aLo + j * 32);
                    const uint64_t dBh = umma_desc(bHi + j * 32), dBl = umma_desc(bLo + j * 32);
                    // The tensor core truncates when it adds into the fp32 accumulator, so the error of a chain grows
                    // linearly with its length (measured: 4.5e-6 of the max-norm per 228 MMAs).  The hi*hi products
                    // rotate over three accumulators and the 2^-11-times-smaller corrections go to a fourth; the
                    // epilogue adds the four in registers (round-to-nearest).
                    umma_tf32(tmem_base + (uint32_t)((it % 3) * BN), dAh, dBh, idesc, !(it < 3 && j == 0));
                                    umma_tf32(tmem_base + (uint32_t)(3 * BN), dAh, dBl, idesc, (it | j) != 0);
                    umma_tf32(tmem_base + (uint32_t)(3 * BN), dAl, dBh, idesc, 1);
                }
                umma_commit(barEmpty + 8 * s);
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

// weight image: [ntile][stage][hi|lo][BN x 32] tiles, pre-split and pre-swizzled so that one bulk copy per stage
// lands a ready-to-use UMMA operand.  K index k = 32*stage + col is the flattened (input, feature) row of wcat.
__global__ void __launch_bounds__(256) kl_tc_pack_kernel(KLLayout L, float* blob) {
    const int BN = L.tc_BN;
    const long long per_tile = (long long)BN * BK;
    const long long n = (long long)L.tc_ntile * L.tc_stages * per_tile;
    const long long KF = (long long)L.I * L.F;
    const float* wcat = blob + L.wcat;
    float* dst = blob + L.tc;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(idx % BK);
        const int row = (int)((idx / BK) % BN);
        const long long ts = idx / per_tile;                      // ntile * stages + stage
        const int stage = (int)(ts % L.tc_stages), nt = (int)(ts / L.tc_stages);
        // position in the CTA K stream -> (input, feature) row of wcat (see the ordering note above tc_spline_features)
        const int CPG = 5 + L.NB / 2;
        const long long kpos = (long long)stage * BK + c;
        const long long pair = kpos / (4 * CPG);
        const int p = (int)(kpos % (4 * CPG));
        int ii, f;
        if (p < 18) { ii = p / 9; f = p % 9; }
        else if (p < 20) { ii = 0; f = -1; }                      // padding slots of the spline side
        else { ii = (p - 20) / (L.NB > 0 ? L.NB : 1); f = 9 + (p - 20) % (L.NB > 0 ? L.NB : 1); }
        const long long k = (pair * 2 + ii) * L.F + f;
        const int o = nt * BN + row;
        float w = 0.f;
        if (f >= 0 && pair * 2 + ii < L.I && k < KF && o < L.O) w = wcat[k * L.O + o];
        const float hi = __uint_as_float(__float_as_uint(w) & 0xffffe000u);
        float* tile = dst + ts * 2 * per_tile;
        tile[swz(row, c)] = hi;
        tile[per_tile + swz(row, c)] = w - hi;
    }
}

template <int NB, int BR>
static int tc_launch_k(const TCArgs& a, dim3 grid, size_t smem, cudaStream_t st) {
    auto k = kl_tc_fwd_kernel<NB, BR>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e) return (int)e;
    k<<<grid, TCShape<NB, BR>::THREADS, smem, st>>>(a);
    return FETODE_LAUNCH_RC();
}

}  // namespace tc

int kl_tc_pack(const KLLayout& L, float* blob, cudaStream_t st) {
    if (!L.tc_ok) return FETODE_OK;
    const long long n = (long long)L.tc_ntile * L.tc_stages * L.tc_BN * tc::BK;
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 8);
    tc::kl_tc_pack_kernel<<<blocks, 256, 0, st>>>(L, blob);
    return FETODE_LAUNCH_RC();
}

// returns FETODE_EUNSUPPORTED when the caller should use the CUDA-core path instead
int kl_tc_fwd(const fetode_kanlayer_t* P, const KLLayout& L, const float* blob, const float* x, const float* prev,
              int prev_is_zero, float* y, long long B, cudaStream_t st) {
    if (!L.tc_ok) return FETODE_EUNSUPPORTED;
    tc::TCArgs a{};
    a.blob = blob; a.wtc = blob + L.tc; a.x = x; a.prev = prev; a.y = y; a.B = B; a.L = L;
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero; a.BN = L.tc_BN; a.units_total = L.tc_units; a.stages_total = L.tc_stages;
    const long long mt = (B + tc::BM - 1) / tc::BM;
    if (mt > 0x7fffffffLL) return FETODE_EUNSUPPORTED;
    long long splits = std::max(1LL, 148LL / (mt * L.tc_ntile));
    splits = std::min<long long>(splits, L.tc_units);
    a.units_per_split = (int)((L.tc_units + splits - 1) / splits);
    a.units_per_split = std::min(a.units_per_split, 6);        // bounds the accumulation chains (see the MMA loop)
    {
        static int ups = -2;                                // tuning / diagnostics: 32-input units per K split; read once
        if (ups == -2) { const char* e = getenv("FETODE_TC_UPS"); ups = e ? atoi(e) : -1; }
        if (ups >= 1) a.units_per_split = std::min(ups, L.tc_units);
    }
    splits = (L.tc_units + a.units_per_split - 1) / a.units_per_split;
    if (splits > 1) {
        cudaError_t e = cudaMemsetAsync(y, 0, (size_t)B * L.O * sizeof(float), st);
        if (e) return (int)e;
    }
    const size_t smem = (size_t)tc::NSTAGE * 2 * (tc::BM * tc::BK * 4 + (size_t)L.tc_BN * tc::BK * 4) + (size_t)2 * 32 * L.TAB * 4 + 1024 + 256;
    if (((uintptr_t)x & 15) || (prev && ((uintptr_t)prev & 15)) || ((uintptr_t)y & 15)) return FETODE_EUNSUPPORTED;
    const dim3 grid((unsigned)mt, (unsigned)L.tc_ntile, (unsigned)splits);
    switch (P->branch) {
        case FETODE_BRANCH_NONE: return tc::tc_launch_k<0, FETODE_BRANCH_NONE>(a, grid, smem, st);
        case FETODE_BRANCH_LOGISTIC:
            if (L.NB == 10) return tc::tc_launch_k<10, FETODE_BRANCH_LOGISTIC>(a, grid, smem, st);
            if (L.NB == 8) return tc::tc_launch_k<8, FETODE_BRANCH_LOGISTIC>(a, grid, smem, st);
            break;
        case FETODE_BRANCH_FERRO2D:
            if (L.NB == 10) return tc::tc_launch_k<10, FETODE_BRANCH_FERRO2D>(a, grid, smem, st);
            break;
    }
    return FETODE_EUNSUPPORTED;
}

}  // namespace fetode
