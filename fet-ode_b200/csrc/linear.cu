// linear.cu -- the small dense head of the ECG right-hand side in fp32 (sm_100a).
//
// Replaces  dh = self.out(z)  (nn.Linear(hidden, latent_dim), train_ecg_kan_fet_nn_ode.py:446, :475) and its
// autograd.  The shapes are small ([4096,128] x [128,64]) and the reference computes them in fp32 (TF32 is off by
// default), where the library's SIMT GEMM launches 64 CTAs and takes 21 us forward and 57 us backward per call --
// 32 calls per training step.  One strided tile kernel serves all three products:
//   forward   y[b,o]  = sum_k x[b,k] W[o,k] + bias[o]
//   dX        gx[b,k] = sum_o gy[b,o] W[o,k]
//   dW        gW[o,k] = sum_b gy[b,o] x[b,k]          (split over b, atomics into a zeroed buffer)
//   dbias     gb[o]   = sum_b gy[b,o]
// FP32 FMA pipe; 32 x 64 output tiles so that a [4096, 64] result still fills 128 CTAs.
#include "common.cuh"

namespace fetode {

constexpr int LG_BM = 32, LG_BN = 64, LG_BK = 32, LG_THREADS = 128;

struct GemmArgs {
    const float* A; const float* B; const float* bias; float* C;
    long long M, sAm, sAk, sBk, sBn;
    int N, K, k_per_split;
    int accumulate;          // C += (the weight-gradient buffer is shared by every call of a step), never overwritten
    // fused head (fetode_head_fwd / _bwd): dh = gain * (silu(z) W^T + b)
    int a_silu, b_silu;      // the A / B operand is silu(stored value)
    const float* scale;      // device scalar multiplying the result (the gain), or null
    const float* emul;       // [M, N]: the result is also multiplied by silu'(emul[m, n]), or null
    float* C2;               // [M, N]: receives the UNSCALED result (before scale / emul), or null
};

__device__ __forceinline__ float lg_silu(float x) { return x / (1.0f + __expf(-x)); }
__device__ __forceinline__ float lg_dsilu(float x) {
    const float s = 1.0f / (1.0f + __expf(-x));
    return s * fmaf(x, 1.0f - s, 1.0f);
}

// C[m,n] (+)= sum_k A[m*sAm + k*sAk] * B[k*sBk + n*sBn] (+ bias[n]);  C row-major [M, N].
// BM x 64 tile (BM = 32 or 16), 128 threads, (BM / 8) x 4 outputs per thread: per k one 128-bit (64-bit) load of the A
// rows (two addresses per warp: broadcast), one of four B columns, 16 (8) FMA.  The next k-tile's global loads are
// issued into registers before the current tile's FMAs (the products are one wave of 4-warp CTAs: nothing else hides
// the load latency); BM = 16 is used while 32-row tiles would leave SMs with a single CTA (the ECG head: 4096 rows).
// HEAD = true compiles the fused-head prologues / epilogues in (a_silu, b_silu, scale, emul, C2); the plain products keep
// their own lean instantiation.
template <int BM, bool HEAD>
__global__ void __launch_bounds__(LG_THREADS) small_gemm_kernel(const GemmArgs a) {
    constexpr int RPT = BM / 8;                             // rows per thread
    constexpr int NA = BM * LG_BK / LG_THREADS, NBL = LG_BK * LG_BN / LG_THREADS;
    __shared__ __align__(16) float As[LG_BK][BM + 4];
    __shared__ __align__(16) float Bs[LG_BK][LG_BN];
    const int tid = threadIdx.x;
    const long long m0 = (long long)blockIdx.x * BM;
    const int n0 = blockIdx.y * LG_BN;
    const int k_begin = blockIdx.z * a.k_per_split, k_end = min(a.K, k_begin + a.k_per_split);
    const int ty = tid / 16, tx = tid % 16;                 // thread -> rows RPT ty .. ; columns 4 tx .. 4 tx + 3
    float acc[RPT][4];
#pragma unroll
    for (int i = 0; i < RPT; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const bool a_k_fast = a.sAk == 1, b_n_fast = a.sBn == 1;
    float ra[NA], rb[NBL];
    auto fetch = [&](int k0) {
#pragma unroll
        for (int u = 0; u < NA; ++u) {
            const int q = tid + u * LG_THREADS;
            const int kk = a_k_fast ? q % LG_BK : q / BM, mm = a_k_fast ? q / LG_BK : q % BM;
            const long long m = m0 + mm;
            const int k = k0 + kk;
            ra[u] = (m < a.M && k < k_end) ? a.A[m * a.sAm + k * a.sAk] : 0.f;
            if (HEAD && a.a_silu) ra[u] = lg_silu(ra[u]);
        }
#pragma unroll
        for (int u = 0; u < NBL; ++u) {
            const int q = tid + u * LG_THREADS;
            const int nn = b_n_fast ? q % LG_BN : q / LG_BK, kk = b_n_fast ? q / LG_BN : q % LG_BK;
            const int n = n0 + nn, k = k0 + kk;
            rb[u] = (n < a.N && k < k_end) ? a.B[k * a.sBk + n * a.sBn] : 0.f;
            if (HEAD && a.b_silu) rb[u] = lg_silu(rb[u]);
        }
    };
    if (k_begin < k_end) fetch(k_begin);
    for (int k0 = k_begin; k0 < k_end; k0 += LG_BK) {
#pragma unroll
        for (int u = 0; u < NA; ++u) {
            const int q = tid + u * LG_THREADS;
            const int kk = a_k_fast ? q % LG_BK : q / BM, mm = a_k_fast ? q / LG_BK : q % BM;
            As[kk][mm] = ra[u];
        }
#pragma unroll
        for (int u = 0; u < NBL; ++u) {
            const int q = tid + u * LG_THREADS;
            const int nn = b_n_fast ? q % LG_BN : q / LG_BK, kk = b_n_fast ? q / LG_BN : q % LG_BK;
            Bs[kk][nn ^ ((kk & 15) << 2)] = rb[u];        // 4-column groups swizzled by k: a k-fast B (the forward's W[o, k])
                                                          // would otherwise store 32 lanes into one bank
        }
        __syncthreads();
        if (k0 + LG_BK < k_end) fetch(k0 + LG_BK);          // in flight during the FMAs below
#pragma unroll
        for (int kk = 0; kk < LG_BK; ++kk) {
            const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][4 * (tx ^ (kk & 15))]);
            float ar[RPT];
            if constexpr (RPT == 4) {
                const float4 av = *reinterpret_cast<const float4*>(&As[kk][4 * ty]);
                ar[0] = av.x; ar[1] = av.y; ar[2] = av.z; ar[3] = av.w;
            } else {
                const float2 av = *reinterpret_cast<const float2*>(&As[kk][2 * ty]);
                ar[0] = av.x; ar[1] = av.y;
            }
#pragma unroll
            for (int i = 0; i < RPT; ++i) {
                acc[i][0] = fmaf(ar[i], b.x, acc[i][0]); acc[i][1] = fmaf(ar[i], b.y, acc[i][1]);
                acc[i][2] = fmaf(ar[i], b.z, acc[i][2]); acc[i][3] = fmaf(ar[i], b.w, acc[i][3]);
            }
        }
        __syncthreads();
    }
    const float gain = (HEAD && a.scale) ? *a.scale : 1.0f;
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
        const long long m = m0 + RPT * ty + i;
        if (m >= a.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + 4 * tx + j;
            if (n >= a.N) continue;
            float v = acc[i][j];
            if (a.bias && blockIdx.z == 0) v += a.bias[n];
            if constexpr (HEAD) {
                if (a.C2) a.C2[m * a.N + n] = v;
                if (a.scale) v *= gain;
                if (a.emul) v *= lg_dsilu(a.emul[m * a.N + n]);
            }
            if (gridDim.z == 1 && !a.accumulate) a.C[m * a.N + n] = v;
            else atomicAdd(a.C + m * a.N + n, v);
        }
    }
}

// gb[o] += scale * sum_b gy[b,o];  optionally *gdot += sum_{b,o} gy[b,o] u[b,o]  (the gain's gradient in the fused head)
__global__ void __launch_bounds__(256) colsum_acc_kernel(const float* gy, float* gb, long long B, int O, const float* scale,
                                                         const float* u, float* gdot) {
    const int o = blockIdx.x * 32 + (threadIdx.x & 31);
    const int w = threadIdx.x >> 5;
    float acc = 0.f, dot = 0.f;
    if (o < O)
        for (long long b = (long long)blockIdx.y * 8 + w; b < B; b += (long long)gridDim.y * 8) {
            const float g = gy[b * O + o];
            acc += g;
            if (u) dot = fmaf(g, u[b * O + o], dot);
        }
    __shared__ float part[8][33];
    part[w][threadIdx.x & 31] = acc;
    __syncthreads();
    if (w == 0 && o < O && gb) {
        float s = 0.f;
        for (int k = 0; k < 8; ++k) s += part[k][threadIdx.x];
        atomicAdd(gb + o, scale ? s * *scale : s);
    }
    if (gdot) {                                           // uniform over the grid
        __syncthreads();
        part[w][threadIdx.x & 31] = dot;
        __syncthreads();
        if (w == 0) {
            float s = 0.f;
            for (int k = 0; k < 8; ++k) s += part[k][threadIdx.x];
#pragma unroll
            for (int q = 16; q > 0; q >>= 1) s += __shfl_xor_sync(0xffffffffu, s, q);
            if (threadIdx.x == 0 && s != 0.f) atomicAdd(gdot, s);
        }
    }
}

static int small_gemm(const GemmArgs& a0, int splits, cudaStream_t st) {
    GemmArgs a = a0;
    splits = std::max(1, std::min(splits, (a.K + LG_BK - 1) / LG_BK));
    a.k_per_split = ((a.K + splits - 1) / splits + LG_BK - 1) / LG_BK * LG_BK;
    splits = (a.K + a.k_per_split - 1) / a.k_per_split;
    const long long nt = (a.N + LG_BN - 1) / LG_BN;
    const long long mt32 = (a.M + LG_BM - 1) / LG_BM;
    const bool half = mt32 * nt * splits < 2 * 148 && a.M > 16;          // 32-row tiles would not give every SM two CTAs
    const long long mt = half ? (a.M + 15) / 16 : mt32;
    if (mt > 0x7fffffffLL) return FETODE_EUNSUPPORTED;
    const dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)splits);
    const bool head = a.a_silu || a.b_silu || a.scale || a.emul || a.C2;
    if (head) {
        if (half) small_gemm_kernel<16, true><<<grid, LG_THREADS, 0, st>>>(a);
        else small_gemm_kernel<LG_BM, true><<<grid, LG_THREADS, 0, st>>>(a);
    } else {
        if (half) small_gemm_kernel<16, false><<<grid, LG_THREADS, 0, st>>>(a);
        else small_gemm_kernel<LG_BM, false><<<grid, LG_THREADS, 0, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_linear_fwd(const float* x, const float* weight, const float* bias, float* y, int64_t B,
                                 int32_t in_features, int32_t out_features, void* stream) {
    FETODE_CHECK_ARG(x && weight && y && B >= 0 && in_features >= 1 && out_features >= 1);
    if (B == 0) return FETODE_OK;
    GemmArgs a{};
    a.A = x; a.sAm = in_features; a.sAk = 1;
    a.B = weight; a.sBk = 1; a.sBn = in_features;           // B(k, n) = W[n, k]
    a.bias = bias; a.C = y; a.M = B; a.N = out_features; a.K = in_features;
    return small_gemm(a, 1, (cudaStream_t)stream);
}

extern "C" int fetode_linear_bwd(const float* x, const float* weight, const float* gy, float* gx, float* gweight,
                                 float* gbias, int64_t B, int32_t in_features, int32_t out_features, void* stream) {
    FETODE_CHECK_ARG(x && weight && gy && B >= 0 && in_features >= 1 && out_features >= 1);
    if (B == 0) return FETODE_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (gx) {                                               // gx[b, k] = sum_o gy[b, o] W[o, k]
        GemmArgs a{};
        a.A = gy; a.sAm = out_features; a.sAk = 1;
        a.B = weight; a.sBk = in_features; a.sBn = 1;
        a.C = gx; a.M = B; a.N = in_features; a.K = out_features;
        const int rc = small_gemm(a, 1, st);
        if (rc) return rc;
    }
    if (gweight) {                                          // gW[o, k] += sum_b gy[b, o] x[b, k]   (caller zeroes)
        if (B > 0x7fffffffLL) return FETODE_EUNSUPPORTED;
        GemmArgs a{};
        a.A = gy; a.sAm = 1; a.sAk = out_features;          // A(m = o, k = b) = gy[b, o]
        a.B = x; a.sBk = in_features; a.sBn = 1;            // B(k = b, n) = x[b, n]
        a.C = gweight; a.M = out_features; a.N = in_features; a.K = (int)B;
        a.accumulate = 1;                                   // also when a tiny batch leaves a single K split
        const long long tiles = ((out_features + LG_BM - 1) / LG_BM) * (long long)((in_features + LG_BN - 1) / LG_BN);
        int splits = (int)std::max(1LL, std::min<long long>(148LL * 2 / tiles, (B + 127) / 128));
        if (splits < 2) splits = 2;                         // always the accumulate path: the buffer is added to
        const int rc = small_gemm(a, splits, st);
        if (rc) return rc;
    }
    if (gbias) {
        const long long by = std::max(1LL, std::min<long long>((B + 63) / 64, 64));
        colsum_acc_kernel<<<dim3((out_features + 31) / 32, (unsigned)by), 256, 0, st>>>(gy, gbias, B, out_features, nullptr, nullptr,
                                                                                  nullptr);
    }
    return FETODE_LAUNCH_RC();
}

// ---- fused tail of the ECG right-hand side: dh = gain * (silu(z) W^T + b) ------------------------------------------------
// Replaces  z = self.act(z); dh = self.out(z); dh = self.scale * alpha * dh  (train_ecg_kan_fet_nn_ode.py:471-480, act =
// nn.SiLU, alpha = softplus(log_alpha)) and its autograd: SiLU is the A-operand prologue of the product, the gain its
// epilogue; in the reverse sweep silu' is the epilogue of the input-gradient product, SiLU the B-operand prologue of the
// weight-gradient product, and the gain's gradient sum(gy * u) rides on the bias-gradient column sum.
extern "C" int fetode_head_fwd(const float* z, const float* weight, const float* bias, const float* gain, float* u, float* y,
                               int64_t B, int32_t in_features, int32_t out_features, void* stream) {
    FETODE_CHECK_ARG(z && weight && gain && y && B >= 0 && in_features >= 1 && out_features >= 1);
    if (B == 0) return FETODE_OK;
    GemmArgs a{};
    a.A = z; a.sAm = in_features; a.sAk = 1; a.a_silu = 1;
    a.B = weight; a.sBk = 1; a.sBn = in_features;
    a.bias = bias; a.C = y; a.C2 = u; a.scale = gain; a.M = B; a.N = out_features; a.K = in_features;
    return small_gemm(a, 1, (cudaStream_t)stream);
}

extern "C" int fetode_head_bwd(const float* z, const float* weight, const float* gain, const float* u, const float* gy,
                               float* gz, float* gweight, float* gbias, float* ggain, int64_t B, int32_t in_features,
                               int32_t out_features, void* stream) {
    FETODE_CHECK_ARG(z && weight && gain && gy && B >= 0 && in_features >= 1 && out_features >= 1 && (!ggain || u));
    if (B == 0) return FETODE_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (gz) {                                               // gz[b, k] = gain * (sum_o gy[b, o] W[o, k]) * silu'(z[b, k])
        GemmArgs a{};
        a.A = gy; a.sAm = out_features; a.sAk = 1;
        a.B = weight; a.sBk = in_features; a.sBn = 1;
        a.C = gz; a.scale = gain; a.emul = z; a.M = B; a.N = in_features; a.K = out_features;
        const int rc = small_gemm(a, 1, st);
        if (rc) return rc;
    }
    if (gweight) {                                          // gW[o, k] += gain * sum_b gy[b, o] silu(z[b, k])
        if (B > 0x7fffffffLL) return FETODE_EUNSUPPORTED;
        GemmArgs a{};
        a.A = gy; a.sAm = 1; a.sAk = out_features;
        a.B = z; a.sBk = in_features; a.sBn = 1; a.b_silu = 1;
        a.C = gweight; a.scale = gain; a.M = out_features; a.N = in_features; a.K = (int)B;
        a.accumulate = 1;
        const long long tiles = ((out_features + LG_BM - 1) / LG_BM) * (long long)((in_features + LG_BN - 1) / LG_BN);
        int splits = (int)std::max(1LL, std::min<long long>(148LL * 2 / tiles, (B + 127) / 128));
        if (splits < 2) splits = 2;
        const int rc = small_gemm(a, splits, st);
        if (rc) return rc;
    }
    if (gbias || ggain) {
        const long long by = std::max(1LL, std::min<long long>((B + 63) / 64, 64));
        colsum_acc_kernel<<<dim3((out_features + 31) / 32, (unsigned)by), 256, 0, st>>>(gy, gbias, B, out_features, gain,
                                                                                  ggain ? u : nullptr, ggain);
    }
    return FETODE_LAUNCH_RC();
}
