// glue.cu -- the normalise-and-bound prologue of the ECG right-hand side as one kernel each way (sm_100a).
//
// Replaces  h = self.ln(h); h = self.h_bound * torch.tanh(h / self.h_bound)   (train_ecg_kan_fet_nn_ode.py:459-461,
// nn.LayerNorm(latent_dim) built at :437) and its autograd.  Stock PyTorch spends 25 us (LayerNorm) + 2 us (tanh)
// forward and 20 + 19 + 2 us backward per call on a [4096, 64] tensor -- 32 calls per training step; the arithmetic is
// 1 MB of traffic.  Here: one warp per row, the row lives in registers (D <= 256), statistics by shuffles, the
// gamma / beta gradients accumulate in registers across the rows a warp owns (lane l always owns features l, l+32, ..).
//
//   z = (x - mean) * rstd * gamma + beta,   mean / var over the feature axis, var biased, rstd = 1/sqrt(var + eps)
//   y = hb * tanh(z / hb)
//   backward:  dz = gy * (1 - (y/hb)^2);  dgamma = sum_b dz * xhat;  dbeta = sum_b dz;
//              dx = rstd * (dxh - mean_d(dxh) - xhat * mean_d(dxh * xhat)),  dxh = dz * gamma
// HBM-bound: 8 B / element forward, 16 B / element backward (x, y, gy read; gx written).
#include "common.cuh"

namespace fetode {

constexpr int LN_MAXV = 8;              // features per lane: D <= 256
constexpr int LN_WARPS = 8;
constexpr int LN_ROWS_PER_WARP = 8;

struct LnArgs {
    const float* x; const float* gamma; const float* beta;
    float* y; float* mean; float* rstd;
    const float* gy; float* gx; float* ggamma; float* gbeta;
    long long B;
    int D;
    float eps, hb, inv_hb;
};

__device__ __forceinline__ float warp_allsum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(LN_WARPS * 32) ln_tanh_fwd_kernel(const LnArgs a) {
    const int lane = threadIdx.x & 31;
    const long long warp = (long long)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * LN_WARPS;
    const int D = a.D;
    const float inv_d = 1.0f / (float)D;
    float g[LN_MAXV], bt[LN_MAXV];
#pragma unroll
    for (int v = 0; v < LN_MAXV; ++v) {
        const int d = lane + 32 * v;
        g[v] = d < D ? a.gamma[d] : 0.f;
        bt[v] = d < D ? a.beta[d] : 0.f;
    }
    for (long long b = warp; b < a.B; b += nwarps) {
        float xv[LN_MAXV];
        float s = 0.f;
#pragma unroll
        for (int v = 0; v < LN_MAXV; ++v) {
            const int d = lane + 32 * v;
            xv[v] = d < D ? a.x[b * D + d] : 0.f;
            s += xv[v];
        }
        const float mean = warp_allsum(s) * inv_d;
        float q = 0.f;
#pragma unroll
        for (int v = 0; v < LN_MAXV; ++v) {
            const int d = lane + 32 * v;
            const float c = d < D ? xv[v] - mean : 0.f;
            q = fmaf(c, c, q);
        }
        const float rstd = rsqrtf(warp_allsum(q) * inv_d + a.eps);
#pragma unroll
        for (int v = 0; v < LN_MAXV; ++v) {
            const int d = lane + 32 * v;
            if (d < D) {
                const float z = fmaf((xv[v] - mean) * rstd, g[v], bt[v]);
                a.y[b * D + d] = a.hb * tanhf(z * a.inv_hb);
            }
        }
        if (lane == 0) { a.mean[b] = mean; a.rstd[b] = rstd; }
    }
}

__global__ void __launch_bounds__(LN_WARPS * 32) ln_tanh_bwd_kernel(const LnArgs a) {
    __shared__ float sg[32 * LN_MAXV], sb[32 * LN_MAXV];
    const int lane = threadIdx.x & 31;
    const long long warp = (long long)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * LN_WARPS;
    const int D = a.D;
    const float inv_d = 1.0f / (float)D;
    for (int q = threadIdx.x; q < 32 * LN_MAXV; q += blockDim.x) { sg[q] = 0.f; sb[q] = 0.f; }
    __syncthreads();
    float g[LN_MAXV], ag[LN_MAXV], ab[LN_MAXV];
#pragma unroll
    for (int v = 0; v < LN_MAXV; ++v) {
        const int d = lane + 32 * v;
        g[v] = d < D ? a.gamma[d] : 0.f;
        ag[v] = 0.f; ab[v] = 0.f;
    }
    for (long long b = warp; b < a.B; b += nwarps) {
        const float mean = a.mean[b], rstd = a.rstd[b];
        float xh[LN_MAXV], dxh[LN_MAXV];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int v = 0; v < LN_MAXV; ++v) {
            const int d = lane + 32 * v;
            xh[v] = 0.f; dxh[v] = 0.f;
            if (d < D) {
                const float t = a.y[b * D + d] * a.inv_hb;
                const float dz = a.gy[b * D + d] * fmaf(-t, t, 1.0f);
                xh[v] = (a.x[b * D + d] - mean) * rstd;
                ag[v] = fmaf(dz, xh[v], ag[v]);
                ab[v] += dz;
                dxh[v] = dz * g[v];
                s1 += dxh[v];
                s2 = fmaf(dxh[v], xh[v], s2);
            }
        }
        if (a.gx) {
            const float m1 = warp_allsum(s1) * inv_d, m2 = warp_allsum(s2) * inv_d;
#pragma unroll
            for (int v = 0; v < LN_MAXV; ++v) {
                const int d = lane + 32 * v;
                if (d < D) a.gx[b * D + d] = rstd * (dxh[v] - m1 - xh[v] * m2);
            }
        }
    }
#pragma unroll
    for (int v = 0; v < LN_MAXV; ++v) {
        const int d = lane + 32 * v;
        if (d < D) { atomicAdd(&sg[d], ag[v]); atomicAdd(&sb[d], ab[v]); }
    }
    __syncthreads();
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        if (a.ggamma) atomicAdd(a.ggamma + d, sg[d]);
        if (a.gbeta) atomicAdd(a.gbeta + d, sb[d]);
    }
}

static unsigned ln_blocks(long long B) {
    const long long warps = (B + LN_ROWS_PER_WARP - 1) / LN_ROWS_PER_WARP;
    return (unsigned)std::max(1LL, std::min<long long>((warps + LN_WARPS - 1) / LN_WARPS, 148LL * 8));
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_ln_tanh_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean,
                                  float* rstd, int64_t B, int32_t D, float eps, float h_bound, void* stream) {
    FETODE_CHECK_ARG(x && gamma && beta && y && mean && rstd && B >= 0 && D >= 1 && h_bound > 0.f);
    FETODE_CHECK_SUPPORTED(D <= 32 * LN_MAXV);
    if (B == 0) return FETODE_OK;
    LnArgs a{};
    a.x = x; a.gamma = gamma; a.beta = beta; a.y = y; a.mean = mean; a.rstd = rstd;
    a.B = B; a.D = D; a.eps = eps; a.hb = h_bound; a.inv_hb = 1.0f / h_bound;
    ln_tanh_fwd_kernel<<<ln_blocks(B), LN_WARPS * 32, 0, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_ln_tanh_bwd(const float* x, const float* gamma, const float* mean, const float* rstd,
                                  const float* y, const float* gy, float* gx, float* ggamma, float* gbeta, int64_t B,
                                  int32_t D, float h_bound, void* stream) {
    FETODE_CHECK_ARG(x && gamma && mean && rstd && y && gy && B >= 0 && D >= 1 && h_bound > 0.f);
    FETODE_CHECK_SUPPORTED(D <= 32 * LN_MAXV);
    if (B == 0) return FETODE_OK;
    LnArgs a{};
    a.x = x; a.gamma = gamma; a.mean = const_cast<float*>(mean); a.rstd = const_cast<float*>(rstd);
    a.y = const_cast<float*>(y); a.gy = gy; a.gx = gx; a.ggamma = ggamma; a.gbeta = gbeta;
    a.B = B; a.D = D; a.hb = h_bound; a.inv_hb = 1.0f / h_bound;
    ln_tanh_bwd_kernel<<<ln_blocks(B), LN_WARPS * 32, 0, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}
