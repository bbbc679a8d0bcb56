// kuramoto.cu -- Kuramoto2D phase lattice (sm_100a): the ODE of the MNIST configuration, forward and
// backward, one CTA per image with the sin/cos planes in shared memory for all Euler steps.
//
// Reference arithmetic being replaced: Kuramoto2D.forward  mnist_kuramoto_kan.py:175-204
//   theta0 = pi (2 x - 1);  `steps` times:  theta += dt (omega + K (cos(theta) conv4(sin theta) - sin(theta) conv4(cos theta)))
//   with conv4 the zero-padded 4-neighbour sum (:167-172, F.conv2d padding=1);  out = [cos theta, sin theta]  [B, 2 H W].
// The reference launches ~6 kernels per step (sin, cos, 2 cuDNN convs, 2 element-wise); here the
// whole loop is one launch and the feature rows are written straight in the layout KANLinear reads.
#include "common.cuh"

namespace fetode {

constexpr int KU_THREADS = 256;
constexpr int KU_MAXPP = 8;            // pixels per thread: H*W <= 2048

struct KuArgs {
    const float *x, *K, *omega, *ckpt_in, *feat_in, *gfeat;
    float *feat, *ckpt, *gx, *gK, *gomega;
    long long B;
    int H, W, steps;
    float dt;
};

__global__ void __launch_bounds__(KU_THREADS) kuramoto_fwd_kernel(const KuArgs a) {
    extern __shared__ float ku_sm[];
    const int H = a.H, W = a.W, HW = H * W, PW = W + 2, PN = (H + 2) * PW;
    float* sp = ku_sm;                 // padded sin plane
    float* cp = ku_sm + PN;            // padded cos plane
    const long long b = blockIdx.x;
    for (int t = threadIdx.x; t < 2 * PN; t += KU_THREADS) ku_sm[t] = 0.f;
    float th[KU_MAXPP], om[KU_MAXPP];
    const float Kc = a.K[0];
    const float pi = 3.14159265358979323846f;
#pragma unroll
    for (int q = 0; q < KU_MAXPP; ++q) {
        const int p = threadIdx.x + q * KU_THREADS;
        th[q] = 0.f; om[q] = 0.f;
        if (p < HW) { th[q] = pi * (2.0f * a.x[b * HW + p] - 1.0f); om[q] = a.omega[p]; }
    }
    __syncthreads();
    for (int s = 0; s < a.steps; ++s) {
#pragma unroll
        for (int q = 0; q < KU_MAXPP; ++q) {
            const int p = threadIdx.x + q * KU_THREADS;
            if (p < HW) {
                float sn, cs;
                sincosf(th[q], &sn, &cs);
                const int r = p / W, c = p % W, pp = (r + 1) * PW + c + 1;
                sp[pp] = sn; cp[pp] = cs;
                if (a.ckpt) a.ckpt[((size_t)s * a.B + b) * HW + p] = th[q];
            }
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < KU_MAXPP; ++q) {
            const int p = threadIdx.x + q * KU_THREADS;
            if (p < HW) {
                const int r = p / W, c = p % W, pp = (r + 1) * PW + c + 1;
                const float S = sp[pp - PW] + sp[pp + PW] + sp[pp - 1] + sp[pp + 1];
                const float C = cp[pp - PW] + cp[pp + PW] + cp[pp - 1] + cp[pp + 1];
                const float coupling = cp[pp] * S - sp[pp] * C;                    // :197
                th[q] = th[q] + a.dt * (om[q] + Kc * coupling);                    // :199
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int q = 0; q < KU_MAXPP; ++q) {
        const int p = threadIdx.x + q * KU_THREADS;
        if (p < HW) {
            float sn, cs;
            sincosf(th[q], &sn, &cs);
            a.feat[b * 2 * HW + p] = cs;                                            // :201-202
            a.feat[b * 2 * HW + HW + p] = sn;
        }
    }
}

// reverse sweep.  d theta_{n+1} / d theta_n [q,p] = delta_qp (1 - dt K (s_p S_p + c_p C_p)) + dt K cos(theta_q - theta_p) [p in N(q)]
__global__ void __launch_bounds__(KU_THREADS) kuramoto_bwd_kernel(const KuArgs a) {
    extern __shared__ float ku_sm[];
    const int H = a.H, W = a.W, HW = H * W, PW = W + 2, PN = (H + 2) * PW;
    float* sp = ku_sm;
    float* cp = ku_sm + PN;
    float* gp = ku_sm + 2 * PN;        // padded upstream gradient plane
    __shared__ float red[KU_THREADS / 32];
    const long long b = blockIdx.x;
    for (int t = threadIdx.x; t < 3 * PN; t += KU_THREADS) ku_sm[t] = 0.f;
    const float Kc = a.K[0];
    float g[KU_MAXPP], gom[KU_MAXPP];
    float gK = 0.f;
#pragma unroll
    for (int q = 0; q < KU_MAXPP; ++q) {
        const int p = threadIdx.x + q * KU_THREADS;
        g[q] = 0.f; gom[q] = 0.f;
        if (p < HW) {
            const float cs = a.feat_in[b * 2 * HW + p], sn = a.feat_in[b * 2 * HW + HW + p];
            g[q] = -sn * a.gfeat[b * 2 * HW + p] + cs * a.gfeat[b * 2 * HW + HW + p];
        }
    }
    __syncthreads();
    for (int s = a.steps - 1; s >= 0; --s) {
#pragma unroll
        for (int q = 0; q < KU_MAXPP; ++q) {
            const int p = threadIdx.x + q * KU_THREADS;
            if (p < HW) {
                float sn, cs;
                sincosf(a.ckpt_in[((size_t)s * a.B + b) * HW + p], &sn, &cs);
                const int r = p / W, c = p % W, pp = (r + 1) * PW + c + 1;
                sp[pp] = sn; cp[pp] = cs; gp[pp] = g[q];
            }
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < KU_MAXPP; ++q) {
            const int p = threadIdx.x + q * KU_THREADS;
            if (p < HW) {
                const int r = p / W, c = p % W, pp = (r + 1) * PW + c + 1;
                const float sn = sp[pp], cs = cp[pp];
                const float S = sp[pp - PW] + sp[pp + PW] + sp[pp - 1] + sp[pp + 1];
                const float C = cp[pp - PW] + cp[pp + PW] + cp[pp - 1] + cp[pp + 1];
                // sum over neighbours q of g_q cos(theta_q - theta_p) = c_p sum g_q c_q + s_p sum g_q s_q
                const float GC = gp[pp - PW] * cp[pp - PW] + gp[pp + PW] * cp[pp + PW] + gp[pp - 1] * cp[pp - 1] +
                                 gp[pp + 1] * cp[pp + 1];
                const float GS = gp[pp - PW] * sp[pp - PW] + gp[pp + PW] * sp[pp + PW] + gp[pp - 1] * sp[pp - 1] +
                                 gp[pp + 1] * sp[pp + 1];
                const float gq = g[q];
                gom[q] += a.dt * gq;
                gK += a.dt * gq * (cs * S - sn * C);
                g[q] = gq * (1.0f - a.dt * Kc * (sn * S + cs * C)) + a.dt * Kc * (cs * GC + sn * GS);
            }
        }
        __syncthreads();
    }
    const float pi = 3.14159265358979323846f;
#pragma unroll
    for (int q = 0; q < KU_MAXPP; ++q) {
        const int p = threadIdx.x + q * KU_THREADS;
        if (p < HW) {
            if (a.gx) a.gx[b * HW + p] = 2.0f * pi * g[q];
            if (a.gomega) atomicAdd(a.gomega + p, gom[q]);
        }
    }
    gK = warp_sum(gK);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = gK;
    __syncthreads();
    if (threadIdx.x == 0 && a.gK) {
        float t = 0.f;
        for (int w = 0; w < KU_THREADS / 32; ++w) t += red[w];
        atomicAdd(a.gK, t);
    }
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_kuramoto_fwd(const float* x, const float* K, const float* omega, int32_t H, int32_t W,
                                   int32_t steps, float dt, float* feat, float* ckpt, int64_t B, void* stream) {
    FETODE_CHECK_ARG(x && K && omega && feat && H >= 1 && W >= 1 && steps >= 0 && B >= 0);
    FETODE_CHECK_SUPPORTED(H * W <= KU_THREADS * KU_MAXPP && B <= 0x7fffffffLL);
    if (B == 0) return FETODE_OK;
    KuArgs a{};
    a.x = x; a.K = K; a.omega = omega; a.feat = feat; a.ckpt = ckpt; a.B = B; a.H = H; a.W = W; a.steps = steps; a.dt = dt;
    const size_t smem = (size_t)2 * (H + 2) * (W + 2) * sizeof(float);
    FETODE_CHECK_SUPPORTED(smem <= 48 * 1024);
    kuramoto_fwd_kernel<<<(unsigned)B, KU_THREADS, smem, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_kuramoto_bwd(const float* ckpt, const float* theta_final_feat, const float* K,
                                   const float* omega, int32_t H, int32_t W, int32_t steps, float dt,
                                   const float* gfeat, float* gx, float* gK, float* gomega, int64_t B, void* stream) {
    FETODE_CHECK_ARG(theta_final_feat && K && gfeat && H >= 1 && W >= 1 && steps >= 0 && B >= 0);
    FETODE_CHECK_ARG(steps == 0 || ckpt != nullptr);
    FETODE_CHECK_SUPPORTED(H * W <= KU_THREADS * KU_MAXPP && B <= 0x7fffffffLL);
    (void)omega;
    if (B == 0) return FETODE_OK;
    KuArgs a{};
    a.ckpt_in = ckpt; a.feat_in = theta_final_feat; a.K = K; a.gfeat = gfeat; a.gx = gx; a.gK = gK; a.gomega = gomega;
    a.B = B; a.H = H; a.W = W; a.steps = steps; a.dt = dt;
    const size_t smem = (size_t)3 * (H + 2) * (W + 2) * sizeof(float);
    FETODE_CHECK_SUPPORTED(smem <= 48 * 1024);
    kuramoto_bwd_kernel<<<(unsigned)B, KU_THREADS, smem, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}
