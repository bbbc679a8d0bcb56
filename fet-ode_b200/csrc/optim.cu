// optim.cu -- fused Adam / AdamW step with gradient-norm clipping on ONE flat fp32 buffer (sm_100a).
//
// what follows the gradient all-reduce in the reference's training loops,
//   torch.optim.Adam(lr=2e-3)                         train_kanfet_node_predprey.py:171, step :257
//   torch.optim.AdamW + clip_grad_norm_(.., 1.0)      train_ecg_kan_fet_nn_ode.py:1093, 1110-1111
// For the Lotka-Volterra net (1 412 parameters in 24 tensors) the per-tensor optimiser launches cost more than the
// fused forward + backward kernels; here the whole update is two launches (sum of squares, update) on the flat
// buffer the all-reduce already uses.  HBM-bound: 16 B read + 12 B written per parameter (+ 4 B when the clipped
// gradient is written back, as clip_grad_norm_ does in place).
//
// Arithmetic follows torch/optim/adam.py (_single_tensor_adam, amsgrad=False, maximize=False) operation by operation:
//   g += wd * p (Adam, L2)            |  p *= 1 - lr * wd (AdamW, decoupled)
//   m = m + (g - m) * (1 - beta1)     (lerp_)
//   v = v * beta2 + (1 - beta2) * g * g
//   p -= (lr / (1 - beta1^t)) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)
// and torch.nn.utils.clip_grad_norm_: coef = min(1, max_norm / (sqrt(sum g^2) + 1e-6)), g *= coef.
#include <cmath>
#include "common.cuh"

namespace fetode {

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, long long n, double* sum) {
    double acc = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float v = g[i];
        acc += (double)v * (double)v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ double part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += part[w];
        atomicAdd(sum, s);
    }
}

struct AdamArgs {
    float* p; float* g; float* m; float* v;
    const double* sumsq;           // device: sum of squared gradients, or null (no clipping)
    long long n;
    float lr, beta1, beta2, eps, wd, max_norm, step_size, bc2_sqrt;
    int decoupled;
    const float* dev_bc;           // device {step_size, bc2_sqrt} (step counter kept on the device), or null
};

// Device-resident step counter for CUDA-graph capture: state = {int64 step; float step_size; float bc2_sqrt}.
__global__ void adam_advance_kernel(long long* state, float lr, float beta1, float beta2) {
    const long long step = ++state[0];
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    float* f = reinterpret_cast<float*>(state + 1);
    f[0] = (float)((double)lr / bc1);
    f[1] = (float)sqrt(bc2);
}

__global__ void __launch_bounds__(256) adam_flat_kernel(const AdamArgs a) {
    float coef = 1.0f;
    if (a.sumsq) {
        const float total = (float)sqrt(*a.sumsq);
        coef = fminf(a.max_norm / (total + 1e-6f), 1.0f);
    }
    const float step_size = a.dev_bc ? a.dev_bc[0] : a.step_size;
    const float bc2_sqrt = a.dev_bc ? a.dev_bc[1] : a.bc2_sqrt;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
        float p = a.p[i], g = a.g[i], m = a.m[i], v = a.v[i];
        if (a.sumsq) { g *= coef; a.g[i] = g; }
        if (a.wd != 0.f) {
            if (a.decoupled) p *= 1.0f - a.lr * a.wd;
            else g = fmaf(a.wd, p, g);
        }
        m = fmaf(g - m, 1.0f - a.beta1, m);
        v = fmaf(v, a.beta2, (1.0f - a.beta2) * g * g);
        const float denom = sqrtf(v) / bc2_sqrt + a.eps;
        p = p - step_size * (m / denom);
        a.p[i] = p; a.m[i] = m; a.v[i] = v;
    }
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_sumsq_f32(const float* g, int64_t n, double* sumsq, void* stream) {
    FETODE_CHECK_ARG(sumsq && n >= 0 && (g || n == 0));
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(sumsq, 0, sizeof(double), st);
    if (e) return (int)e;
    if (n == 0) return FETODE_OK;
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 4);
    sumsq_kernel<<<blocks, 256, 0, st>>>(g, n, sumsq);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_adam_flat(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                                float eps, float weight_decay, int decoupled, int64_t step, float max_norm,
                                const double* sumsq, void* stream) {
    FETODE_CHECK_ARG(p && g && m && v && n >= 0 && step >= 1);
    FETODE_CHECK_ARG(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f);
    if (n == 0) return FETODE_OK;
    AdamArgs a;
    a.p = p; a.g = g; a.m = m; a.v = v; a.sumsq = sumsq; a.n = n;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay; a.max_norm = max_norm;
    a.decoupled = decoupled;
    a.dev_bc = nullptr;
    // python-float (double) bias corrections rounded once, as torch does with scalar step counts
    const double bc1 = 1.0 - std::pow((double)beta1, (double)step);
    const double bc2 = 1.0 - std::pow((double)beta2, (double)step);
    a.step_size = (float)((double)lr / bc1);
    a.bc2_sqrt = (float)std::sqrt(bc2);
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 8);
    adam_flat_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_adam_flat_graph(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1,
                                      float beta2, float eps, float weight_decay, int decoupled, int64_t* state,
                                      float max_norm, const double* sumsq, void* stream) {
    FETODE_CHECK_ARG(p && g && m && v && n >= 0 && state);
    FETODE_CHECK_ARG(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f);
    cudaStream_t st = (cudaStream_t)stream;
    adam_advance_kernel<<<1, 1, 0, st>>>(reinterpret_cast<long long*>(state), lr, beta1, beta2);
    int rc = FETODE_LAUNCH_RC();
    if (rc || n == 0) return rc;
    AdamArgs a;
    a.p = p; a.g = g; a.m = m; a.v = v; a.sumsq = sumsq; a.n = n;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay; a.max_norm = max_norm;
    a.decoupled = decoupled;
    a.step_size = 0.f; a.bc2_sqrt = 1.f;
    a.dev_bc = reinterpret_cast<const float*>(state + 1);
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 8);
    adam_flat_kernel<<<blocks, 256, 0, st>>>(a);
    return FETODE_LAUNCH_RC();
}
