// rkstage.cu -- Runge-Kutta stage algebra for right-hand sides that are not fused into one kernel
// (arbitrary callables behind odeint, the wide ECG / ETT right-hand sides): one launch per
// stage input / step update instead of the reference's chain of axpy launches.
//
// Reference arithmetic being replaced:
//   integrate_euler / rk2 / rk4   train_ecg_kan_fet_nn_ode.py:670-705  (h + dt/6 (k1+2k2+2k3+k4) ...)
//   odeint_rk4                    train_kan_fet_ett.py:51-83
//   torchdiffeq fixed-grid steps and the dopri5 stage / error / dense-output combinations
//   (call sites train_kanfet_node_predprey.py:252, train_ecg_kan_fet_nn_ode.py:1034-1041)
#include "common.cuh"

namespace fetode {

struct CombineArgs {
    const float* y;
    const float* k[FETODE_RK_MAX_TERMS];
    float c[FETODE_RK_MAX_TERMS];
    float* out;
    long long n;
    int ns;
};

// out = y + sum_i c_i k_i ; HBM-bound: (ns + 2) * 4 bytes per element, 128-bit accesses.
template <int VEC>
__global__ void __launch_bounds__(256) rk_combine_kernel(const CombineArgs a) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long nv = a.n / VEC;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
        if constexpr (VEC == 4) {
            float4 acc = a.y ? reinterpret_cast<const float4*>(a.y)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int s = 0; s < FETODE_RK_MAX_TERMS; ++s) {
                if (s < a.ns) {
                    const float4 kv = reinterpret_cast<const float4*>(a.k[s])[i];
                    const float c = a.c[s];
                    acc.x = fmaf(c, kv.x, acc.x); acc.y = fmaf(c, kv.y, acc.y);
                    acc.z = fmaf(c, kv.z, acc.z); acc.w = fmaf(c, kv.w, acc.w);
                }
            }
            reinterpret_cast<float4*>(a.out)[i] = acc;
        } else {
            float acc = a.y ? a.y[i] : 0.f;
#pragma unroll
            for (int s = 0; s < FETODE_RK_MAX_TERMS; ++s)
                if (s < a.ns) acc = fmaf(a.c[s], a.k[s][i], acc);
            a.out[i] = acc;
        }
    }
}

// partial[blockIdx] = sum over the block's elements of (err / (atol + rtol max(|y0|,|y1|)))^2
__global__ void __launch_bounds__(256) rk_errnorm_kernel(const float* err, const float* y0, const float* y1,
                                                         float atol, float rtol, long long n, double* sum) {
    double acc = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float tol = atol + rtol * fmaxf(fabsf(y0[i]), fabsf(y1[i]));
        const float r = err[i] / tol;
        acc += (double)r * (double)r;
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

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_rk_combine(const float* y, const float* const* ks, const float* coef, int32_t ns,
                                 float* out, int64_t n, void* stream) {
    FETODE_CHECK_ARG(out && n >= 0 && ns >= 0 && ns <= FETODE_RK_MAX_TERMS);
    FETODE_CHECK_ARG(ns == 0 || (ks && coef));
    if (n == 0) return FETODE_OK;
    CombineArgs a;
    a.y = y; a.out = out; a.n = n; a.ns = ns;
    bool vec = (((uintptr_t)out | (uintptr_t)y) & 15) == 0 && (n % 4) == 0;
    for (int s = 0; s < FETODE_RK_MAX_TERMS; ++s) {
        a.k[s] = s < ns ? ks[s] : nullptr;
        a.c[s] = s < ns ? coef[s] : 0.f;
        if (s < ns) { FETODE_CHECK_ARG(ks[s] != nullptr); vec = vec && (((uintptr_t)ks[s]) & 15) == 0; }
    }
    const long long work = vec ? n / 4 : n;
    const int blocks = (int)std::min<long long>((work + 255) / 256, 148LL * 8);
    if (vec) rk_combine_kernel<4><<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
    else rk_combine_kernel<1><<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_rk_error_sumsq(const float* err, const float* y0, const float* y1, float atol, float rtol,
                                     int64_t n, double* sumsq, void* stream) {
    FETODE_CHECK_ARG(err && y0 && y1 && sumsq && n >= 0);
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(sumsq, 0, sizeof(double), st);
    if (e) return (int)e;
    if (n == 0) return FETODE_OK;
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 4);
    rk_errnorm_kernel<<<blocks, 256, 0, st>>>(err, y0, y1, atol, rtol, n, sumsq);
    return FETODE_LAUNCH_RC();
}
