// kanrnn.cu -- the logistic-KAN recurrent encoder of the ETT "KAN-FET" forecaster as ONE launch per direction (sm_100a).
//
// Replaces (reference checkout 704a0df4):
//   FullyNonlinearKANCell.forward   train_kan_fet_ett.py:791-796   (same cell: train_ecg.py:203-217)
//   KANRNNEncoder.forward's loop    train_kan_fet_ett.py:810-815   (336 sequential cell calls, ~10 launches each)
// and their autograd.
//
// Cell: phi = [LogisticBasis_in(x_t) | LogisticBasis_h(h_prev)] flattened (i, j) row-major, out = sigmoid(phi),
// h_t = out[:, :H].  Column c < F*nb reads x_t[c / nb]; column c >= F*nb reads h_prev[(c - F*nb) / nb].  With the
// script's sizes (F = 7, nb = 10, H = 64) every kept column reads x_t only: h_T is a function of the LAST context row
// and the 335 earlier cell calls are dead code -- the kernel then evaluates one step.  The general case (H > F*nb) runs
// the T steps in-kernel with h in shared memory.
//
// Mapping: CTA = RN_ROWS batch rows, thread = column c (its a, b in registers for the whole launch).
#include "common.cuh"

namespace fetode {

constexpr int RN_ROWS = 4;

struct RnnArgs {
    const float *x, *a_in, *b_in, *a_h, *b_h;     // x [B,T,F]; a_in,b_in [F,nb]; a_h,b_h [H,nb]
    const float* h0;                               // [B,H] initial hidden state or null (zeros: KANRNNEncoder :811)
    float* g_h0;                                   // backward: dL/dh_0 [B,H] or null
    float* h_out;                                  // [B,H]
    float* hist;                                   // [T+1,B,H] (h_0 = 0 .. h_T) or null; only written when H > F*nb
    const float* g_h;                              // backward: dL/dh_T [B,H]
    float *g_x, *g_a_in, *g_b_in, *g_a_h, *g_b_h;  // g_x [B,T,F] zero-filled by the caller
    long long B;
    int T, F, H, nb;
};

__device__ __forceinline__ float rn_phi(float v, float a, float b) {
    return 2.0f * rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * a * (v - b)));      // LogisticBasis :748-749
}
__device__ __forceinline__ float rn_sig(float z) { return rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * z)); }

__global__ void __launch_bounds__(256) kanrnn_fwd_kernel(const RnnArgs a) {
    extern __shared__ float rn_sm[];                 // [RN_ROWS][2][H]
    const int c = threadIdx.x, H = a.H, nb = a.nb, Fn = a.F * nb;
    const bool live = c < H;
    const bool from_x = c < Fn;
    const int i = from_x ? c / nb : (c - Fn) / nb, j = from_x ? c % nb : (c - Fn) % nb;
    float pa = 0.f, pb = 0.f;
    if (live) { pa = from_x ? a.a_in[i * nb + j] : a.a_h[i * nb + j]; pb = from_x ? a.b_in[i * nb + j] : a.b_h[i * nb + j]; }
    const bool recurrent = H > Fn;                   // some kept column reads h_prev
    const long long b0 = (long long)blockIdx.x * RN_ROWS;
    for (int r = 0; r < RN_ROWS; ++r) {
        const long long b = b0 + r;
        if (b >= a.B) break;                         // uniform over the block
        float* hb = rn_sm + r * 2 * H;
        if (!recurrent) {
            if (live) a.h_out[b * H + c] = rn_sig(rn_phi(a.x[(b * a.T + (a.T - 1)) * a.F + i], pa, pb));
            continue;
        }
        if (live) { const float hi = a.h0 ? a.h0[b * H + c] : 0.f; hb[c] = hi; if (a.hist) a.hist[b * H + c] = hi; }
        __syncthreads();
        int cur = 0;
        for (int t = 0; t < a.T; ++t) {
            if (live) {
                const float v = from_x ? a.x[(b * a.T + t) * a.F + i] : hb[cur * H + i];
                const float hn = rn_sig(rn_phi(v, pa, pb));
                hb[(cur ^ 1) * H + c] = hn;
                if (a.hist) a.hist[((size_t)(t + 1) * a.B + b) * H + c] = hn;
            }
            __syncthreads();
            cur ^= 1;
        }
        if (live) a.h_out[b * H + c] = hb[cur * H + c];
        __syncthreads();
    }
}

// reverse sweep; per-column parameter gradients accumulate in registers over rows and steps, one atomic pair per CTA
__global__ void __launch_bounds__(256) kanrnn_bwd_kernel(const RnnArgs a) {
    extern __shared__ float rn_sm[];                 // [2][H] adjoint of h (ping-pong)
    const int c = threadIdx.x, H = a.H, nb = a.nb, Fn = a.F * nb;
    const bool live = c < H;
    const bool from_x = c < Fn;
    const int i = from_x ? c / nb : (c - Fn) / nb, j = from_x ? c % nb : (c - Fn) % nb;
    float pa = 0.f, pb = 0.f;
    if (live) { pa = from_x ? a.a_in[i * nb + j] : a.a_h[i * nb + j]; pb = from_x ? a.b_in[i * nb + j] : a.b_h[i * nb + j]; }
    const bool recurrent = H > Fn;
    const int t_stop = recurrent ? 0 : a.T - 1;      // non-recurrent: only the last step is live
    float ga = 0.f, gb = 0.f;
    const long long b0 = (long long)blockIdx.x * RN_ROWS;
    for (int r = 0; r < RN_ROWS; ++r) {
        const long long b = b0 + r;
        if (b >= a.B) break;
        if (live) { rn_sm[c] = a.g_h[b * H + c]; rn_sm[H + c] = 0.f; }
        __syncthreads();
        int cur = 0;
        for (int t = a.T - 1; t >= t_stop; --t) {
            if (live) {
                const float g = rn_sm[cur * H + c];
                const float v = from_x ? a.x[(b * a.T + t) * a.F + i] : a.hist[((size_t)t * a.B + b) * H + i];
                const float phi = rn_phi(v, pa, pb);
                const float s = rn_sig(phi);
                const float e = g * s * (1.0f - s) * phi * (1.0f - 0.5f * phi);   // dL/dz, z = a (v - b)
                ga = fmaf(e, v - pb, ga);
                gb = fmaf(-e, pa, gb);
                const float dv = e * pa;
                if (from_x) { if (a.g_x) atomicAdd(a.g_x + (b * a.T + t) * a.F + i, dv); }
                else atomicAdd(&rn_sm[(cur ^ 1) * H + i], dv);
            }
            __syncthreads();
            if (live) rn_sm[cur * H + c] = 0.f;       // becomes the accumulator of the step after next
            __syncthreads();
            cur ^= 1;
        }
        if (live && a.g_h0) a.g_h0[b * H + c] = recurrent ? rn_sm[cur * H + c] : 0.f;
        __syncthreads();
    }
    if (live) {
        float* da = from_x ? a.g_a_in : a.g_a_h;
        float* db = from_x ? a.g_b_in : a.g_b_h;
        if (da) atomicAdd(da + i * nb + j, ga);
        if (db) atomicAdd(db + i * nb + j, gb);
    }
}

}  // namespace fetode

using namespace fetode;

static int rn_check(const float* x, int64_t B, int T, int F, int H, int nb) {
    if (!x || B < 0 || T < 1 || F < 1 || H < 1 || nb < 1) return FETODE_EINVAL;
    if (H > 256) return FETODE_EUNSUPPORTED;
    if (H > (F + H) * nb) return FETODE_ESHAPE;
    return FETODE_OK;
}

extern "C" int fetode_kanrnn_fwd(const float* x, const float* h0, const float* a_in, const float* b_in,
                                 const float* a_h, const float* b_h, float* h_out, float* hist, int64_t B, int32_t T,
                                 int32_t F, int32_t H, int32_t nb, void* stream) {
    int rc = rn_check(x, B, T, F, H, nb);
    if (rc) return rc;
    FETODE_CHECK_ARG(a_in && b_in && a_h && b_h && h_out);
    if (B == 0) return FETODE_OK;
    RnnArgs a{};
    a.x = x; a.h0 = h0; a.a_in = a_in; a.b_in = b_in; a.a_h = a_h; a.b_h = b_h; a.h_out = h_out; a.hist = hist;
    a.B = B; a.T = T; a.F = F; a.H = H; a.nb = nb;
    const int threads = (H + 31) / 32 * 32;
    const unsigned grid = (unsigned)((B + RN_ROWS - 1) / RN_ROWS);
    kanrnn_fwd_kernel<<<grid, threads, (size_t)RN_ROWS * 2 * H * sizeof(float), (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_kanrnn_bwd(const float* x, const float* a_in, const float* b_in, const float* a_h,
                                 const float* b_h, const float* hist, const float* g_h, float* g_x, float* g_h0,
                                 float* g_a_in, float* g_b_in, float* g_a_h, float* g_b_h, int64_t B, int32_t T,
                                 int32_t F, int32_t H, int32_t nb, void* stream) {
    int rc = rn_check(x, B, T, F, H, nb);
    if (rc) return rc;
    FETODE_CHECK_ARG(a_in && b_in && a_h && b_h && g_h);
    FETODE_CHECK_ARG(H <= F * nb || hist != nullptr);
    if (B == 0) return FETODE_OK;
    RnnArgs a{};
    a.x = x; a.a_in = a_in; a.b_in = b_in; a.a_h = a_h; a.b_h = b_h; a.hist = const_cast<float*>(hist); a.g_h = g_h;
    a.g_x = g_x; a.g_h0 = g_h0; a.g_a_in = g_a_in; a.g_b_in = g_b_in; a.g_a_h = g_a_h; a.g_b_h = g_b_h;
    a.B = B; a.T = T; a.F = F; a.H = H; a.nb = nb;
    const int threads = (H + 31) / 32 * 32;
    const unsigned grid = (unsigned)((B + RN_ROWS - 1) / RN_ROWS);
    kanrnn_bwd_kernel<<<grid, threads, (size_t)2 * H * sizeof(float), (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}
