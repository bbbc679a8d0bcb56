// ferro.cu -- stand-alone ferroelectric hysteresis neurons (sm_100a), forward and backward.
//
// Reference arithmetic being replaced (file:line relative to the reference checkout):
//   FULL     FerroelectricBasis.forward        ferro_class.py:368-420  (branch_sign never written => br == 1)
//            NoisyFerroelectricBasis.forward   ferro_class.py:466-519  (noise_std = 0; branch_sign <- target_sign)
//   TWO_D    TwoDimensionFerroelectricBasis    ferro_class.py:547-592
//   HARDGATE hysteretic LogisticBasis.forward  train_ecg_kan_fet_nn_ode.py:102-134 (+ KANFeatureMixer sigmoid :419)
//
// The reference materialises a [B,I,O,K] tensor per elementary op (~30 launches); here one thread owns
// one (row, output) pair and reduces over (input, basis) in registers, the previous-input state is the
// compact [B,I] row block, and the optional branch_sign state of the Noisy variant is the only
// [B,I,O,K] array that touches HBM (one read + one write per call).
//
// The hard gate  (sigmoid(5 (x - prev)) > 0.5)  is evaluated with a correctly rounded exp and IEEE add / divide
// so that the branch taken is bit-identical to torch's CPU result (the threshold is a rounding artefact of
// 1/(1+exp(-z)): true from z >= ~8.94e-8, false at z == 0); everything else uses MUFU ex2 / rcp.
#include "common.cuh"

namespace fetode {

__device__ __forceinline__ float sigm(float z) { return rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * z)); }
__device__ __forceinline__ float tanh_fast(float z) {
    // 1 - 2/(1+e^{2z}); saturates cleanly for |z| large
    return 1.0f - 2.0f * rcp_approx(1.0f + ex2_approx(fminf(2.0f * FETODE_LOG2E * z, 80.0f)));
}

struct SoftBranch {          // forward values the backward needs
    float up, cp, cn, su, sl, tgt, m, T;
};

// ferro_class.py:384-408.  br_one: branch_sign == 1 for every element (FerroelectricBasis, TwoDimension...)
__device__ __forceinline__ SoftBranch soft_branch(float x, float up, float k, float Ec, float br, float g,
                                                  float alpha) {
    SoftBranch s;
    s.up = up;
    s.cp = sigm(g * (x - Ec));
    s.cn = sigm(g * (-x - Ec));
    s.su = up * s.cp;
    s.sl = (1.0f - up) * s.cn;
    s.tgt = s.su - s.sl + (1.0f - s.su - s.sl) * br;
    s.m = alpha * br + (1.0f - alpha) * s.tgt;
    s.T = tanh_fast(k * fmaf(Ec, s.m, x));
    return s;
}

// d(momentum)/dx and d(momentum)/dEc   (prev and br are constants of the graph, ferro_class.py:381-382)
__device__ __forceinline__ void soft_branch_grad(const SoftBranch& s, float br, float g, float alpha,
                                                 float& dm_dx, float& dm_dEc) {
    const float dup = g * s.up * (1.0f - s.up);
    const float dcp = g * s.cp * (1.0f - s.cp);          // d cp / dx  = - d cp / dEc
    const float dcn = g * s.cn * (1.0f - s.cn);          // d cn / dx  = -dcn ; d cn / dEc = -dcn
    const float dsu_dx = dup * s.cp + s.up * dcp;
    const float dsu_dEc = -s.up * dcp;
    const float dsl_dx = -dup * s.cn - (1.0f - s.up) * dcn;
    const float dsl_dEc = -(1.0f - s.up) * dcn;
    const float a = (1.0f - alpha) * (1.0f - br), b = (1.0f - alpha) * (-1.0f - br);
    dm_dx = a * dsu_dx + b * dsl_dx;
    dm_dEc = a * dsu_dEc + b * dsl_dEc;
}

// ------------------------------------------------------------------------------------------
// FULL: params [I,O,K], out [B,O]
// ------------------------------------------------------------------------------------------
constexpr int FF_THREADS = 128;     // 4 warps: lanes = 32 consecutive outputs, warps split the inputs

struct FerroArgs {
    const float *k, *Ec, *Ps, *bias, *coef;
    const float *x, *prev, *br, *gout;
    float *out, *basis_out, *br_out, *gx;
    float *gk, *gEc, *gPs, *gbias, *gcoef;
    unsigned char* gate_out;
    long long B;
    int I, O, K, prev_is_zero, apply_sigmoid;
    float g, alpha;
    // fused right-hand-side glue of KANFetODEFunc.forward (train_ecg.py:997-1013), FULL variant only
    float pre_bound;          // > 0: the neuron sees x_eff = pre_bound * tanh(x / pre_bound)            :1000
    int post_op;              // 1: out = tanh(out) :1003 ; 2: nan_to_num(nan=0, +-inf=+-1e3) then clamp(+-50) :1008-1011
    float* xeff_out;          // [B,I] <- x_eff (the next call's prev_x) or null
    const float* yout;        // backward: the forward's (post-op) output, for d(post)/d(out)
    // implicit im2col (FerroelectricBasisConv2d, ferro_class.py:689-779): rows are (image b, patch l), input i is
    // (channel, ky, kx); x / prev / gx are IMAGES [B,Cin,H,W], out / gout are [B,Cout,L] (+ out_bias)
    fetode_conv2d_t cv;
    int conv, br_momentum;    // br_momentum: br_out <- branch_momentum (:763-767) instead of target_sign
    const float* out_bias;
    float* g_out_bias;
};

// image offset of input i of patch row `row`, or -1 in the zero padding (F.unfold semantics, :699)
__device__ __forceinline__ long long conv_index(const fetode_conv2d_t& g, long long row, int i) {
    const int L = g.Hout * g.Wout;
    const long long b = row / L;
    const int l = (int)(row % L), oy = l / g.Wout, ox = l % g.Wout;
    const int kk = g.kH * g.kW, c = i / kk, r = i % kk, ky = r / g.kW, kx = r % g.kW;
    const int y = oy * g.stride - g.padding + ky, x = ox * g.stride - g.padding + kx;
    if (y < 0 || y >= g.H || x < 0 || x >= g.W) return -1;
    return ((b * g.Cin + c) * g.H + y) * (long long)g.W + x;
}

__device__ __forceinline__ float ferro_post(float v, int op) {
    if (op == 1) return tanh_fast(v);
    if (op == 2) {
        if (v != v) v = 0.f;
        v = fminf(fmaxf(v, -1e3f), 1e3f);                 // +-inf -> +-1e3 (nan_to_num), then
        return fminf(fmaxf(v, -50.0f), 50.0f);            // torch.clamp(dh, -50, 50)
    }
    return v;
}
// d(post)/d(out) from the post-op value y: tanh' = 1 - y^2; the clamp passes the gradient strictly inside (-50, 50)
// (torch also passes it AT the bound: a measure-zero difference) -- a NaN that nan_to_num zeroed poisons the chain
// in torch as well (0 * NaN upstream)
__device__ __forceinline__ float ferro_post_grad(float y, int op) {
    if (op == 1) return 1.0f - y * y;
    if (op == 2) return fabsf(y) < 50.0f ? 1.0f : 0.f;
    return 1.0f;
}

template <bool HAS_BR>
__global__ void __launch_bounds__(FF_THREADS) ferro_full_fwd_kernel(const FerroArgs a) {
    extern __shared__ float ff_sm[];
    float* xs = ff_sm;                 // [I] x row
    float* ups = ff_sm + a.I;          // [I] is_moving_up
    __shared__ float part[4][32];
    const long long b = blockIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int o = blockIdx.y * 32 + lane;
    const int I = a.I, O = a.O, K = a.K;
    for (int i = threadIdx.x; i < I; i += FF_THREADS) {
        const long long src = a.conv ? conv_index(a.cv, b, i) : b * I + i;
        float x = src >= 0 ? a.x[src] : 0.f;
        if (a.pre_bound > 0.f) x = a.pre_bound * tanh_fast(x / a.pre_bound);
        const float p = a.prev_is_zero == 1 ? 0.f : (a.prev_is_zero == 2 ? x : (src >= 0 ? a.prev[src] : 0.f));
        xs[i] = x;
        ups[i] = sigm(a.g * (x - p));
        if (a.xeff_out && blockIdx.y == 0) a.xeff_out[b * I + i] = x;
    }
    __syncthreads();
    float acc = 0.f;
    if (o < O) {
        for (int i = w; i < I; i += 4) {
            const float x = xs[i], up = ups[i];
            const size_t base = ((size_t)i * O + o) * K;
            const size_t sbase = (((size_t)b * I + i) * O + o) * K;
            for (int j = 0; j < K; ++j) {
                const float br = HAS_BR ? a.br[sbase + j] : 1.0f;
                const SoftBranch s = soft_branch(x, up, a.k[base + j], a.Ec[base + j], br, a.g, a.alpha);
                const float P = fmaf(a.Ps[base + j], s.T, a.bias[base + j]);
                acc = fmaf(a.coef[base + j], P, acc);
                if (a.basis_out) a.basis_out[sbase + j] = P;
                if (HAS_BR && a.br_out) a.br_out[sbase + j] = a.br_momentum ? s.m : s.tgt;   // ferro_class.py:512-514 / :767
            }
        }
    }
    part[w][lane] = acc;
    __syncthreads();
    if (w == 0 && o < O) {
        const float v = part[0][lane] + part[1][lane] + part[2][lane] + part[3][lane];
        if (a.conv) {
            const int L = a.cv.Hout * a.cv.Wout;
            a.out[((b / L) * O + o) * L + b % L] = v + (a.out_bias ? a.out_bias[o] : 0.f);     // (B, Cout, L) :770-773
        } else {
            a.out[b * O + o] = ferro_post(v, a.post_op);
        }
    }
}

// backward: one CTA per (input i, 32 outputs, tile of FB_BT rows); warps split the basis index j.
constexpr int FB_BT = 64;

template <bool HAS_BR>
__global__ void __launch_bounds__(FF_THREADS) ferro_full_bwd_kernel(const FerroArgs a) {
    __shared__ float xs[FB_BT], ups[FB_BT], dxs[FB_BT], dpre[FB_BT];
    __shared__ long long srcs[FB_BT];
    const int i = blockIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int o = blockIdx.y * 32 + lane;
    const long long b0 = (long long)blockIdx.z * FB_BT;
    const int nb = (int)min((long long)FB_BT, a.B - b0);
    const int I = a.I, O = a.O, K = a.K;
    for (int t = threadIdx.x; t < FB_BT; t += FF_THREADS) {
        float x = 0.f, p = 0.f;
        float dp = 1.0f;
        long long src = -1;
        if (t < nb) {
            src = a.conv ? conv_index(a.cv, b0 + t, i) : (b0 + t) * I + i;
            x = src >= 0 ? a.x[src] : 0.f;
            if (a.pre_bound > 0.f) {
                const float th = tanh_fast(x / a.pre_bound);
                x = a.pre_bound * th;
                dp = 1.0f - th * th;
            }
            p = a.prev_is_zero == 1 ? 0.f : (a.prev_is_zero == 2 ? x : (src >= 0 ? a.prev[src] : 0.f));
        }
        srcs[t] = src;
        xs[t] = x;
        ups[t] = sigm(a.g * (x - p));
        dxs[t] = 0.f;
        dpre[t] = dp;
    }
    __syncthreads();
    for (int j = w; j < K; j += 4) {
        float k = 0.f, Ec = 0.f, Ps = 0.f, bias = 0.f, coef = 0.f;
        const size_t base = ((size_t)i * O + o) * K + j;
        if (o < O) { k = a.k[base]; Ec = a.Ec[base]; Ps = a.Ps[base]; bias = a.bias[base]; coef = a.coef[base]; }
        float ak = 0.f, aEc = 0.f, aPs = 0.f, ab = 0.f, ac = 0.f, gob = 0.f;
        for (int t = 0; t < nb; ++t) {
            float dx = 0.f;
            if (o < O) {
                const float x = xs[t];
                const float br = HAS_BR ? a.br[(((size_t)(b0 + t) * I + i) * O + o) * K + j] : 1.0f;
                const SoftBranch s = soft_branch(x, ups[t], k, Ec, br, a.g, a.alpha);
                float dm_dx, dm_dEc;
                soft_branch_grad(s, br, a.g, a.alpha, dm_dx, dm_dEc);
                float go;
                if (a.conv) {
                    const int L = a.cv.Hout * a.cv.Wout;
                    go = a.gout[(((b0 + t) / L) * O + o) * L + (b0 + t) % L];
                    if (i == 0 && j == 0) gob += go;
                } else {
                    go = a.gout[(b0 + t) * O + o];
                }
                if (a.post_op) go *= ferro_post_grad(a.yout[(b0 + t) * O + o], a.post_op);
                ac = fmaf(go, fmaf(Ps, s.T, bias), ac);
                const float gP = go * coef;
                aPs = fmaf(gP, s.T, aPs);
                ab += gP;
                const float dz = gP * Ps * (1.0f - s.T * s.T);
                ak = fmaf(dz, fmaf(Ec, s.m, x), ak);
                aEc = fmaf(dz * k, fmaf(Ec, dm_dEc, s.m), aEc);
                dx = dz * k * fmaf(Ec, dm_dx, 1.0f);
            }
            if (a.gx) {
                dx = warp_sum(dx);
                if (lane == 0) atomicAdd(&dxs[t], dx);
            }
        }
        if (o < O) {
            if (a.gk) atomicAdd(a.gk + base, ak);
            if (a.gEc) atomicAdd(a.gEc + base, aEc);
            if (a.gPs) atomicAdd(a.gPs + base, aPs);
            if (a.gbias) atomicAdd(a.gbias + base, ab);
            if (a.gcoef) atomicAdd(a.gcoef + base, ac);
            if (a.conv && a.g_out_bias && i == 0 && j == 0) atomicAdd(a.g_out_bias + o, gob);
        }
    }
    __syncthreads();
    if (a.gx)
        for (int t = threadIdx.x; t < nb; t += FF_THREADS)
            if (srcs[t] >= 0) atomicAdd(a.gx + srcs[t], dxs[t] * dpre[t]);
}

// ------------------------------------------------------------------------------------------
// TWO_D and HARDGATE: params [I,K], out [B,I,K] (element-wise over the basis axis)
// ------------------------------------------------------------------------------------------
// torch's CPU  sigmoid(z) > 0.5  evaluated literally: e = fl32(exp(-z)) (torch's vectorised and scalar float exp are
// both correctly rounded for the arguments that matter here; CUDA's expf is not -- it is 2 ulp -- so the exponential
// is taken in double and rounded once), s = fl32(1 + e), r = fl32(1 / s), r > 0.5.  Net effect: true iff
// e <= 1 - 2^-23, i.e. z > ~8.94e-8; false at z == 0.  One evaluation per (row, input) -- dx does not depend on the
// basis index -- so the double-precision exp is off the critical path.
// Away from the threshold the outcome is decided without the exponential: z >= 1e-6 gives exp(-z) <= 1 - 1e-6 + 5e-13,
// far below 1 - 2^-23 (true); z <= 0 gives e >= 1, s >= 2, r <= 0.5 (false); NaN compares false both ways, as
// NaN > 0.5 does.  Only 0 < z < 1e-6 (x within 2e-7 of prev) takes the double-precision path.
__device__ __forceinline__ bool hard_gate(float z) {
    if (z >= 1e-6f) return true;
    if (!(z > 0.f)) return false;
    const float e = (float)exp(-(double)z);
    const float s = __fadd_rn(1.0f, e);
    return __fdiv_rn(1.0f, s) > 0.5f;
}

// thread = one parameter (i, j), held in registers; the CTA walks a tile of rows: the x load is a broadcast within each
// run of K lanes, the stores are contiguous, and no per-element index division or parameter reload remains
constexpr int FEF_ROWS = 16;

template <int VARIANT>
__global__ void __launch_bounds__(256) ferro_elem_fwd_kernel(const FerroArgs a) {
    const int IK = a.I * a.K;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= IK) return;
    const int i = p / a.K;
    const float k = a.k[p], Ec = a.Ec[p];
    const float Ps = VARIANT == FETODE_FERRO_LOGISTIC ? 0.f : a.Ps[p], bias = VARIANT == FETODE_FERRO_LOGISTIC ? 0.f : a.bias[p];
    const float coef = VARIANT == FETODE_FERRO_TWO_D ? a.coef[p] : 1.0f;
    const float pv_shared = (VARIANT == FETODE_FERRO_HARDGATE && !a.prev_is_zero) ? a.prev[i] : 0.f;   // ONE row shared by the batch :93
    for (long long b0 = (long long)blockIdx.y * FEF_ROWS; b0 < a.B; b0 += (long long)gridDim.y * FEF_ROWS) {
        const int nb = (int)min((long long)FEF_ROWS, a.B - b0);
#pragma unroll 4
        for (int t = 0; t < nb; ++t) {
            const long long b = b0 + t;
            const long long e = b * IK + p;
            const float x = a.x[b * a.I + i];
            float out;
            if constexpr (VARIANT == FETODE_FERRO_LOGISTIC) {
                out = 2.0f * sigm(k * (x - Ec));                                    // efficientkan.py:24
                if (a.basis_out) a.basis_out[e] = out;
            } else if constexpr (VARIANT == FETODE_FERRO_TWO_D) {
                const float pv = a.prev_is_zero ? 0.f : a.prev[b * a.I + i];
                const SoftBranch s = soft_branch(x, sigm(a.g * (x - pv)), k, Ec, 1.0f, a.g, a.alpha);
                const float P = fmaf(Ps, s.T, bias);
                if (a.basis_out) a.basis_out[e] = P;
                out = coef * P;
            } else {
                const bool up = hard_gate(a.g * (x - pv_shared));                   // :122-123
                const float s = sigm(k * (up ? x - Ec : x + Ec));                   // :116-117
                out = fmaf(2.0f * Ps, s, -Ps) + bias;                               // :126
                if (a.gate_out) a.gate_out[e] = up ? 1 : 0;
                if (a.basis_out) a.basis_out[e] = out;
                if (a.apply_sigmoid) out = sigm(out);                               // KANFeatureMixer :419
                if (a.apply_sigmoid == 2 && out != out) out = 0.f;                  // + nan_to_num(nan=0) :465 (sigmoid is never +-inf)
            }
            a.out[e] = out;
        }
    }
}

// one thread per parameter (i,j), loop over a tile of rows; the K basis contributions to gx[b,i] sit in adjacent lanes
// and are summed with a segmented shuffle scan, so each warp issues one atomic per (row, input) it touches
constexpr int FE_BT = 32;

template <int VARIANT>
__global__ void __launch_bounds__(128) ferro_elem_bwd_kernel(const FerroArgs a) {
    const int IK = a.I * a.K;
    const int p_raw = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = p_raw < IK;                           // dead lanes stay for the shuffles, contribute nothing
    const int p = live ? p_raw : IK - 1;
    const int i = live ? p / a.K : -1;
    const int il = live ? i : 0;
    const int lane = threadIdx.x & 31;
    // segment structure of this warp (lanes with equal i are contiguous): fixed for the whole row loop
    const int i_prev = __shfl_up_sync(0xffffffffu, i, 1);
    const bool head = lane == 0 || i_prev != i;
    bool same[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        const int other = __shfl_down_sync(0xffffffffu, i, 1 << q);
        same[q] = lane + (1 << q) < 32 && other == i;
    }
    const float k = a.k[p], Ec = a.Ec[p];
    const float Ps = VARIANT == FETODE_FERRO_LOGISTIC ? 0.f : a.Ps[p], bias = VARIANT == FETODE_FERRO_LOGISTIC ? 0.f : a.bias[p];
    const float coef = VARIANT == FETODE_FERRO_TWO_D ? a.coef[p] : 1.0f;
    float ak = 0.f, aEc = 0.f, aPs = 0.f, ab = 0.f, ac = 0.f;
    for (long long b0 = (long long)blockIdx.y * FE_BT; b0 < a.B; b0 += (long long)gridDim.y * FE_BT) {
    const int nb = (int)min((long long)FE_BT, a.B - b0);
#pragma unroll 4
    for (int t = 0; t < nb; ++t) {
        const long long b = b0 + t;
        const float x = a.x[b * a.I + il];
        float go = live ? a.gout[b * IK + p] : 0.f;
        float dx;
        if constexpr (VARIANT == FETODE_FERRO_LOGISTIC) {
            const float s = sigm(k * (x - Ec));
            const float dz = go * 2.0f * s * (1.0f - s);                        // phi (1 - phi/2)
            ak = fmaf(dz, x - Ec, ak);
            aEc = fmaf(-dz, k, aEc);
            dx = dz * k;
        } else if constexpr (VARIANT == FETODE_FERRO_TWO_D) {
            const float pv = a.prev_is_zero ? 0.f : a.prev[b * a.I + il];
            const SoftBranch s = soft_branch(x, sigm(a.g * (x - pv)), k, Ec, 1.0f, a.g, a.alpha);
            float dm_dx, dm_dEc;
            soft_branch_grad(s, 1.0f, a.g, a.alpha, dm_dx, dm_dEc);
            ac = fmaf(go, fmaf(Ps, s.T, bias), ac);
            const float gP = go * coef;
            aPs = fmaf(gP, s.T, aPs);
            ab += gP;
            const float dz = gP * Ps * (1.0f - s.T * s.T);
            ak = fmaf(dz, fmaf(Ec, s.m, x), ak);
            aEc = fmaf(dz * k, fmaf(Ec, dm_dEc, s.m), aEc);
            dx = dz * k * fmaf(Ec, dm_dx, 1.0f);
        } else {
            const float pv = a.prev_is_zero ? 0.f : a.prev[il];
            const bool up = hard_gate(a.g * (x - pv));
            const float arg = up ? x - Ec : x + Ec;
            const float s = sigm(k * arg);
            if (a.apply_sigmoid) {
                const float y = sigm(fmaf(2.0f * Ps, s, -Ps) + bias);
                go *= y * (1.0f - y);
                if (a.apply_sigmoid == 2 && y != y) go = 0.f;                   // nan_to_num passes no gradient at a NaN
            }
            aPs = fmaf(go, 2.0f * s - 1.0f, aPs);
            ab += go;
            const float dz = go * 2.0f * Ps * s * (1.0f - s);
            ak = fmaf(dz, arg, ak);
            aEc += up ? -dz * k : dz * k;
            dx = dz * k;
        }
        if (a.gx) {
            // segmented inclusive scan from the right: after log2 steps the head lane of every run of equal i holds
            // the run's sum (runs are <= K long; K > 32 simply spans several warps' heads)
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const float w = __shfl_down_sync(0xffffffffu, dx, 1 << q);
                if (same[q]) dx += w;
            }
            if (head && live) atomicAdd(a.gx + b * a.I + i, dx);
        }
    }
    }
    if (!live) return;
    if (a.gk) atomicAdd(a.gk + p, ak);
    if (a.gEc) atomicAdd(a.gEc + p, aEc);
    if (VARIANT != FETODE_FERRO_LOGISTIC && a.gPs) atomicAdd(a.gPs + p, aPs);
    if (VARIANT != FETODE_FERRO_LOGISTIC && a.gbias) atomicAdd(a.gbias + p, ab);
    if (VARIANT == FETODE_FERRO_TWO_D && a.gcoef) atomicAdd(a.gcoef + p, ac);
}

// prev_out for HARDGATE: the LAST batch row (train_ecg_kan_fet_nn_ode.py:131-132)
__global__ void copy_row_kernel(const float* src, float* dst, int n) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) dst[t] = src[t];
}

static int ferro_check(const fetode_ferro_t* F) {
    if (!F) return FETODE_EINVAL;
    if (F->in_dim < 1 || F->num_basis < 1 || F->out_dim < 1) return FETODE_EINVAL;
    if (F->variant < FETODE_FERRO_FULL || F->variant > FETODE_FERRO_LOGISTIC) return FETODE_EINVAL;
    if (!(F->k && F->Ec)) return FETODE_EINVAL;
    if (F->variant != FETODE_FERRO_LOGISTIC && !(F->Ps && F->bias)) return FETODE_EINVAL;
    if (F->variant <= FETODE_FERRO_TWO_D && !F->coef) return FETODE_EINVAL;
    if (F->variant != FETODE_FERRO_FULL && F->out_dim != 1) return FETODE_ESHAPE;
    return FETODE_OK;
}

static FerroArgs ferro_args(const fetode_ferro_t* F, int64_t B) {
    FerroArgs a{};
    a.k = F->k; a.Ec = F->Ec; a.Ps = F->Ps; a.bias = F->bias; a.coef = F->coef;
    a.B = B; a.I = F->in_dim; a.O = F->out_dim; a.K = F->num_basis;
    a.g = F->gate_slope; a.alpha = F->alpha; a.apply_sigmoid = F->apply_sigmoid;
    a.pre_bound = F->variant == FETODE_FERRO_FULL ? F->pre_bound : 0.f;
    a.post_op = F->variant == FETODE_FERRO_FULL ? F->post_op : 0;
    return a;
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_ferro_fwd(const fetode_ferro_t* F, const float* x, const float* prev, int prev_is_zero,
                                const float* br, float* out, float* basis_out, float* br_out,
                                uint8_t* gate_out, float* prev_out, int64_t B, void* stream) {
    int rc = ferro_check(F);
    if (rc) return rc;
    FETODE_CHECK_ARG(x && out && B >= 0);
    if (B == 0) return FETODE_OK;
    if (!prev_is_zero) FETODE_CHECK_ARG(prev != nullptr);
    FETODE_CHECK_ARG(prev_is_zero >= 0 && prev_is_zero <= 2 && (prev_is_zero != 2 || F->variant == FETODE_FERRO_FULL));
    FETODE_CHECK_ARG(F->post_op >= 0 && F->post_op <= 2);
    cudaStream_t st = (cudaStream_t)stream;
    FerroArgs a = ferro_args(F, B);
    a.x = x; a.prev = prev; a.prev_is_zero = prev_is_zero; a.br = br; a.out = out; a.basis_out = basis_out;
    a.br_out = br_out; a.gate_out = gate_out;
    if (F->variant == FETODE_FERRO_FULL && a.pre_bound > 0.f) { a.xeff_out = prev_out; prev_out = nullptr; }
    if (F->variant == FETODE_FERRO_FULL) {
        FETODE_CHECK_SUPPORTED(B <= 0x7fffffffLL);
        const dim3 grid((unsigned)B, (unsigned)((a.O + 31) / 32));
        const size_t smem = (size_t)2 * a.I * sizeof(float);
        FETODE_CHECK_SUPPORTED(smem <= 48 * 1024);
        if (br) ferro_full_fwd_kernel<true><<<grid, FF_THREADS, smem, st>>>(a);
        else ferro_full_fwd_kernel<false><<<grid, FF_THREADS, smem, st>>>(a);
    } else {
        const long long by = std::min<long long>((B + FEF_ROWS - 1) / FEF_ROWS, 65535);
        const int IK = a.I * a.K;
        const int threads = std::min(256, (IK + 31) / 32 * 32);        // narrow layers (LV: I*K = 20) get narrow CTAs
        const dim3 blocks((unsigned)((IK + threads - 1) / threads), (unsigned)by);
        if (F->variant == FETODE_FERRO_TWO_D) ferro_elem_fwd_kernel<FETODE_FERRO_TWO_D><<<blocks, threads, 0, st>>>(a);
        else if (F->variant == FETODE_FERRO_LOGISTIC) ferro_elem_fwd_kernel<FETODE_FERRO_LOGISTIC><<<blocks, threads, 0, st>>>(a);
        else ferro_elem_fwd_kernel<FETODE_FERRO_HARDGATE><<<blocks, threads, 0, st>>>(a);
    }
    rc = FETODE_LAUNCH_RC();
    if (rc) return rc;
    if (prev_out) {
        if (F->variant == FETODE_FERRO_HARDGATE) {
            copy_row_kernel<<<(a.I + 127) / 128, 128, 0, st>>>(x + (B - 1) * a.I, prev_out, a.I);
            rc = FETODE_LAUNCH_RC();
        } else if (prev_out != x && F->variant != FETODE_FERRO_LOGISTIC) {
            rc = (int)cudaMemcpyAsync(prev_out, x, (size_t)B * a.I * sizeof(float), cudaMemcpyDeviceToDevice, st);
        }
    }
    return rc;
}

extern "C" int fetode_ferro_bwd(const fetode_ferro_t* F, const float* x, const float* prev, int prev_is_zero,
                                const float* br, const float* gout, float* gx,
                                const fetode_ferro_grads_t* grads, int64_t B, void* stream) {
    return fetode_ferro_bwd_fused(F, x, prev, prev_is_zero, br, gout, nullptr, gx, grads, B, stream);
}

extern "C" int fetode_ferro_bwd_fused(const fetode_ferro_t* F, const float* x, const float* prev, int prev_is_zero,
                                      const float* br, const float* gout, const float* yout, float* gx,
                                      const fetode_ferro_grads_t* grads, int64_t B, void* stream) {
    int rc = ferro_check(F);
    if (rc) return rc;
    FETODE_CHECK_ARG(x && gout && B >= 0);
    if (!prev_is_zero) FETODE_CHECK_ARG(prev != nullptr);
    if (F->variant == FETODE_FERRO_FULL && F->post_op) FETODE_CHECK_ARG(yout != nullptr);
    cudaStream_t st = (cudaStream_t)stream;
    FerroArgs a = ferro_args(F, B);
    a.yout = yout;
    a.x = x; a.prev = prev; a.prev_is_zero = prev_is_zero; a.br = br; a.gout = gout; a.gx = gx;
    if (grads) { a.gk = grads->k; a.gEc = grads->Ec; a.gPs = grads->Ps; a.gbias = grads->bias; a.gcoef = grads->coef; }
    if (gx) {
        cudaError_t e = cudaMemsetAsync(gx, 0, (size_t)B * a.I * sizeof(float), st);
        if (e) return (int)e;
    }
    if (B == 0) return FETODE_OK;
    if (F->variant == FETODE_FERRO_FULL) {
        const long long bz = (B + FB_BT - 1) / FB_BT;
        FETODE_CHECK_SUPPORTED(bz <= 65535 && (a.O + 31) / 32 <= 65535);
        const dim3 grid((unsigned)a.I, (unsigned)((a.O + 31) / 32), (unsigned)bz);
        if (br) ferro_full_bwd_kernel<true><<<grid, FF_THREADS, 0, st>>>(a);
        else ferro_full_bwd_kernel<false><<<grid, FF_THREADS, 0, st>>>(a);
    } else {
        const long long by = std::min<long long>((B + FE_BT - 1) / FE_BT, 65535);
        const dim3 grid((unsigned)((a.I * a.K + 127) / 128), (unsigned)by);
        if (F->variant == FETODE_FERRO_TWO_D) ferro_elem_bwd_kernel<FETODE_FERRO_TWO_D><<<grid, 128, 0, st>>>(a);
        else if (F->variant == FETODE_FERRO_LOGISTIC) ferro_elem_bwd_kernel<FETODE_FERRO_LOGISTIC><<<grid, 128, 0, st>>>(a);
        else ferro_elem_bwd_kernel<FETODE_FERRO_HARDGATE><<<grid, 128, 0, st>>>(a);
    }
    return FETODE_LAUNCH_RC();
}

// ---- OriginalBatchedFerroelectricBasis (ferro_class.py:9-104) -----------------------------------------------------
// A sequential scan over the batch: sample b sees prev_x = x[b-1] and the branch momentum left by sample b-1, and --
// unlike every other variant -- that momentum IS the carried branch_sign (:73) and stays on the autograd graph, so the
// gradient of sample b reaches Ec and x of all earlier samples through a linear recurrence
//     br_b = A_b br_{b-1} + C_b ,  A_b = alpha + (1-alpha)(1 - su_b - sl_b).
// CTA = one output o; thread = elements (i, j) of that output, their br / prev in registers; the B steps run in-kernel
// with one block reduction per sample for out[b, o].  The backward walks the samples in reverse with the adjoint of br
// in a register; every parameter gradient is private to its thread (no reduction), gx needs one atomic per element.
constexpr int FS_THREADS = 256, FS_EPT = 4;               // up to 1024 (i, j) elements per output

struct FerroSeqArgs {
    const float *k, *Ec, *Ps, *bias, *coef;
    const float *x, *prev0, *br0, *gout, *br_hist_in;
    float *out, *basis_out, *br_hist, *br_final, *gx;
    float *gk, *gEc, *gPs, *gbias, *gcoef;
    long long B;
    int I, O, K;
    float g, alpha;
};

__global__ void __launch_bounds__(FS_THREADS) ferro_seq_fwd_kernel(const FerroSeqArgs a) {
    __shared__ float part[2][FS_THREADS / 32];
    const int o = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int I = a.I, O = a.O, K = a.K, n = I * K;
    float k[FS_EPT], Ec[FS_EPT], Ps[FS_EPT], bs[FS_EPT], cf[FS_EPT], br[FS_EPT], pv[FS_EPT];
    int ii[FS_EPT];
    size_t base[FS_EPT];
#pragma unroll
    for (int e = 0; e < FS_EPT; ++e) {
        const int q = tid + e * FS_THREADS;
        ii[e] = -1;
        if (q < n) {
            const int i = q / K, j = q % K;
            ii[e] = i;
            base[e] = ((size_t)i * O + o) * K + j;
            k[e] = a.k[base[e]]; Ec[e] = a.Ec[base[e]]; Ps[e] = a.Ps[base[e]]; bs[e] = a.bias[base[e]]; cf[e] = a.coef[base[e]];
            br[e] = a.br0 ? a.br0[base[e]] : 1.0f;
            pv[e] = a.prev0 ? a.prev0[i] : 0.f;
        }
    }
    const size_t IOK = (size_t)I * O * K;
    for (long long b = 0; b < a.B; ++b) {
        float acc = 0.f;
#pragma unroll
        for (int e = 0; e < FS_EPT; ++e) {
            if (ii[e] < 0) continue;
            const float x = a.x[b * I + ii[e]];
            if (a.br_hist) a.br_hist[b * IOK + base[e]] = br[e];                 // branch_sign BEFORE this sample
            const SoftBranch s = soft_branch(x, sigm(a.g * (x - pv[e])), k[e], Ec[e], br[e], a.g, a.alpha);
            const float P = fmaf(Ps[e], s.T, bs[e]);
            acc = fmaf(cf[e], P, acc);
            if (a.basis_out) a.basis_out[b * IOK + base[e]] = P;
            br[e] = s.m;                                                       // :73  branch_sign <- momentum
            pv[e] = x;                                                         // :87
        }
        acc = warp_sum(acc);
        if (lane == 0) part[b & 1][w] = acc;
        __syncthreads();
        if (tid == 0) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < FS_THREADS / 32; ++q) t += part[b & 1][q];
            a.out[b * O + o] = t;
        }
    }
#pragma unroll
    for (int e = 0; e < FS_EPT; ++e)
        if (ii[e] >= 0 && a.br_final) a.br_final[base[e]] = br[e];
}

__global__ void __launch_bounds__(FS_THREADS) ferro_seq_bwd_kernel(const FerroSeqArgs a) {
    const int o = blockIdx.x, tid = threadIdx.x;
    const int I = a.I, O = a.O, K = a.K, n = I * K;
    const size_t IOK = (size_t)I * O * K;
#pragma unroll 1
    for (int e = 0; e < FS_EPT; ++e) {
        const int q = tid + e * FS_THREADS;
        if (q >= n) break;
        const int i = q / K, j = q % K;
        const size_t base = ((size_t)i * O + o) * K + j;
        const float k = a.k[base], Ec = a.Ec[base], Ps = a.Ps[base], bias = a.bias[base], coef = a.coef[base];
        float ak = 0.f, aEc = 0.f, aPs = 0.f, ab = 0.f, ac = 0.f, lam = 0.f;     // lam: adjoint of the carried branch_sign
        for (long long b = a.B - 1; b >= 0; --b) {
            const float x = a.x[b * I + i];
            const float pv = b > 0 ? a.x[(b - 1) * I + i] : (a.prev0 ? a.prev0[i] : 0.f);
            const float br = a.br_hist_in[b * IOK + base];
            const SoftBranch s = soft_branch(x, sigm(a.g * (x - pv)), k, Ec, br, a.g, a.alpha);
            float dm_dx, dm_dEc;
            soft_branch_grad(s, br, a.g, a.alpha, dm_dx, dm_dEc);
            const float go = a.gout[b * O + o];
            ac = fmaf(go, fmaf(Ps, s.T, bias), ac);
            const float gP = go * coef;
            aPs = fmaf(gP, s.T, aPs);
            ab += gP;
            const float dz = gP * Ps * (1.0f - s.T * s.T);
            ak = fmaf(dz, fmaf(Ec, s.m, x), ak);
            const float mu = fmaf(dz * k, Ec, lam);                             // adjoint of this sample's momentum
            aEc = fmaf(dz * k, s.m, fmaf(mu, dm_dEc, aEc));
            if (a.gx) atomicAdd(a.gx + b * I + i, fmaf(mu, dm_dx, dz * k));
            lam = mu * (a.alpha + (1.0f - a.alpha) * (1.0f - s.su - s.sl));     // d m / d br_{b-1}
        }
        if (a.gk) atomicAdd(a.gk + base, ak);
        if (a.gEc) atomicAdd(a.gEc + base, aEc);
        if (a.gPs) atomicAdd(a.gPs + base, aPs);
        if (a.gbias) atomicAdd(a.gbias + base, ab);
        if (a.gcoef) atomicAdd(a.gcoef + base, ac);
    }
}

// ---- FerroelectricBasisConv2d / MemEfficient_FerroelectricBasisConv2d (ferro_class.py:601-944) -------------------
static int conv_check(const fetode_ferro_t* F, const fetode_conv2d_t* G) {
    if (!G) return FETODE_EINVAL;
    if (F->variant != FETODE_FERRO_FULL) return FETODE_EINVAL;
    if (G->Cin < 1 || G->H < 1 || G->W < 1 || G->kH < 1 || G->kW < 1 || G->stride < 1 || G->padding < 0) return FETODE_EINVAL;
    if (G->Hout != (G->H + 2 * G->padding - G->kH) / G->stride + 1) return FETODE_ESHAPE;
    if (G->Wout != (G->W + 2 * G->padding - G->kW) / G->stride + 1) return FETODE_ESHAPE;
    if (G->Hout < 1 || G->Wout < 1) return FETODE_ESHAPE;
    if (F->in_dim != G->Cin * G->kH * G->kW) return FETODE_ESHAPE;
    return FETODE_OK;
}

extern "C" int fetode_ferro_conv2d_fwd(const fetode_ferro_t* F, const fetode_conv2d_t* G, const float* x,
                                       const float* prev, int prev_mode, const float* br, const float* out_bias,
                                       float* out, float* basis_out, float* br_out, int br_momentum, int64_t B,
                                       void* stream) {
    int rc = ferro_check(F);
    if (rc) return rc;
    rc = conv_check(F, G);
    if (rc) return rc;
    FETODE_CHECK_ARG(x && out && B >= 0 && prev_mode >= 0 && prev_mode <= 2 && (prev_mode != 0 || prev));
    if (B == 0) return FETODE_OK;
    const long long rows = (long long)B * G->Hout * G->Wout;
    FETODE_CHECK_SUPPORTED(rows <= 0x7fffffffLL);
    FerroArgs a = ferro_args(F, rows);
    a.pre_bound = 0.f; a.post_op = 0;
    a.x = x; a.prev = prev; a.prev_is_zero = prev_mode; a.br = br; a.out = out; a.basis_out = basis_out;
    a.br_out = br_out; a.br_momentum = br_momentum; a.cv = *G; a.conv = 1; a.out_bias = out_bias;
    const dim3 grid((unsigned)rows, (unsigned)((a.O + 31) / 32));
    const size_t smem = (size_t)2 * a.I * sizeof(float);
    FETODE_CHECK_SUPPORTED(smem <= 48 * 1024);
    cudaStream_t st = (cudaStream_t)stream;
    if (br) ferro_full_fwd_kernel<true><<<grid, FF_THREADS, smem, st>>>(a);
    else ferro_full_fwd_kernel<false><<<grid, FF_THREADS, smem, st>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_ferro_conv2d_bwd(const fetode_ferro_t* F, const fetode_conv2d_t* G, const float* x,
                                       const float* prev, int prev_mode, const float* br, const float* gout,
                                       float* gx, const fetode_ferro_grads_t* grads, float* g_out_bias, int64_t B,
                                       void* stream) {
    int rc = ferro_check(F);
    if (rc) return rc;
    rc = conv_check(F, G);
    if (rc) return rc;
    FETODE_CHECK_ARG(x && gout && B >= 0 && prev_mode >= 0 && prev_mode <= 2 && (prev_mode != 0 || prev));
    cudaStream_t st = (cudaStream_t)stream;
    if (gx) {
        cudaError_t e = cudaMemsetAsync(gx, 0, (size_t)B * G->Cin * G->H * G->W * sizeof(float), st);
        if (e) return (int)e;
    }
    if (B == 0) return FETODE_OK;
    const long long rows = (long long)B * G->Hout * G->Wout;
    FerroArgs a = ferro_args(F, rows);
    a.pre_bound = 0.f; a.post_op = 0;
    a.x = x; a.prev = prev; a.prev_is_zero = prev_mode; a.br = br; a.gout = gout; a.gx = gx;
    a.cv = *G; a.conv = 1; a.g_out_bias = g_out_bias;
    if (grads) { a.gk = grads->k; a.gEc = grads->Ec; a.gPs = grads->Ps; a.gbias = grads->bias; a.gcoef = grads->coef; }
    const long long bz = (rows + FB_BT - 1) / FB_BT;
    FETODE_CHECK_SUPPORTED(bz <= 65535 && (a.O + 31) / 32 <= 65535);
    const dim3 grid((unsigned)a.I, (unsigned)((a.O + 31) / 32), (unsigned)bz);
    if (br) ferro_full_bwd_kernel<true><<<grid, FF_THREADS, 0, st>>>(a);
    else ferro_full_bwd_kernel<false><<<grid, FF_THREADS, 0, st>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_ferro_seq_fwd(const fetode_ferro_t* F, const float* x, const float* prev0, const float* br0,
                                    float* out, float* basis_out, float* br_hist, float* br_final, int64_t B,
                                    void* stream) {
    int rc = ferro_check(F);
    if (rc) return rc;
    FETODE_CHECK_ARG(F->variant == FETODE_FERRO_FULL && x && out && B >= 0);
    FETODE_CHECK_SUPPORTED(F->in_dim * F->num_basis <= FS_THREADS * FS_EPT);
    if (B == 0) return FETODE_OK;
    FerroSeqArgs a{};
    a.k = F->k; a.Ec = F->Ec; a.Ps = F->Ps; a.bias = F->bias; a.coef = F->coef;
    a.x = x; a.prev0 = prev0; a.br0 = br0; a.out = out; a.basis_out = basis_out; a.br_hist = br_hist; a.br_final = br_final;
    a.B = B; a.I = F->in_dim; a.O = F->out_dim; a.K = F->num_basis; a.g = F->gate_slope; a.alpha = F->alpha;
    ferro_seq_fwd_kernel<<<(unsigned)a.O, FS_THREADS, 0, (cudaStream_t)stream>>>(a);
    return FETODE_LAUNCH_RC();
}

extern "C" int fetode_ferro_seq_bwd(const fetode_ferro_t* F, const float* x, const float* prev0, const float* br_hist,
                                    const float* gout, float* gx, const fetode_ferro_grads_t* grads, int64_t B,
                                    void* stream) {
    int rc = ferro_check(F);
    if (rc) return rc;
    FETODE_CHECK_ARG(F->variant == FETODE_FERRO_FULL && x && br_hist && gout && B >= 0);
    FETODE_CHECK_SUPPORTED(F->in_dim * F->num_basis <= FS_THREADS * FS_EPT);
    cudaStream_t st = (cudaStream_t)stream;
    if (gx) {
        cudaError_t e = cudaMemsetAsync(gx, 0, (size_t)B * F->in_dim * sizeof(float), st);
        if (e) return (int)e;
    }
    if (B == 0) return FETODE_OK;
    FerroSeqArgs a{};
    a.k = F->k; a.Ec = F->Ec; a.Ps = F->Ps; a.bias = F->bias; a.coef = F->coef;
    a.x = x; a.prev0 = prev0; a.br_hist_in = br_hist; a.gout = gout; a.gx = gx;
    if (grads) { a.gk = grads->k; a.gEc = grads->Ec; a.gPs = grads->Ps; a.gbias = grads->bias; a.gcoef = grads->coef; }
    a.B = B; a.I = F->in_dim; a.O = F->out_dim; a.K = F->num_basis; a.g = F->gate_slope; a.alpha = F->alpha;
    ferro_seq_bwd_kernel<<<(unsigned)a.O, FS_THREADS, 0, st>>>(a);
    return FETODE_LAUNCH_RC();
}
