// kanlinear_bwd.cu -- backward of KANLinear of any width on the FP32 CUDA-core pipe (sm_100a).
//
// Replaces the autograd graph PyTorch records for KANLinear.forward (efficient_kan/efficientkan.py:160-182,
// b_splines :117-131, LogisticBasis.forward :20-24; KANFET branch TwoDimensionFerroelectricBasis.forward
// ferro_class.py:547-592; MNIST copy mnist_kuramoto_kan.py:112-142).  Gradients follow SURVEY appendix A.3/A.4:
// the hysteresis snapshot `prev` is a constant of the graph (ferro_class.py:553-554), `grid` gets none.
//
// Three launches over the folded contraction  y = A(x) . wcat + bias   (A never touches HBM):
//   1. kl_bwd_dw_kernel   gW[(i,f),o] = sum_b A[b,(i,f)] gy[b,o]       (features rebuilt in the prologue)
//      kl_colsum_kernel   gsum[o]     = sum_b gy[b,o]                  (gradient of every folded constant)
//   2. kl_bwd_dx_kernel   G[b,(i,f)]  = sum_o gy[b,o] wcat[(i,f),o] ;  gx[b,i] = sum_f G dA/dx ; the branch-basis
//                         parameter gradients (a,b | k,Ec) are reduced over the rows of the tile here
//   3. kl_unfold_kernel   chain rule from the folded weights back to base_weight / spline_weight /
//                         spline_scaler / logistic_weight / logistic_scaler / logistic_bias / Ps,bias,coef
#include <mutex>

#include "kanlinear.cuh"

namespace fetode {

constexpr int KB_BM = 64, KB_THREADS = 256, KB_KCMAX = 128, KB_BN = 64, KB_OC = 64;

struct KLBwdArgs {
    fetode_kanlayer_t P;
    fetode_kanlayer_grads_t g;
    const float* blob;
    const float* x;
    const float* prev;
    const float* gy;
    float* gx;
    float* ws;           // [I*F, O] folded-weight gradient, then gsum [O]
    long long B;
    KLLayout L;
    float gl;
    int prev_is_zero, KI, split;
    int accumulate;      // 1: ws already holds a partial folded gradient (deferred unfold): add to it, do not unfold
};

// ------------------------------------------------------------------------------------------
// 1. gW = A^T gy
// ------------------------------------------------------------------------------------------
template <int BR>
__global__ void __launch_bounds__(KB_THREADS) kl_bwd_dw_kernel(const KLBwdArgs a) {
    extern __shared__ float4 kb_smem4[];
    float* As = reinterpret_cast<float*>(kb_smem4);       // [KC][64]
    const KLLayout& L = a.L;
    const int F = L.F, I = L.I, O = L.O, KI = a.KI, KC = KI * F;
    float* Gs = As + KB_KCMAX * KB_BM;                    // [64][KB_BN]
    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * KI, n0 = blockIdx.y * KB_BN;
    const int ty = tid / 16, tx = tid % 16;
    const int frow = tid % KB_BM, fii = tid / KB_BM;

    float acc[8][4];
#pragma unroll
    for (int q = 0; q < 8; ++q)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[q][c] = 0.f;

    const long long ntile = (a.B + KB_BM - 1) / KB_BM;
    for (long long tile = blockIdx.z; tile < ntile; tile += gridDim.z) {
        const long long row0 = tile * KB_BM;
        for (int ii = fii; ii < KI; ii += KB_THREADS / KB_BM) {
            const int i = i0 + ii;
            const long long b = row0 + frow;
            float* col = As + (ii * F) * KB_BM + frow;
            if (i < I && b < a.B) {
                const float x = a.x[b * I + i];
                float prev = 0.f;
                if constexpr (BR == FETODE_BRANCH_FERRO2D)
                    if (!a.prev_is_zero) prev = a.prev[b * I + i];
                kl_features<BR>(x, prev, a.gl, a.blob + L.tab + (size_t)i * L.TAB, L, col, KB_BM);
            } else {
                for (int f = 0; f < F; ++f) col[f * KB_BM] = 0.f;
            }
        }
        for (int idx = tid; idx < KB_BM * KB_BN; idx += KB_THREADS) {
            const int r = idx / KB_BN, n = idx % KB_BN;
            float v = 0.f;
            if (row0 + r < a.B && n0 + n < O) v = a.gy[(row0 + r) * O + n0 + n];
            Gs[r * KB_BN + n] = v;
        }
        __syncthreads();
#pragma unroll 4
        for (int r = 0; r < KB_BM; ++r) {
            const float4 gv = *reinterpret_cast<const float4*>(Gs + r * KB_BN + tx * 4);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int kk = ty + 16 * q;
                if (kk < KC) {
                    const float av = As[kk * KB_BM + r];
                    acc[q][0] = fmaf(av, gv.x, acc[q][0]);
                    acc[q][1] = fmaf(av, gv.y, acc[q][1]);
                    acc[q][2] = fmaf(av, gv.z, acc[q][2]);
                    acc[q][3] = fmaf(av, gv.w, acc[q][3]);
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int kk = ty + 16 * q;
        if (kk >= KC) continue;
        const int i = i0 + kk / F;
        if (i >= I) continue;
        float* dst = a.ws + ((size_t)i * F + kk % F) * O;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int n = n0 + tx * 4 + c;
            if (n < O) {
                if (gridDim.z == 1 && !a.accumulate) dst[n] = acc[q][c];
                else atomicAdd(dst + n, acc[q][c]);
            }
        }
    }
}

// gsum[o] = sum_b gy[b,o]
__global__ void __launch_bounds__(256) kl_colsum_kernel(const float* gy, float* gsum, long long B, int O) {
    const int o = blockIdx.x * 32 + (threadIdx.x & 31);
    const int w = threadIdx.x >> 5;
    float acc = 0.f;
    if (o < O)
        for (long long b = (long long)blockIdx.y * 8 + w; b < B; b += (long long)gridDim.y * 8) acc += gy[b * O + o];
    __shared__ float part[8][33];
    part[w][threadIdx.x & 31] = acc;
    __syncthreads();
    if (w == 0 && o < O) {
        float s = 0.f;
        for (int k = 0; k < 8; ++k) s += part[k][threadIdx.x];
        atomicAdd(gsum + o, s);
    }
}

// ------------------------------------------------------------------------------------------
// 2. G = gy wcat^T ;  gx = sum_f G dA/dx ;  branch-basis parameter gradients
// ------------------------------------------------------------------------------------------
constexpr int KB_GST = KB_BM + 4;      // row stride of the transposed gy tile (keeps float4 alignment)

template <int BR>
__global__ void __launch_bounds__(KB_THREADS) kl_bwd_dx_kernel(const KLBwdArgs a) {
    extern __shared__ float4 kb_smem4[];
    float* GsT = reinterpret_cast<float*>(kb_smem4);       // [KB_OC][KB_GST]   gy tile, [o][row]
    float* Ws = GsT + KB_OC * KB_GST;                      // [KC][KB_OC+1]
    float* Gt = Ws + KB_KCMAX * (KB_OC + 1);               // [KC][64]
    const KLLayout& L = a.L;
    const int F = L.F, I = L.I, O = L.O, S = L.S, NB = L.NB, KI = a.KI, KC = KI * F;
    const int tid = threadIdx.x;
    const int ty = tid / 16, tx = tid % 16;
    const int frow = tid % KB_BM, fii = tid / KB_BM;
    const long long row0 = (long long)blockIdx.x * KB_BM;
    const float* wcat = a.blob + L.wcat;
    const int ngroups = (I + KI - 1) / KI;

    for (int grp = blockIdx.y; grp < ngroups; grp += gridDim.y) {
        const int i0 = grp * KI;
        float acc[4][8];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[r][q] = 0.f;
        for (int o0 = 0; o0 < O; o0 += KB_OC) {
            for (int idx = tid; idx < KB_BM * KB_OC; idx += KB_THREADS) {
                const int r = idx / KB_OC, n = idx % KB_OC;
                float v = 0.f;
                if (row0 + r < a.B && o0 + n < O) v = a.gy[(row0 + r) * O + o0 + n];
                GsT[n * KB_GST + r] = v;
            }
            for (int idx = tid; idx < KC * KB_OC; idx += KB_THREADS) {
                const int kk = idx / KB_OC, n = idx % KB_OC;
                const int i = i0 + kk / F;
                float w = 0.f;
                if (i < I && o0 + n < O) w = wcat[((size_t)i * F + kk % F) * O + o0 + n];
                Ws[kk * (KB_OC + 1) + n] = w;
            }
            __syncthreads();
#pragma unroll 4
            for (int n = 0; n < KB_OC; ++n) {
                const float4 gv = *reinterpret_cast<const float4*>(GsT + n * KB_GST + ty * 4);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int kk = tx + 16 * q;
                    if (kk < KC) {
                        const float w = Ws[kk * (KB_OC + 1) + n];
                        acc[0][q] = fmaf(gv.x, w, acc[0][q]);
                        acc[1][q] = fmaf(gv.y, w, acc[1][q]);
                        acc[2][q] = fmaf(gv.z, w, acc[2][q]);
                        acc[3][q] = fmaf(gv.w, w, acc[3][q]);
                    }
                }
            }
            __syncthreads();
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int kk = tx + 16 * q;
            if (kk < KC) {
#pragma unroll
                for (int r = 0; r < 4; ++r) Gt[kk * KB_BM + ty * 4 + r] = acc[r][q];
            }
        }
        __syncthreads();
        // ---- owner phase: thread (frow, ii) consumes the F values G[row, (i, .)]
        for (int ii = fii; ii < KI; ii += KB_THREADS / KB_BM) {
            const int i = i0 + ii;
            const long long b = row0 + frow;
            const bool live = i < I && b < a.B;      // whole warps share `i`; rows may be ragged
            if (i >= I) continue;
            const float* G = Gt + (ii * F) * KB_BM + frow;
            const float* tab = a.blob + L.tab + (size_t)i * L.TAB;
            const float x = live ? a.x[b * I + i] : 0.f;
            float gx = 0.f;
            {   // silu'(x) = s (1 + x (1 - s))
                const float s = rcp_approx(1.0f + ex2_approx(-FETODE_LOG2E * x));
                gx = G[0] * s * fmaf(x, 1.0f - s, 1.0f);
            }
            const float4 gi = *reinterpret_cast<const float4*>(tab + L.t_gi);
            if (x >= gi.x && x < gi.z) {
                const float* kp = tab + L.t_kn;
                const float* rd1 = tab + L.t_rd1; const float* rd2 = tab + L.t_rd2; const float* rd3 = tab + L.t_rd3;
                const int m = bspline_interval(x, kp, gi.x, gi.y, L.NK);
                const int mp = m + 3;
                // order-2 values N^2_{m-2..m}, then N'^3_j = 3 (N^2_j rd3[j] - N^2_{j+1} rd3[j+1])
                const float left1 = x - kp[mp], right1 = kp[mp + 1] - x;
                const float left2 = x - kp[mp - 1], right2 = kp[mp + 2] - x;
                float t = rd1[mp];
                const float a0 = right1 * t, a1 = left1 * t;
                t = a0 * rd2[mp - 1];
                const float b0 = right1 * t;
                const float saved = left2 * t;
                t = a1 * rd2[mp];
                const float b1 = saved + right2 * t;
                const float b2 = left1 * t;
                const float e0 = 3.0f * b0 * rd3[mp - 2], e1 = 3.0f * b1 * rd3[mp - 1], e2 = 3.0f * b2 * rd3[mp];
                const float dv[4] = {-e0, e0 - e1, e1 - e2, e2};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int s = m - 3 + c;
                    if (s >= 0 && s < S) gx = fmaf(G[(1 + s) * KB_BM], dv[c], gx);
                }
            }
            if constexpr (BR == FETODE_BRANCH_LOGISTIC) {
                const float* pa = a.P.bp[0] + (size_t)i * NB;
                const float* pb = a.P.bp[1] + (size_t)i * NB;
                for (int j = 0; j < NB; ++j) {
                    const float aj = pa[j], bj = pb[j];
                    const float r = rcp_approx(1.0f + ex2_approx(fminf(-FETODE_LOG2E * aj * (x - bj), FETODE_EXP_CLAMP)));
                    float e = live ? G[(1 + S + j) * KB_BM] * r * (1.0f - r) : 0.f;
                    gx = fmaf(e, aj, gx);
                    if (a.g.bp[0]) { const float v = warp_sum(e * (x - bj)); if ((tid & 31) == 0) atomicAdd(a.g.bp[0] + (size_t)i * NB + j, v); }
                    if (a.g.bp[1]) { const float v = warp_sum(-e * aj); if ((tid & 31) == 0) atomicAdd(a.g.bp[1] + (size_t)i * NB + j, v); }
                }
            } else if constexpr (BR == FETODE_BRANCH_FERRO2D) {
                const float g = a.P.gate_slope, al2 = 2.0f * (1.0f - a.P.alpha);
                float prev = 0.f;
                if (live && !a.prev_is_zero) prev = a.prev[b * I + i];
                const float omu = rcp_approx(1.0f + ex2_approx(a.gl * (x - prev)));        // 1 - is_moving_up
                const float domu = -g * omu * (1.0f - omu);
                const float* pk = a.P.bp[0] + (size_t)i * NB;
                const float* pE = a.P.bp[1] + (size_t)i * NB;
                for (int j = 0; j < NB; ++j) {
                    const float k = pk[j], Ec = pE[j];
                    const float cn = rcp_approx(1.0f + ex2_approx(fminf(a.gl * (x + Ec), FETODE_EXP_CLAMP)));
                    const float dcn = -g * cn * (1.0f - cn);                                // d/dx == d/dEc
                    const float sl = omu * cn;
                    const float m = 1.0f - al2 * sl;
                    const float dm_dx = -al2 * (domu * cn + omu * dcn);
                    const float dm_dE = -al2 * omu * dcn;
                    const float z2 = 2.0f * FETODE_LOG2E * k * fmaf(Ec, m, x);
                    const float r = rcp_approx(1.0f + ex2_approx(fminf(z2, FETODE_EXP_CLAMP)));
                    float e = live ? G[(1 + S + j) * KB_BM] * (-2.0f) * r * (1.0f - r) : 0.f;   // dL/dz
                    gx = fmaf(e * k, fmaf(Ec, dm_dx, 1.0f), gx);
                    if (a.g.bp[0]) { const float v = warp_sum(e * fmaf(Ec, m, x)); if ((tid & 31) == 0) atomicAdd(a.g.bp[0] + (size_t)i * NB + j, v); }
                    if (a.g.bp[1]) { const float v = warp_sum(e * k * fmaf(Ec, dm_dE, m)); if ((tid & 31) == 0) atomicAdd(a.g.bp[1] + (size_t)i * NB + j, v); }
                }
            }
            if (live && a.gx) a.gx[b * I + i] = gx;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// 3. unfold: folded-weight gradient -> the reference's parameters
// ------------------------------------------------------------------------------------------
// CTA = 8 inputs x 32 outputs.  The folded gradient is read row-wise (output index contiguous), transposed through
// shared memory, and the reference-layout gradients ([O, I, S], [O, I*NB], [O, I]) are written with the (input, s | j)
// index contiguous, so both sides of the transpose are coalesced.
constexpr int UF_TI = 8, UF_TO = 32, UF_THREADS = 256, UF_PAD = UF_TO + 1;

__global__ void __launch_bounds__(UF_THREADS) kl_unfold_kernel(const KLBwdArgs a) {
    extern __shared__ float uf_smem[];
    const KLLayout& L = a.L;
    const int I = L.I, O = L.O, F = L.F, S = L.S, NB = L.NB;
    const fetode_kanlayer_t& P = a.P;
    float* tile = uf_smem;                                   // [UF_TI][F][UF_PAD]
    float* red_ls = tile + UF_TI * F * UF_PAD;               // [UF_TO]           branch_scaler partial sums
    float* red_bp = red_ls + UF_TO;                          // [3][UF_TI][NB]    Ps, bias, coef partial sums
    const int tid = threadIdx.x;
    const int o0 = blockIdx.x * UF_TO, i0 = blockIdx.y * UF_TI;
    const float* gsum = a.ws + (size_t)I * F * O;
    for (int q = tid; q < UF_TO + 3 * UF_TI * NB; q += UF_THREADS) red_ls[q] = 0.f;
    for (int q = tid; q < UF_TI * F * UF_TO; q += UF_THREADS) {
        const int oo = q % UF_TO, r = q / UF_TO;             // r = ii * F + f
        const int i = i0 + r / F, o = o0 + oo;
        tile[r * UF_PAD + oo] = (i < I && o < O) ? a.ws[((size_t)i0 * F + r) * O + o] : 0.f;
    }
    __syncthreads();
    // ---- base_weight [O, I]
    for (int q = tid; q < UF_TO * UF_TI; q += UF_THREADS) {
        const int ii = q % UF_TI, oo = q / UF_TI;
        const int i = i0 + ii, o = o0 + oo;
        if (i < I && o < O && a.g.base_weight) a.g.base_weight[(size_t)o * I + i] += tile[(ii * F) * UF_PAD + oo];
    }
    // ---- spline_weight [O, I, S] and spline_scaler [O, I] (S == 8 lanes per (o, i): shuffle reduction)
    for (int q0 = 0; q0 < UF_TO * UF_TI * S; q0 += UF_THREADS) {
        const int q = q0 + tid;
        const int s = q % S, ii = (q / S) % UF_TI, oo = q / (S * UF_TI);
        const int i = i0 + ii, o = o0 + oo;
        const bool ok = oo < UF_TO && i < I && o < O;
        float gsc = 0.f;
        if (ok) {
            const size_t oi = (size_t)o * I + i;
            const float g = tile[(ii * F + 1 + s) * UF_PAD + oo];
            const float sc = P.spline_scaler ? P.spline_scaler[oi] : 1.0f;
            if (a.g.spline_weight) a.g.spline_weight[oi * S + s] += g * sc;
            gsc = g * P.spline_weight[oi * S + s];
        }
        if (S == 8) {
            gsc += __shfl_xor_sync(0xffffffffu, gsc, 1);
            gsc += __shfl_xor_sync(0xffffffffu, gsc, 2);
            gsc += __shfl_xor_sync(0xffffffffu, gsc, 4);
            if (ok && s == 0 && a.g.spline_scaler && P.spline_scaler) a.g.spline_scaler[(size_t)o * I + i] += gsc;
        } else if (ok && a.g.spline_scaler && P.spline_scaler) {
            atomicAdd(a.g.spline_scaler + (size_t)o * I + i, gsc);
        }
    }
    // ---- branch_weight [O, I*NB], branch_scaler [O], branch_bias [O], Ps / bias / coef [I, NB]
    if (P.branch != FETODE_BRANCH_NONE) {
        for (int q = tid; q < UF_TO * UF_TI * NB; q += UF_THREADS) {
            const int j = q % NB, ii = (q / NB) % UF_TI, oo = q / (NB * UF_TI);
            const int i = i0 + ii, o = o0 + oo;
            if (i >= I || o >= O) continue;
            const float lsc = P.scale_logistic * (P.branch_scaler ? P.branch_scaler[o] : 1.0f);
            const float go = gsum[o];
            const float g = tile[(ii * F + 1 + S + j) * UF_PAD + oo];
            const size_t wi = (size_t)o * I * NB + (size_t)i * NB + j;
            const float wraw = P.branch_weight[wi];
            float dw;                                   // dL / d(w = wraw * lsc)
            if (P.branch == FETODE_BRANCH_FERRO2D) {
                const int ij = i * NB + j;
                const float Ps = P.bp[2][ij], bias = P.bp[3][ij], coef = P.bp[4][ij];
                dw = g * (-2.0f * coef * Ps) + go * coef * (Ps + bias);
                const float w = wraw * lsc;
                float* rb = red_bp + ii * NB + j;
                if (a.g.bp[2]) atomicAdd(rb, g * (-2.0f * w * coef) + go * w * coef);
                if (a.g.bp[3]) atomicAdd(rb + UF_TI * NB, go * w * coef);
                if (a.g.bp[4]) atomicAdd(rb + 2 * UF_TI * NB, g * (-2.0f * w * Ps) + go * w * (Ps + bias));
            } else {
                dw = 2.0f * g;
            }
            if (a.g.branch_weight) a.g.branch_weight[wi] += dw * lsc;
            if (a.g.branch_scaler && P.branch_scaler) atomicAdd(red_ls + oo, dw * wraw * P.scale_logistic);
        }
        __syncthreads();
        if (a.g.branch_scaler && P.branch_scaler)
            for (int oo = tid; oo < UF_TO; oo += UF_THREADS)
                if (o0 + oo < O) atomicAdd(a.g.branch_scaler + o0 + oo, red_ls[oo]);
        if (P.branch == FETODE_BRANCH_FERRO2D)
            for (int q = tid; q < 3 * UF_TI * NB; q += UF_THREADS) {
                const int which = q / (UF_TI * NB), r = q % (UF_TI * NB);
                const int i = i0 + r / NB;
                if (i < I && a.g.bp[2 + which]) atomicAdd(a.g.bp[2 + which] + (size_t)i * NB + r % NB, red_bp[q]);
            }
        if (a.g.branch_bias && P.branch_bias && blockIdx.y == 0)
            for (int oo = tid; oo < UF_TO; oo += UF_THREADS)
                if (o0 + oo < O) a.g.branch_bias[o0 + oo] += gsum[o0 + oo];
    }
}

static void kl_unfold_run(const KLBwdArgs& a, cudaStream_t st) {
    const KLLayout& L = a.L;
    const size_t smem = (size_t)(UF_TI * L.F * UF_PAD + UF_TO + 3 * UF_TI * std::max(L.NB, 1)) * 4;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kl_unfold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kl_unfold_kernel<<<dim3((L.O + UF_TO - 1) / UF_TO, (L.I + UF_TI - 1) / UF_TI), UF_THREADS, smem, st>>>(a);
}

template <int BR>
static int kl_bwd_launch(KLBwdArgs& a, cudaStream_t st) {
    const KLLayout& L = a.L;
    const long long ntile = (a.B + KB_BM - 1) / KB_BM;
    const int ngroups = (L.I + a.KI - 1) / a.KI;
    cudaError_t e;
    // 1. folded-weight gradient
    {
        const int gy_ = (L.O + KB_BN - 1) / KB_BN;
        long long split = (148LL * 2 + (long long)ngroups * gy_ - 1) / ((long long)ngroups * gy_);
        split = std::max(1LL, std::min(split, ntile));
        const size_t smem = (size_t)(KB_KCMAX * KB_BM + KB_BM * KB_BN) * 4;
        auto k = kl_bwd_dw_kernel<BR>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e) return (int)e;
        if (!a.accumulate) {
            e = cudaMemsetAsync(a.ws, 0, ((size_t)L.I * L.F * L.O + L.O) * 4, st);
            if (e) return (int)e;
        }
        int rc_tc = FETODE_EUNSUPPORTED;
        if (a.L.tc_ok && a.B >= kl_tc_min_rows())               // wide layer: tcgen05 weight gradient (kanlinear_tc_bwd.cu)
            rc_tc = kl_tc_dw(&a.P, a.L, a.blob, a.x, a.prev, a.prev_is_zero, a.gy, a.ws, a.B, a.accumulate, st);
        if (rc_tc != FETODE_OK && rc_tc != FETODE_EUNSUPPORTED) return rc_tc;
        if (rc_tc == FETODE_EUNSUPPORTED)
            k<<<dim3(ngroups, gy_, (unsigned)split), KB_THREADS, smem, st>>>(a);
        const long long by = std::max(1LL, std::min<long long>((a.B + 63) / 64, 64));
        kl_colsum_kernel<<<dim3((L.O + 31) / 32, (unsigned)by), 256, 0, st>>>(a.gy, a.ws + (size_t)L.I * L.F * L.O, a.B, L.O);
    }
    // 2. input gradient + branch-basis parameter gradients
    const bool need_bp = a.g.bp[0] || a.g.bp[1];
    if (a.gx || (BR != FETODE_BRANCH_NONE && need_bp)) {
        long long gy_ = (148LL * 2 + ntile - 1) / ntile;
        gy_ = std::max(1LL, std::min<long long>(gy_, ngroups));
        if (ntile > 0x7fffffffLL) return FETODE_EUNSUPPORTED;
        const size_t smem = (size_t)(KB_OC * KB_GST + KB_KCMAX * (KB_OC + 1) + KB_KCMAX * KB_BM) * 4;
        int rc_tc = FETODE_EUNSUPPORTED;
        if (a.L.tc_ok && a.B >= kl_tc_min_rows())              // wide layer: tcgen05 input gradient (kanlinear_tc_dx.cu)
            rc_tc = kl_tc_dx(&a.P, a.L, a.blob, a.x, a.prev, a.prev_is_zero, a.gy, a.gx, a.g.bp[0], a.g.bp[1], a.B, st);
        if (rc_tc != FETODE_OK && rc_tc != FETODE_EUNSUPPORTED) return rc_tc;
        if (rc_tc == FETODE_EUNSUPPORTED) {
            auto k = kl_bwd_dx_kernel<BR>;
            e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e) return (int)e;
            k<<<dim3((unsigned)ntile, (unsigned)gy_), KB_THREADS, smem, st>>>(a);
        }
    }
    // 3. unfold (deferred by the caller when the same layer is applied many times per step)
    if (!a.accumulate) kl_unfold_run(a, st);
    return FETODE_LAUNCH_RC();
}

// chain rule from a folded-weight gradient `ws` ([I*F, O] then gsum [O]) to the reference's parameters; shared with
// the fused small-width reverse sweep (smallnet_bwd.cu)
int kl_unfold_launch(const fetode_kanlayer_t* P, const fetode_kanlayer_grads_t* grads, float* ws, cudaStream_t st) {
    int rc = kl_check(P);
    if (rc) return rc;
    KLBwdArgs a{};
    a.P = *P;
    if (grads) a.g = *grads;
    a.ws = ws;
    a.L = kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis);
    kl_unfold_run(a, st);
    return FETODE_LAUNCH_RC();
}

}  // namespace fetode

using namespace fetode;

extern "C" int64_t fetode_kanlinear_bwd_ws_floats(const fetode_kanlayer_t* P) {
    if (kl_check(P) != FETODE_OK) return -1;
    const KLLayout L = kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis);
    if (L.F > KB_KCMAX) return -1;
    return ((int64_t)L.I * L.F * L.O + L.O + 3) / 4 * 4;
}

extern "C" int fetode_kanlinear_bwd(const fetode_kanlayer_t* P, const float* wcat, const float* x,
                                    const float* prev, int prev_is_zero, const float* gy, float* gx,
                                    const fetode_kanlayer_grads_t* grads, float* ws, int64_t B, void* stream) {
    int rc = kl_check(P);
    if (rc) return rc;
    FETODE_CHECK_ARG(wcat && x && gy && ws && B >= 0);
    if (B == 0) return FETODE_OK;
    if (P->branch == FETODE_BRANCH_FERRO2D && !prev_is_zero) FETODE_CHECK_ARG(prev != nullptr);
    KLBwdArgs a{};
    a.P = *P;
    if (grads) a.g = *grads;
    a.blob = wcat; a.x = x; a.prev = prev; a.gy = gy; a.gx = gx; a.ws = ws; a.B = B;
    a.L = kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis);
    FETODE_CHECK_SUPPORTED(a.L.F <= KB_KCMAX);
    a.KI = std::max(1, std::min(4, KB_KCMAX / a.L.F));
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero;
    cudaStream_t st = (cudaStream_t)stream;
    switch (P->branch) {
        case FETODE_BRANCH_NONE: return kl_bwd_launch<FETODE_BRANCH_NONE>(a, st);
        case FETODE_BRANCH_LOGISTIC: return kl_bwd_launch<FETODE_BRANCH_LOGISTIC>(a, st);
        case FETODE_BRANCH_FERRO2D: return kl_bwd_launch<FETODE_BRANCH_FERRO2D>(a, st);
    }
    return FETODE_EUNSUPPORTED;
}

// Deferred-unfold pair: a layer applied many times per optimiser step (every Runge-Kutta stage) accumulates its folded
// gradient over all the calls and is unfolded to the reference's parameters once.
extern "C" int fetode_kanlinear_bwd_acc(const fetode_kanlayer_t* P, const float* wcat, const float* x,
                                        const float* prev, int prev_is_zero, const float* gy, float* gx,
                                        float* gbp0, float* gbp1, float* ws_acc, int64_t B, void* stream) {
    int rc = kl_check(P);
    if (rc) return rc;
    FETODE_CHECK_ARG(wcat && x && gy && ws_acc && B >= 0);
    if (B == 0) return FETODE_OK;
    if (P->branch == FETODE_BRANCH_FERRO2D && !prev_is_zero) FETODE_CHECK_ARG(prev != nullptr);
    KLBwdArgs a{};
    a.P = *P;
    a.g.bp[0] = gbp0; a.g.bp[1] = gbp1;
    a.blob = wcat; a.x = x; a.prev = prev; a.gy = gy; a.gx = gx; a.ws = ws_acc; a.B = B;
    a.L = kl_layout(P->in_features, P->out_features, P->grid_size, P->num_basis);
    FETODE_CHECK_SUPPORTED(a.L.F <= KB_KCMAX);
    a.KI = std::max(1, std::min(4, KB_KCMAX / a.L.F));
    a.gl = P->gate_slope * FETODE_LOG2E;
    a.prev_is_zero = prev_is_zero;
    a.accumulate = 1;
    cudaStream_t st = (cudaStream_t)stream;
    switch (P->branch) {
        case FETODE_BRANCH_NONE: return kl_bwd_launch<FETODE_BRANCH_NONE>(a, st);
        case FETODE_BRANCH_LOGISTIC: return kl_bwd_launch<FETODE_BRANCH_LOGISTIC>(a, st);
        case FETODE_BRANCH_FERRO2D: return kl_bwd_launch<FETODE_BRANCH_FERRO2D>(a, st);
    }
    return FETODE_EUNSUPPORTED;
}

extern "C" int fetode_kanlinear_unfold(const fetode_kanlayer_t* P, float* ws_acc, const fetode_kanlayer_grads_t* grads,
                                       void* stream) {
    FETODE_CHECK_ARG(ws_acc && grads);
    return kl_unfold_launch(P, grads, ws_acc, (cudaStream_t)stream);
}

// ---- fetode_kanlinear_bwd with library-owned scratch (no allocation anywhere on the call path) ----------------------
// No counterpart in the reference: added for callers that cannot allocate (CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The folded-gradient workspace of
// fetode_kanlinear_bwd lives in ONE file-scope device array per GPU that loads the library (FETODE_STATIC_WS_FLOATS
// floats = 1 MiB, held until exit).  Every call shares it and it keeps the last call's contents: the launch clears the
// part it uses first (cudaMemsetAsync on the caller's stream, as fetode_kanlinear_bwd does for a caller's buffer), and
// calls that overlap on the GPU (different streams, or replays of two captured graphs at once) RACE on it.
static __device__ __align__(256) float g_kl_static_ws[FETODE_STATIC_WS_FLOATS];

static int kl_static_ws_address(float** out) {
    // In host code `g_kl_static_ws` is a host-side stand-in, not a device address: ask the runtime, once per device,
    // with that device current.
    constexpr int MAXDEV = 64;
    static float* addr[MAXDEV] = {};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAXDEV) return FETODE_EUNSUPPORTED;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_kl_static_ws);
        if (e != cudaSuccess) return (int)e;
        if (!p) return FETODE_EINVAL;
        addr[dev] = p;
    }
    *out = addr[dev];
    return FETODE_OK;
}

extern "C" int64_t fetode_kanlinear_bwd_static_max_floats(void) { return FETODE_STATIC_WS_FLOATS; }

extern "C" int fetode_kanlinear_bwd_static(const fetode_kanlayer_t* P, const float* wcat, const float* x,
                                           const float* prev, int prev_is_zero, const float* gy, float* gx,
                                           const fetode_kanlayer_grads_t* grads, int64_t B, void* stream) {
    const int64_t need = fetode_kanlinear_bwd_ws_floats(P);
    if (need < 0) return kl_check(P) != FETODE_OK ? kl_check(P) : FETODE_EUNSUPPORTED;
    if (need > FETODE_STATIC_WS_FLOATS) return FETODE_ETOOLARGE;     // layer needs more scratch than the library holds
    float* ws = nullptr;
    int rc = kl_static_ws_address(&ws);
    if (rc) return rc;
    return fetode_kanlinear_bwd(P, wcat, x, prev, prev_is_zero, gy, gx, grads, ws, B, stream);
}
