// smallnet.cu -- fused small-width KAN / KANFET network + Runge-Kutta loop (sm_100a).
//
// One thread integrates one trajectory; all n_steps RK steps (each 1/2/4 right-hand-side
// evaluations of the two-layer KAN) run inside one launch with the state in registers.  The
// network's parameters are folded once per parameter update (fetode_smallnet_pack) into a
// ~9 KB image that every CTA stages in shared memory; all weight reads are warp-uniform
// (broadcast) 128-bit LDS except the four spline coefficients selected by the knot interval.
//
// Reference arithmetic being replaced (file:line relative to the reference checkout):
//   KAN.forward / KANLinear.forward           efficient_kan/efficientkan.py:160-182, 274-279
//   b_splines (Cox-de Boor)                   efficient_kan/efficientkan.py:117-131
//   LogisticBasis.forward                     efficient_kan/efficientkan.py:20-24
//   TwoDimensionFerroelectricBasis.forward    ferro_class.py:547-592      (KANFET branch)
//   integrate_euler / rk2 / rk4               train_ecg_kan_fet_nn_ode.py:670-705
//   odeint_rk4                                train_kan_fet_ett.py:51-83
//   torchdiffeq fixed-grid euler/midpoint/rk4 (call sites train_kanfet_node_predprey.py:252)
//
// Folding (all exact in real arithmetic; fp32 rounding differs from the reference by O(1e-7)):
//   sigmoid(z) = 1/(1+2^(-z*log2e)),  tanh(z) = 1 - 2/(1+2^(2z*log2e)),  branch_sign == 1
//   (never written by TwoDimensionFerroelectricBasis) => momentum = 1 - 2(1-alpha)*to_lower.
//   With r = 1/(1+2^t3):  coef*(Ps*tanh+bias)*W = W*coef*(Ps+bias) + (-2*W*coef*Ps)*r ; the first
//   term is input independent and summed into a per-output bias at pack time.
#include <cstdlib>
#include "smallnet.cuh"
#include "smallnet_v2.cuh"

namespace fetode {

// ------------------------------------------------------------------------------------------
// pack: fold the two layers' parameters into the image
// ------------------------------------------------------------------------------------------
__global__ void smallnet_pack_kernel(fetode_kanlayer_t P1, fetode_kanlayer_t P2, SNLayout L, float* img) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const int D = L.D, H = L.H, NB = L.NB, S = L.S;
    for (int i = tid; i < L.total_ext; i += nt) img[i] = 0.f;
    __syncthreads();
    if (tid == 0) {
        img[L.hdr + 0] = P1.gate_slope * FETODE_LOG2E;
        img[L.hdr + 1] = P2.gate_slope * FETODE_LOG2E;
        img[L.hdr + 2] = 2.0f * (1.0f - P1.alpha);
        img[L.hdr + 3] = 2.0f * (1.0f - P2.alpha);
    }
    if (NB > 0) {
        for (int idx = tid; idx < D * NB; idx += nt) {
            img[L.raw1 + idx] = P1.bp[0][idx];
            img[L.raw1 + D * NB + idx] = P1.bp[1][idx];
        }
        for (int idx = tid; idx < H * NB; idx += nt) {
            img[L.raw2 + idx] = P2.bp[0][idx];
            img[L.raw2 + H * NB + idx] = P2.bp[1][idx];
        }
    }
    for (int d = tid; d < D; d += nt)
        pack_knots_row(P1.grid + (size_t)d * L.NK, L.NK, L.NKP, img + L.l1_kn + d * L.NKP,
                       img + L.l1_rd1 + d * L.NKP, img + L.l1_rd2 + d * L.NKP, img + L.l1_rd3 + d * L.NKP,
                       img + L.l1_gi + d * 4);
    for (int idx = tid; idx < D * NB; idx += nt)
    {
        float4 q = branch_params(P1, idx);
        q.x = bq_c1(P1, q.x);                               // FERRO2D: 2^(c1) (see smallnet.cuh, bq_c1)
        reinterpret_cast<float4*>(img + L.l1_bp)[idx] = q;
    }
    for (int idx = tid; idx < D * (NB / 2); idx += nt) {                      // pair (p, p + NB/2) of input d
        const int d = idx / (NB / 2), p = idx % (NB / 2);
        const float4 lo = branch_params(P1, d * NB + p), hi = branch_params(P1, d * NB + p + NB / 2);
        float4* q = reinterpret_cast<float4*>(img + L.l1_bq) + d * NB + 2 * p;
        q[0] = make_float4(bq_c1(P1, lo.x), bq_c1(P1, hi.x), lo.y, hi.y);
        q[1] = make_float4(lo.z, hi.z, lo.w, hi.w);
    }
    for (int o = tid; o < D; o += nt) {
        double b = P2.branch_bias ? (double)P2.branch_bias[o] : 0.0;
        for (int h = 0; h < H; ++h)
            for (int j = 0; j < NB; ++j) { double wr, wc; branch_fold(P2, o, h, j, wr, wc); b += wc; }
        img[L.b2 + o] = (float)b;
    }
    for (int h = tid; h < H; h += nt) {
        float* rec = img + L.rec0 + h * L.REC;
        double b = P1.branch_bias ? (double)P1.branch_bias[h] : 0.0;
        for (int d = 0; d < D; ++d) {
            float* wu = rec + L.r_wu1 + d * L.FU;
            wu[0] = P1.base_weight[h * D + d];
            for (int j = 0; j < NB; ++j) {
                double wr, wc;
                branch_fold(P1, h, d, j, wr, wc);
                wu[1 + j] = (float)wr;
                b += wc;
            }
            float* wv = rec + L.r_wv1 + d * L.FU;
            for (int p = 0; p < NB / 2; ++p) { wv[2 * p] = wu[1 + p]; wv[2 * p + 1] = wu[1 + p + NB / 2]; }
            wv[2 * (NB / 2)] = wu[0];
            float* ws = rec + L.r_ws1 + d * L.SPP;
            const float sc = P1.spline_scaler ? P1.spline_scaler[h * D + d] : 1.0f;
            for (int s = 0; s < S; ++s) ws[s + 3] = P1.spline_weight[(h * D + d) * S + s] * sc;
        }
        rec[L.r_b1] = (float)b;
        pack_knots_row(P2.grid + (size_t)h * L.NK, L.NK, L.NKP, rec + L.r_kn, rec + L.r_rd1, rec + L.r_rd2,
                       rec + L.r_rd3, rec + L.r_gi);
        for (int j = 0; j < NB; ++j) {
            float4 q = branch_params(P2, h * NB + j);
            q.x = bq_c1(P2, q.x);
            reinterpret_cast<float4*>(rec + L.r_bp)[j] = q;
        }
        for (int p = 0; p < NB / 2; ++p) {
            const float4 lo = branch_params(P2, h * NB + p), hi = branch_params(P2, h * NB + p + NB / 2);
            float4* q = reinterpret_cast<float4*>(rec + L.r_bq) + 2 * p;
            q[0] = make_float4(bq_c1(P2, lo.x), bq_c1(P2, hi.x), lo.y, hi.y);
            q[1] = make_float4(lo.z, hi.z, lo.w, hi.w);
        }
        for (int o = 0; o < D; ++o) {
            float* wu = rec + L.r_wu2 + o;
            wu[0] = P2.base_weight[o * H + h];
            for (int j = 0; j < NB; ++j) { double wr, wc; branch_fold(P2, o, h, j, wr, wc); wu[(1 + j) * L.DV] = (float)wr; }
            float* ws = rec + L.r_ws2 + o;
            const float sc = P2.spline_scaler ? P2.spline_scaler[o * H + h] : 1.0f;
            for (int s = 0; s < S; ++s) ws[(s + 3) * L.DV] = P2.spline_weight[(o * H + h) * S + s] * sc;
        }
    }
}


template <int D, int NB, int G, int BR>
__global__ void __launch_bounds__(SN_THREADS, 8) smallnet_rk_kernel(const SNArgs a) {
    extern __shared__ float4 sn_smem4[];
    float* sm = reinterpret_cast<float*>(sn_smem4);
    constexpr SNLayout L0 = sn_layout(D, 0, NB, G);
    const int H = a.H;
    const int total = L0.total + H * L0.REC;
    const int tid = threadIdx.x;
    for (int i = tid; i < total / 4; i += SN_THREADS)
        sn_smem4[i] = reinterpret_cast<const float4*>(a.packed)[i];
    float* prev2 = sm + total + tid;          // [H][SN_THREADS] column of this thread
    __syncthreads();

    const long long b = (long long)blockIdx.x * SN_THREADS + tid;
    if (b >= a.B) return;
    const int SF = (BR == FETODE_BRANCH_FERRO2D) ? D + H : 0;

    float y[D], prev1[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { y[d] = a.y0[b * D + d]; prev1[d] = 0.f; }
    if constexpr (BR == FETODE_BRANCH_FERRO2D) {
        if (!a.state_is_zero) {
#pragma unroll
            for (int d = 0; d < D; ++d) prev1[d] = a.state[b * SF + d];
#pragma unroll 1
            for (int h = 0; h < H; ++h) prev2[h * SN_THREADS] = a.state[b * SF + D + h];
        } else {
#pragma unroll 1
            for (int h = 0; h < H; ++h) prev2[h * SN_THREADS] = 0.f;
        }
    }

    const int tab = a.tab;
    if (tab == TAB_RHS_ONLY) {
        float out[D];
        sn_rhs<D, NB, G, BR>(sm, H, y, prev1, prev2, out);
#pragma unroll
        for (int d = 0; d < D; ++d) y[d] = out[d];
    } else {
        const int nst = tab == FETODE_TABLEAU_EULER ? 1 : (tab == FETODE_TABLEAU_MIDPOINT ? 2 : 4);
        const int NC = D + SF;
#pragma unroll 1
        for (int step = 0; step < a.n_steps; ++step) {
            if (a.ckpt) {
                float* c = a.ckpt + ((size_t)step * NC) * a.B + b;
#pragma unroll
                for (int d = 0; d < D; ++d) c[(size_t)d * a.B] = y[d];
                if constexpr (BR == FETODE_BRANCH_FERRO2D) {
#pragma unroll
                    for (int d = 0; d < D; ++d) c[(size_t)(D + d) * a.B] = prev1[d];
#pragma unroll 1
                    for (int h = 0; h < H; ++h) c[(size_t)(2 * D + h) * a.B] = prev2[h * SN_THREADS];
                }
            }
            float k1[D], k2[D], acc[D], ys[D], kn[D];
#pragma unroll
            for (int d = 0; d < D; ++d) { k1[d] = k2[d] = acc[d] = 0.f; ys[d] = y[d]; }
#pragma unroll 1
            for (int s = 0; s < nst; ++s) {
                sn_rhs<D, NB, G, BR>(sm, H, ys, prev1, prev2, kn);
                if (tab == FETODE_TABLEAU_RK4) {
                    // (k1 + 2*k2 + 2*k3 + k4), stage inputs h + c*k_prev   train_ecg_kan_fet_nn_ode.py:698-703
                    const float c = (s == 2) ? a.dt : a.hdt;
                    const float w = (s == 0 || s == 3) ? 1.0f : 2.0f;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        acc[d] = (s == 0) ? kn[d] : fmaf(w, kn[d], acc[d]);
                        ys[d] = fmaf(c, kn[d], y[d]);
                    }
                } else if (tab == FETODE_TABLEAU_RK4_38) {
                    // torchdiffeq rk4_alt_step_func (3/8 rule); acc carries k3 after stage 2
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        if (s == 0) { k1[d] = kn[d]; ys[d] = y[d] + a.dt * kn[d] * (1.0f / 3.0f); }
                        else if (s == 1) { k2[d] = kn[d]; ys[d] = y[d] + a.dt * (kn[d] - k1[d] * (1.0f / 3.0f)); }
                        else if (s == 2) { acc[d] = kn[d]; ys[d] = y[d] + a.dt * (k1[d] - k2[d] + kn[d]); }
                        else acc[d] = (k1[d] + 3.0f * (k2[d] + acc[d]) + kn[d]) * a.dt * 0.125f;
                    }
                } else {
                    // euler: h + dt*f(h) :676 ; midpoint: k2 = f(h + 0.5*dt*k1), h + dt*k2 :686-689
#pragma unroll
                    for (int d = 0; d < D; ++d) { ys[d] = fmaf(a.hdt, kn[d], y[d]); acc[d] = kn[d]; }
                }
            }
#pragma unroll
            for (int d = 0; d < D; ++d) {
                if (tab == FETODE_TABLEAU_RK4) y[d] = fmaf(a.dt6, acc[d], y[d]);
                else if (tab == FETODE_TABLEAU_RK4_38) y[d] = y[d] + acc[d];
                else y[d] = fmaf(a.dt, acc[d], y[d]);
            }
            if (a.traj && (step + 1) % a.dump_every == 0) {
                float* t = a.traj + ((size_t)((step + 1) / a.dump_every - 1) * a.B + b) * D;
#pragma unroll
                for (int d = 0; d < D; ++d) t[d] = y[d];
            }
        }
    }
#pragma unroll
    for (int d = 0; d < D; ++d) a.y_out[b * D + d] = y[d];
    if constexpr (BR == FETODE_BRANCH_FERRO2D) {
        if (a.state) {
#pragma unroll
            for (int d = 0; d < D; ++d) a.state[b * SF + d] = prev1[d];
#pragma unroll 1
            for (int h = 0; h < H; ++h) a.state[b * SF + D + h] = prev2[h * SN_THREADS];
        }
    }
}

// ------------------------------------------------------------------------------------------
// host dispatch
// ------------------------------------------------------------------------------------------
int sn_check(const fetode_smallnet_t* net) {
    if (!net) return FETODE_EINVAL;
    if (net->n_layers != 2) return FETODE_EUNSUPPORTED;
    const int D = net->widths[0], H = net->widths[1];
    if (net->widths[2] != D) return FETODE_EUNSUPPORTED;
    if (D < 1 || D > 4 || H < 1 || H > 64) return FETODE_EUNSUPPORTED;
    if (net->spline_order != 3 || net->grid_size != 5) return FETODE_EUNSUPPORTED;
    switch (net->branch) {
        case FETODE_BRANCH_NONE: if (net->num_basis != 0) return FETODE_EINVAL; break;
        case FETODE_BRANCH_LOGISTIC:
        case FETODE_BRANCH_FERRO2D:
            if (net->num_basis != 10) return FETODE_EUNSUPPORTED;
            break;
        default: return FETODE_EINVAL;
    }
    return FETODE_OK;
}

template <int D, int NB, int BR>
static int sn_launch_k(const SNArgs& a, cudaStream_t st) {
    constexpr SNLayout L0 = sn_layout(D, 0, NB, 5);
    const size_t smem = (size_t)(L0.total + a.H * L0.REC) * 4 + (size_t)a.H * SN_THREADS * 4;
    const unsigned grid = (unsigned)((a.B + SN_THREADS - 1) / SN_THREADS);
    auto kern = smallnet_rk_kernel<D, NB, 5, BR>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<grid, SN_THREADS, smem, st>>>(a);
    return FETODE_LAUNCH_RC();
}

template <int D>
static int sn_launch_d(const fetode_smallnet_t* net, const SNArgs& a, cudaStream_t st) {
    switch (net->branch) {
        case FETODE_BRANCH_NONE: return sn_launch_k<D, 0, FETODE_BRANCH_NONE>(a, st);
        case FETODE_BRANCH_LOGISTIC: return sn_launch_k<D, 10, FETODE_BRANCH_LOGISTIC>(a, st);
        case FETODE_BRANCH_FERRO2D: return sn_launch_k<D, 10, FETODE_BRANCH_FERRO2D>(a, st);
    }
    return FETODE_EUNSUPPORTED;
}

// FETODE_SN_V1=1 keeps the first-generation kernel for shapes the second one covers (A/B timing only; read once)
static bool sn_force_v1() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("FETODE_SN_V1"); v = (e && e[0] == '1') ? 1 : 0; }
    return v == 1;
}

bool snw_prefers(const fetode_smallnet_t* net, long long B, bool wants_v1_ckpt);      // smallnet_wide.cu
int snw_forward(const fetode_smallnet_t* net, const SNArgs& s, cudaStream_t st);

static int sn_launch(const fetode_smallnet_t* net, const SNArgs& a, cudaStream_t st) {
    // small batches: hidden units on the lanes (a dependency chain of 3 feature evaluations per right-hand side)
    if (a.tab != TAB_RHS_ONLY && !sn_force_v1() && snw_prefers(net, a.B, a.ckpt != nullptr)) return snw_forward(net, a, st);
    if (sn2_has(net) && !sn_force_v1()) {
        const int off = sn_layout(net->widths[0], net->widths[1], net->num_basis, net->grid_size).total_ext;
        return sn2_launch(net, a, a.packed + off, st);
    }
    switch (net->widths[0]) {
        case 1: return sn_launch_d<1>(net, a, st);
        case 2: return sn_launch_d<2>(net, a, st);
        case 3: return sn_launch_d<3>(net, a, st);
        case 4: return sn_launch_d<4>(net, a, st);
    }
    return FETODE_EUNSUPPORTED;
}

}  // namespace fetode

using namespace fetode;

extern "C" int fetode_smallnet_supported(const fetode_smallnet_t* net) { return sn_check(net); }

extern "C" int64_t fetode_smallnet_packed_floats(const fetode_smallnet_t* net) {
    if (sn_check(net) != FETODE_OK) return -1;
    return sn_layout(net->widths[0], net->widths[1], net->num_basis, net->grid_size).total_ext + sn2_image_floats(net);
}

extern "C" int64_t fetode_smallnet_state_floats(const fetode_smallnet_t* net) {
    if (sn_check(net) != FETODE_OK) return -1;
    return net->branch == FETODE_BRANCH_FERRO2D ? net->widths[0] + net->widths[1] : 0;
}

extern "C" int fetode_smallnet_pack(const fetode_smallnet_t* net, const fetode_kanlayer_t* layers,
                                    float* packed, void* stream) {
    int rc = sn_check(net);
    if (rc) return rc;
    FETODE_CHECK_ARG(layers && packed);
    const int D = net->widths[0], H = net->widths[1];
    for (int l = 0; l < 2; ++l) {
        const fetode_kanlayer_t& P = layers[l];
        FETODE_CHECK_SHAPE(P.in_features == (l ? H : D) && P.out_features == (l ? D : H));
        FETODE_CHECK_SHAPE(P.grid_size == net->grid_size && P.spline_order == 3);
        FETODE_CHECK_SHAPE(P.branch == net->branch && P.num_basis == net->num_basis);
        FETODE_CHECK_ARG(P.grid && P.base_weight && P.spline_weight);
        if (P.branch != FETODE_BRANCH_NONE) {
            FETODE_CHECK_ARG(P.branch_weight && P.bp[0] && P.bp[1]);
            if (P.branch == FETODE_BRANCH_FERRO2D) FETODE_CHECK_ARG(P.bp[2] && P.bp[3] && P.bp[4]);
        }
    }
    const SNLayout L = sn_layout(D, H, net->num_basis, net->grid_size);
    smallnet_pack_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(layers[0], layers[1], L, packed);
    rc = FETODE_LAUNCH_RC();
    if (rc) return rc;
    return sn2_pack(net, layers, packed + L.total_ext, (cudaStream_t)stream);    // second-generation image, appended
}

static int sn_run(const fetode_smallnet_t* net, const float* packed, int tab, const float* y0, float* y_out,
                  float* state, int state_is_zero, int64_t B, double dt, int32_t n_steps, float* traj,
                  int32_t dump_every, float* ckpt, void* stream) {
    int rc = sn_check(net);
    if (rc) return rc;
    FETODE_CHECK_ARG(packed && y0 && y_out && B >= 0);
    if (B == 0) return FETODE_OK;
    if (net->branch == FETODE_BRANCH_FERRO2D && !state_is_zero) FETODE_CHECK_ARG(state != nullptr);
    if (traj) FETODE_CHECK_ARG(dump_every >= 1);
    SNArgs a;
    a.packed = packed; a.y0 = y0; a.y_out = y_out; a.state = state; a.traj = traj; a.ckpt = ckpt;
    a.B = B; a.dt = (float)dt; a.hdt = (float)(0.5 * dt); a.dt6 = (float)(dt / 6.0);
    a.H = net->widths[1]; a.tab = tab;
    a.n_steps = n_steps; a.dump_every = dump_every > 0 ? dump_every : 1; a.state_is_zero = state_is_zero;
    return sn_launch(net, a, (cudaStream_t)stream);
}

extern "C" int fetode_smallnet_rk_fwd(const fetode_smallnet_t* net, const float* packed, int tableau,
                                      const float* y0, float* y_out, float* state, int state_is_zero,
                                      int64_t B, double dt, int32_t n_steps, float* traj,
                                      int32_t dump_every, float* ckpt, void* stream) {
    FETODE_CHECK_ARG(tableau >= FETODE_TABLEAU_EULER && tableau <= FETODE_TABLEAU_RK4_38 && n_steps >= 0);
    return sn_run(net, packed, tableau, y0, y_out, state, state_is_zero, B, dt, n_steps, traj, dump_every,
                  ckpt, stream);
}

extern "C" int fetode_smallnet_rhs(const fetode_smallnet_t* net, const float* packed, const float* y,
                                   float* dy, float* state, int state_is_zero, int64_t B, void* stream) {
    return sn_run(net, packed, TAB_RHS_ONLY, y, dy, state, state_is_zero, B, 0.0, 0, nullptr, 1, nullptr,
                  stream);
}
