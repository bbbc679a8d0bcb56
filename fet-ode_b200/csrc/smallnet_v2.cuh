// smallnet_v2.cuh -- second-generation fused small-width KAN / KANFET Runge-Kutta kernel: image layout and hooks.
//
// Same arithmetic as smallnet.cuh (the reference lines are cited there and in smallnet_v2.cu); what changes is
// the mapping onto the SM:
//   * a lane integrates 2 trajectories (TP = 1 packed pair; TP = 2 is kept in the template but loses to occupancy);
//     every FP32 operation is a packed fma/mul/add.f32x2 over the TWO TRAJECTORIES of a pair, so the weight operand is
//     the same in both halves: the image stores every weight twice (w, w) and one 128-bit shared-memory load feeds
//     two FFMA2 -- half the shared-memory instructions per trajectory of the first kernel;
//   * layer 1 is evaluated for all H hidden units before layer 2 (accumulators in registers, the hidden
//     pre-activations then go to a per-thread shared-memory column that doubles as the next call's previous input
//     of layer 2): the 2*NB layer-1 basis values no longer live across the hidden-unit loop -> no register spills;
//   * the spline term is a piecewise cubic: per (input, knot interval, output) the four monomial coefficients in
//     s = x - t_m are tabulated at pack time (exact in double), so an input costs one branch-free interval lookup, one
//     128-bit load and three FMA per output instead of the Cox-de Boor triangle (efficientkan.py:117-131); inputs
//     outside the knot span are masked, not branched around (divergent branches cost more than the arithmetic);
//   * the reciprocals of silu, 1 - is_moving_up and the NB crossed_neg factors share batched MUFU.RCP groups
//     (19 MUFU per input and trajectory instead of 23).
#pragma once
#include "smallnet.cuh"

namespace fetode {

// Image (floats).  "dup" = every value stored twice in a row (the two halves of a packed pair).
//   hdr[8]                  gl1 gl1 gl2 gl2 | 0 0 0 0                 (gate_slope * log2e per layer, dup)
//   b2[4]                   layer-2 output bias, dup (DO <= 2 -> 4 floats; DO <= 4 -> 8)
//   b1[2H]                  layer-1 bias, dup
//   per layer-1 input d:    gi[4] = {t_0, (NK-1)/(t_last-t_0), t_last, near-uniform flag} | kn[NKQ] knots (kn[1+j] = t_j) | bc[NB*8] branch constants
//                           (FERRO2D: [C_j dup] x NB | [k'_j dup, c2'_j dup] x NB | -2(1-alpha) dup ; LOGISTIC: - | [A_j dup, B_j dup] x NB | -) |
//                           w1[H][W1Q] ([base | r_0..r_{NB-1}] multipliers of output h, dup, padded to W1Q)
//   l1tab[D][NI][H][4]      cubic coefficients (c0..c3) of the layer-1 spline of (d -> h) on knot interval m
//   per hidden unit h:      gi[4] | kn[NKQ] | bc[NB*8] | w2[(1+NB)][DOP*2] ([base | r_j] multipliers of every output,
//                           dup) | tab[NI][DOP][4]
struct SN2Layout {
    int D, H, NB, NK, NI, NKQ, W1Q, DOP;
    int hdr, b2, b1, l1, L1S, o_gi, o_kn, o_bc, o_w1, l1tab, rec0, REC, r_gi, r_kn, r_bc, r_w2, r_tab;
    int total;
};

__host__ __device__ constexpr SN2Layout sn2_layout(int D, int H, int NB, int G) {
    SN2Layout L{};
    L.D = D; L.H = H; L.NB = NB;
    L.NK = G + 7;
    L.NI = L.NK - 1;
    L.NKQ = round_up(L.NK + 1, 4);      // kn[0] = t_0 pad, kn[1 + j] = t_j
    L.W1Q = round_up(2 * (1 + NB), 4);
    L.DOP = D <= 2 ? 2 : 4;
    int o = 0;
    L.hdr = o; o += 8;
    L.b2 = o; o += 2 * L.DOP;
    L.b1 = o; o += round_up(2 * H, 4);
    L.o_gi = 0; L.o_kn = 4; L.o_bc = 4 + L.NKQ; L.o_w1 = L.o_bc + NB * 8;
    L.L1S = round_up(L.o_w1 + H * L.W1Q, 4);
    L.l1 = o; o += D * L.L1S;
    L.l1tab = o; o += D * L.NI * H * 4;
    L.r_gi = 0; L.r_kn = 4; L.r_bc = 4 + L.NKQ; L.r_w2 = L.r_bc + NB * 8;
    L.r_tab = L.r_w2 + (1 + NB) * L.DOP * 2;
    L.REC = round_up(L.r_tab + L.NI * L.DOP * 4, 4);
    L.rec0 = o; o += H * L.REC;
    L.total = round_up(o, 4);
    return L;
}

// shapes the v2 kernel is instantiated for
__host__ __device__ inline bool sn2_has(const fetode_smallnet_t* net) {
    return net->n_layers == 2 && net->widths[0] == 2 && net->widths[1] == 10 && net->widths[2] == 2 &&
           net->grid_size == 5 && net->spline_order == 3 && net->num_basis == 10 &&
           (net->branch == FETODE_BRANCH_FERRO2D || net->branch == FETODE_BRANCH_LOGISTIC);
}

int64_t sn2_image_floats(const fetode_smallnet_t* net);                      // 0 when !sn2_has
int sn2_pack(const fetode_smallnet_t* net, const fetode_kanlayer_t* layers, float* img2, cudaStream_t st);
int sn2_launch(const fetode_smallnet_t* net, const SNArgs& a, const float* img2, cudaStream_t st);

}  // namespace fetode
