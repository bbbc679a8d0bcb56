/*
 * fetode.h -- C ABI of libfetode.so: the B200 (sm_100a) implementation of the KAN-FET
 * neural-ODE right-hand side.
 *
 * This is the drop-in boundary.  The reference (sallywang147/FET-ODE @ 704a0df4) has no FFI of
 * its own: the path sits behind torch.nn.Module.forward calls and the torchdiffeq-style
 * odeint(func, y0, t) call.  Each entry point below therefore names the reference *method*
 * (file:line, relative to the reference checkout) whose arithmetic it replaces; INTEGRATION.md
 * shows the ctypes stub a maintainer adds on the reference side.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to contiguous row-major fp32 unless marked "host";
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - every function returns 0 on success, a negative FETODE_E* code on a usage error, or a
 *     positive cudaError_t; fetode_strerror() renders all three;
 *   - nothing allocates device memory: scratch comes from the caller (sizes from the
 *     *_floats / *_bytes queries); launches are asynchronous on `stream`;
 *   - re-entrant, no global state.
 */
#ifndef FETODE_H_
#define FETODE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FETODE_VERSION 100
#define FETODE_API __attribute__((visibility("default")))

/* ---- error codes ------------------------------------------------------------------- */
#define FETODE_OK            0
#define FETODE_EINVAL       (-1)   /* bad argument (null pointer, non-positive size, ...)  */
#define FETODE_EUNSUPPORTED (-2)   /* shape / option outside what the kernels implement    */
#define FETODE_ENOGPU       (-3)   /* no sm_100 device                                     */
#define FETODE_ESHAPE       (-4)   /* shape mismatch (reference raises AssertionError)     */
#define FETODE_ETOOLARGE    (-5)   /* needs more scratch than the library's static array holds (fetode_kanlinear_bwd_static) */

FETODE_API const char* fetode_strerror(int rc);
FETODE_API int fetode_version(void);
/* host out-params; any may be NULL */
FETODE_API int fetode_device_info(int* sm_count, int* cc_major, int* cc_minor, int* clock_khz);

/* ---- branch (per-input basis in the "logistic" slot of KANLinear) ------------------ */
#define FETODE_BRANCH_NONE     0   /* enable_logistic_basis=False                           */
#define FETODE_BRANCH_LOGISTIC 1   /* LogisticBasis              efficientkan.py:7-24       */
#define FETODE_BRANCH_FERRO2D  2   /* TwoDimensionFerroelectricBasis  ferro_class.py:526-596 (KANFET) */

/*
 * One KANLinear layer (efficient_kan/efficientkan.py:27-182; MNIST copy
 * mnist_kuramoto_kan.py:25-142).  Field names follow the reference's parameter names.
 * bp[]: branch-basis parameters, each [in_features, num_basis]:
 *   LOGISTIC: bp[0]=a, bp[1]=b                          (efficientkan.py:17-18)
 *   FERRO2D : bp[0]=k, bp[1]=Ec, bp[2]=Ps, bp[3]=bias, bp[4]=coef   (ferro_class.py:537-541)
 */
typedef struct fetode_kanlayer {
    int32_t in_features, out_features;
    int32_t grid_size, spline_order;          /* spline_order must be 3                      */
    int32_t num_basis;                        /* 0 when branch == NONE                       */
    int32_t branch;                           /* FETODE_BRANCH_*                             */
    const float* grid;                        /* [I, G+2k+1]            efficientkan.py:55-61 */
    const float* base_weight;                 /* [O, I]                                :64   */
    const float* spline_weight;               /* [O, I, G+k]                           :65   */
    const float* spline_scaler;               /* [O, I] or NULL                        :67   */
    const float* branch_weight;               /* logistic_weight [O, I*nb] or NULL     :86   */
    const float* branch_scaler;               /* logistic_scaler [O] or NULL           :88   */
    const float* branch_bias;                 /* logistic_bias [O] or NULL  mnist_kuramoto_kan.py:78 */
    const float* bp[5];
    float scale_logistic;                     /* efficientkan.py:45,153                      */
    float gate_slope, alpha;                  /* ferro_class.py:527  (10.0, 0.8)             */
} fetode_kanlayer_t;

/* Gradient destinations, same shapes as the parameters; NULL = not wanted.  Kernels ACCUMULATE
 * (+=) into these buffers, so the caller zeroes them first (torch's .grad convention). */
typedef struct fetode_kanlayer_grads {
    float* base_weight;
    float* spline_weight;
    float* spline_scaler;
    float* branch_weight;
    float* branch_scaler;
    float* branch_bias;
    float* bp[5];
} fetode_kanlayer_grads_t;

/* ---- KANLinear, any width (CUDA-core path; the tcgen05 path is selected inside) ------
 * Replaces KANLinear.forward efficientkan.py:160-182 (b_splines :117-131,
 * scaled_spline_weight :145-147, scaled_logistic_weight :149-158, LogisticBasis.forward :20-24)
 * and its autograd backward.
 *
 * `wcat` is the folded weight matrix [I*F, O], F = 1 + (G+k) + nb, rows ordered
 * (i, [silu | spline s | branch j]); it is rebuilt by fetode_kanlinear_pack whenever the
 * parameters change.  `prev` is the hysteresis state of the FERRO2D branch: the previous input
 * [B, I]; with prev_is_zero != 0 it is taken as all-zero and not read (first call after
 * reset_state, ferro_class.py:544,594-596).  prev_out (NULL, or a buffer that does NOT alias prev)
 * receives x, the next call's prev (ferro_class.py:584-590).
 */
/* Kernel selection for the wide path: -1 = automatic (tcgen05 from 64 rows up; the default, or FETODE_TC=0|1 read once at
 * first use), 0 = CUDA-core kernels only, 1 = tcgen05 for any batch.  Process-wide; for tests and A/B timing. */
FETODE_API int fetode_set_tc_mode(int mode);
FETODE_API int64_t fetode_kanlinear_wcat_floats(const fetode_kanlayer_t* L);
FETODE_API int fetode_kanlinear_pack(const fetode_kanlayer_t* L, float* wcat, void* stream);
FETODE_API int fetode_kanlinear_fwd(const fetode_kanlayer_t* L, const float* wcat, const float* x,
                         const float* prev, int prev_is_zero, float* prev_out,
                         float* y, int64_t B, void* stream);
/* scratch for bwd: fetode_kanlinear_bwd_ws_floats(L) floats, need not be zeroed */
FETODE_API int64_t fetode_kanlinear_bwd_ws_floats(const fetode_kanlayer_t* L);
FETODE_API int fetode_kanlinear_bwd(const fetode_kanlayer_t* L, const float* wcat, const float* x,
                         const float* prev, int prev_is_zero, const float* gy,
                         float* gx /* [B,I] or NULL, overwritten */,
                         const fetode_kanlayer_grads_t* grads, float* ws,
                         int64_t B, void* stream);
/* fetode_kanlinear_bwd for callers that cannot allocate: NO COUNTERPART IN THE REFERENCE -- added
 * (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments and
 * results as fetode_kanlinear_bwd minus `ws`: the scratch is a static device array inside the library,
 * FETODE_STATIC_WS_FLOATS floats (1 MiB) per GPU that loads the library, held until exit.  Nothing on the call path
 * allocates (no cudaMalloc / cudaFree / allocating library routine).  Layers with
 * fetode_kanlinear_bwd_ws_floats(L) > FETODE_STATIC_WS_FLOATS are rejected with FETODE_ETOOLARGE.  Every call shares
 * the array, which keeps the last call's contents; the call clears what it uses on `stream` first.  Calls that OVERLAP
 * on the GPU (two streams, two graphs replayed at once) race on it: serialise them. */
#define FETODE_STATIC_WS_FLOATS 262144
FETODE_API int64_t fetode_kanlinear_bwd_static_max_floats(void);
FETODE_API int fetode_kanlinear_bwd_static(const fetode_kanlayer_t* L, const float* wcat, const float* x,
                         const float* prev, int prev_is_zero, const float* gy,
                         float* gx /* [B,I] or NULL, overwritten */,
                         const fetode_kanlayer_grads_t* grads, int64_t B, void* stream);
/* The same backward split in two for a layer that is applied many times per optimiser step (once per Runge-Kutta
 * stage: 32 times per step in train_ecg_kan_fet_nn_ode.py's RK4 x 8 loop).  _bwd_acc ADDS this call's folded-weight
 * gradient ([I*F, O] then the output-gradient column sums [O]) to ws_acc (fetode_kanlinear_bwd_ws_floats(L) floats,
 * zeroed by the caller before the first call) and the branch-basis parameter gradients to gbp0 / gbp1 ([I, nb] each,
 * or NULL); gx is overwritten.  _unfold then applies the chain rule from the accumulated folded gradient to the
 * reference's parameters (every non-NULL member of grads except bp[0], bp[1] is added to) -- once per step. */
FETODE_API int fetode_kanlinear_bwd_acc(const fetode_kanlayer_t* L, const float* wcat, const float* x,
                         const float* prev, int prev_is_zero, const float* gy, float* gx /* [B,I] or NULL */,
                         float* gbp0, float* gbp1, float* ws_acc, int64_t B, void* stream);
FETODE_API int fetode_kanlinear_unfold(const fetode_kanlayer_t* L, float* ws_acc, const fetode_kanlayer_grads_t* grads,
                         void* stream);

/* ---- normalise-and-bound prologue of the ECG right-hand side ------------------------------
 * Replaces  h = self.ln(h); h = self.h_bound * torch.tanh(h / self.h_bound)  (train_ecg_kan_fet_nn_ode.py:459-461;
 * nn.LayerNorm(latent_dim), eps 1e-5, affine) and its autograd.  x, y, gy, gx [B, D] row-major, D <= 256;
 * mean / rstd [B] are written by _fwd and read by _bwd; ggamma / gbeta [D] are ADDED to (caller zeroes), any of
 * gx / ggamma / gbeta may be NULL. */
FETODE_API int fetode_ln_tanh_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean,
                       float* rstd, int64_t B, int32_t D, float eps, float h_bound, void* stream);
FETODE_API int fetode_ln_tanh_bwd(const float* x, const float* gamma, const float* mean, const float* rstd,
                       const float* y, const float* gy, float* gx, float* ggamma, float* gbeta, int64_t B,
                       int32_t D, float h_bound, void* stream);

/* ---- dense head of the ECG right-hand side -----------------------------------------------
 * Replaces  dh = self.out(z)  (nn.Linear(hidden, latent_dim), train_ecg_kan_fet_nn_ode.py:446,475; also self.proj
 * :497,505) and its autograd, in fp32 FMA arithmetic like the reference.  x [B, in], weight [out, in] (torch layout),
 * bias [out] or NULL, y [B, out].  _bwd: gx [B, in] overwritten (or NULL); gweight [out, in] and gbias [out] must be
 * zero-filled by the caller and receive the sums over the batch (or NULL). */
FETODE_API int fetode_linear_fwd(const float* x, const float* weight, const float* bias, float* y, int64_t B,
                      int32_t in_features, int32_t out_features, void* stream);
FETODE_API int fetode_linear_bwd(const float* x, const float* weight, const float* gy, float* gx, float* gweight,
                      float* gbias, int64_t B, int32_t in_features, int32_t out_features, void* stream);
/* The tail of the ECG right-hand side as one product:  dh = gain * (silu(z) W^T + b)
 * Replaces  z = self.act(z); dh = self.out(z); dh = self.scale * alpha * dh  (train_ecg_kan_fet_nn_ode.py:471-480; act =
 * nn.SiLU, alpha = softplus(log_alpha)) and its autograd.  gain: DEVICE scalar (scale * alpha, formed by the caller:
 * its own autograd stays with the caller).  z [B, in], weight [out, in], bias [out] or NULL, y [B, out];
 * u [B, out] (or NULL) receives the unscaled product, which _bwd needs for the gain's gradient.
 * _bwd: gz [B, in] overwritten (or NULL); gweight [out, in], gbias [out], ggain [1] are ADDED to (caller zeroes; any may
 * be NULL; ggain needs u). */
FETODE_API int fetode_head_fwd(const float* z, const float* weight, const float* bias, const float* gain, float* u,
                       float* y, int64_t B, int32_t in_features, int32_t out_features, void* stream);
FETODE_API int fetode_head_bwd(const float* z, const float* weight, const float* gain, const float* u, const float* gy,
                       float* gz, float* gweight, float* gbias, float* ggain, int64_t B, int32_t in_features,
                       int32_t out_features, void* stream);

/* B-spline basis values alone: KANLinear.b_splines efficientkan.py:117-131 (used by update_grid :184-221 and
 * curve2coeff :133-143 on CUDA inputs).  grid [I, G+2k+1], x [B,I] -> out [B, I, G+k]; spline_order must be 3. */
FETODE_API int fetode_bspline_basis(const float* grid, const float* x, float* out, int64_t B, int32_t I,
                         int32_t grid_size, int32_t spline_order, void* stream);

/* ---- fused small-width KAN / KANFET network + Runge-Kutta loop -----------------------
 * Replaces, for a stack of KANLinear layers whose widths are all <= 32 (the Lotka-Volterra
 * KANFET([2,10,2]) of train_kanfet_node_predprey.py:146), the whole of
 *   KAN.forward efficientkan.py:274-279  inside  integrate_euler/rk2/rk4
 *   train_ecg_kan_fet_nn_ode.py:670-705 (= odeint_rk4 train_kan_fet_ett.py:51-83 with
 *   n_substeps folded into n_steps) and torchdiffeq's fixed-grid euler/midpoint/rk4(3/8).
 * One thread integrates one trajectory; n_steps RK steps run inside one launch.
 */
#define FETODE_TABLEAU_EULER    0
#define FETODE_TABLEAU_MIDPOINT 1   /* integrate_rk2 :681-691 / torchdiffeq "midpoint"       */
#define FETODE_TABLEAU_RK4      2   /* classic, integrate_rk4 :693-705                      */
#define FETODE_TABLEAU_RK4_38   3   /* torchdiffeq "rk4" (3/8 rule)                          */

#define FETODE_SMALLNET_MAX_LAYERS 2
typedef struct fetode_smallnet {
    int32_t n_layers;                          /* currently exactly 2                        */
    int32_t widths[FETODE_SMALLNET_MAX_LAYERS + 1];   /* [D, H, D]: in == out (ODE state)    */
    int32_t grid_size, spline_order, num_basis, branch;
} fetode_smallnet_t;

/* 0 if the fused kernel has an instantiation for this net, FETODE_EUNSUPPORTED otherwise */
FETODE_API int fetode_smallnet_supported(const fetode_smallnet_t* net);
FETODE_API int64_t fetode_smallnet_packed_floats(const fetode_smallnet_t* net);
/* fold the layers' parameters into the shared-memory image the kernels stage */
FETODE_API int fetode_smallnet_pack(const fetode_smallnet_t* net, const fetode_kanlayer_t* layers,
                         float* packed, void* stream);
/* hysteresis state per trajectory: [B, state_floats] = previous input of every layer */
FETODE_API int64_t fetode_smallnet_state_floats(const fetode_smallnet_t* net);
/*
 * y0 [B,D] -> y_out [B,D] after n_steps steps of size dt.  state (in/out, [B, state_floats]) is
 * ignored on input when state_is_zero != 0; may be NULL when branch != FERRO2D.
 * traj (or NULL): [n_steps/dump_every, B, D] receives y after every dump_every-th step.
 * ckpt (or NULL): [n_steps, D + state_floats, B] receives (y, state) at the START of every step
 * (what fetode_smallnet_rk_bwd replays from).
 */
FETODE_API int fetode_smallnet_rk_fwd(const fetode_smallnet_t* net, const float* packed, int tableau,
                           const float* y0, float* y_out, float* state, int state_is_zero,
                           int64_t B, double dt, int32_t n_steps,
                           float* traj, int32_t dump_every, float* ckpt, void* stream);
/*
 * Reverse sweep of fetode_smallnet_rk_fwd (what loss.backward() does in train_kanfet_node_predprey.py:254-256:
 * plain autograd through every solver stage; hysteresis snapshots are constants).  ckpt is the buffer the forward
 * call filled ([n_steps, D + state_floats, B]).  g_yout [B,D] (or NULL) is d(loss)/d(y_out); g_traj (or NULL)
 * [n_steps/dump_every, B, D] is d(loss)/d(traj).  g_y0 [B,D] is overwritten.  Parameter gradients are ADDED to
 * gfold (fetode_smallnet_gfold_floats floats, zeroed by the caller before the first launch of a backward pass;
 * several launches -- one per run of equal step sizes -- may accumulate into it) in folded form;
 * fetode_smallnet_unfold then applies the chain rule and ACCUMULATES into grads[2] (one entry per layer).
 */
FETODE_API int64_t fetode_smallnet_gfold_floats(const fetode_smallnet_t* net);
FETODE_API int fetode_smallnet_rk_bwd(const fetode_smallnet_t* net, const float* packed, int tableau,
                           const float* ckpt, const float* g_yout, const float* g_traj, int32_t dump_every,
                           float* g_y0, float* gfold, int64_t B, double dt, int32_t n_steps, void* stream);
FETODE_API int fetode_smallnet_unfold(const fetode_smallnet_t* net, const fetode_kanlayer_t* layers, float* gfold,
                           const fetode_kanlayer_grads_t* grads, void* stream);
/* one RHS evaluation y -> f(y) (KAN.forward), same state rules */
/* The same fused integration and its reverse sweep with the HIDDEN UNITS on the lanes (csrc/smallnet_wide.cu): a warp
 * holds three trajectories, ten lanes each.  Built for training (train_kanfet_node_predprey.py:246-257: the reference's
 * run is ONE trajectory): the dependency chain of a right-hand side is 3 feature evaluations instead of 12, and every
 * parameter-gradient column is accumulated privately by the lane that owns the hidden unit, reduced once per launch.
 * Supported: widths [2,10,2], grid 5, order 3, num_basis 10, branch LOGISTIC | FERRO2D; `packed` as fetode_smallnet_pack.
 * ckpt: fetode_smallnet_wide_ckpt_floats(net, B, n_steps) floats = [n_steps][5][B][12] (slot 0: hysteresis snapshots at
 * the step start, slot 1+s: input and hidden pre-activations of stage s); written by _wide_fwd (NULL: not kept), read by
 * _wide_bwd.  Every other argument as fetode_smallnet_rk_fwd / _rk_bwd; gfold is the same folded-gradient buffer. */
FETODE_API int fetode_smallnet_wide_supported(const fetode_smallnet_t* net);
FETODE_API int64_t fetode_smallnet_wide_ckpt_floats(const fetode_smallnet_t* net, int64_t B, int32_t n_steps);
FETODE_API int fetode_smallnet_wide_fwd(const fetode_smallnet_t* net, const float* packed, int tableau, const float* y0,
                         float* y_out, float* state, int state_is_zero, int64_t B, double dt, int32_t n_steps,
                         float* traj, int32_t dump_every, float* ckpt, void* stream);
FETODE_API int fetode_smallnet_wide_bwd(const fetode_smallnet_t* net, const float* packed, int tableau, const float* ckpt,
                         const float* g_yout, const float* g_traj, int32_t dump_every, float* g_y0, float* gfold,
                         int64_t B, double dt, int32_t n_steps, void* stream);
/* ONE attempted step of an explicit Runge-Kutta method with a caller-supplied tableau, as one launch: the building block
 * of the adaptive solver.  Replaces, for the recognised small KAN / KANFET right-hand side, what torchdiffeq's
 * _runge_kutta_step + _compute_error_ratio do per attempt at train_kanfet_node_predprey.py:252 (dopri5 is that script's
 * default method): six right-hand sides, the stage sums, the 5th-order solution and the error norm.
 *   k_0 is given (first_eval = 1: FSAL) or evaluated (first_eval = 0);
 *   stage s input  ys_s = y + dt sum_{r<s} a[s][r] k_r,  k_s = f(ys_s)  for first_eval <= s < n_stages (<= 7);
 *   y1 = ys_{n_stages-1} when y1_is_last_input (dopri5: the last row of the tableau is the solution), else y;
 *   ks [n_stages, B, D] receives every k_s (k_0 copied when given);  y1 [B, D] or NULL;
 *   *errsq += sum_{rows, d} ( dt sum_s e[s] k_s / (atol + rtol max(|y|, |y1|)) )^2   (caller zeroes; NULL: skipped);
 *   state advances with EVERY evaluated stage, also on attempts the controller then rejects (the reference's stateful
 *   modules do the same);  ckpt: fetode_smallnet_attempt_ckpt_floats floats or NULL.
 * _attempt_bwd is the adjoint of one ACCEPTED attempt: g_y1 [B,D] / g_ks [n_stages,B,D] (either may be NULL) are the
 * adjoints of the outputs; x_y1, x_klast, x_y are optional [B,D] addends to g_y1, g_ks[n_stages-1] and the returned
 * g_y (the carries of the next step, so that the host adds nothing between launches); g_y [B,D], g_k0 [B,D] (when
 * first_eval) are written, parameter adjoints are ADDED to gfold (fetode_smallnet_gfold_floats, fetode_smallnet_unfold).
 * Step sizes are constants of the differentiation, as in torchdiffeq's discretise-then-optimise backward.
 * Same shape support as fetode_smallnet_wide_*. */
typedef struct fetode_rk_attempt {
    int32_t n_stages, first_eval, y1_is_last_input, _pad;
    float a[7][7];
    float e[7];
    float dt, rtol, atol;
} fetode_rk_attempt_t;
FETODE_API int64_t fetode_smallnet_attempt_ckpt_floats(const fetode_smallnet_t* net, const fetode_rk_attempt_t* rk,
                                                       int64_t B);
FETODE_API int fetode_smallnet_attempt_fwd(const fetode_smallnet_t* net, const float* packed,
                         const fetode_rk_attempt_t* rk, const float* y, const float* k0, float* y1, float* ks,
                         float* state, int state_is_zero, int64_t B, float* ckpt, double* errsq, void* stream);
FETODE_API int fetode_smallnet_attempt_bwd(const fetode_smallnet_t* net, const float* packed,
                         const fetode_rk_attempt_t* rk, const float* ckpt, const float* g_y1, const float* g_ks,
                         const float* x_y1, const float* x_klast, const float* x_y, float* g_y, float* g_k0,
                         float* gfold, int64_t B, void* stream);
/* The WHOLE dopri5 solve on the device for B <= fetode_smallnet_dp5_max_rows() (12) trajectories -- the reference's
 * Lotka-Volterra run is one (train_kanfet_node_predprey.py:146-147, :252): torchdiffeq's _adaptive_step loop (attempt,
 * error ratio, accept / reject, _optimal_step_size with safety 0.9 / ifactor 10 / dfactor 0.2, _interp_evaluate at the
 * requested times) runs inside one CTA; the host launches once per `cap` accepted steps.
 *   y, fy [B,D]: solution and FSAL slope f(y) at ctrl[0], in/out;  times [T] (device doubles), out [T,B,D] (row 0 is the
 *   caller's);  ctrl [8] device doubles, in/out: tb, ta, dt, next output index (start at 1), accepted, rejected,
 *   status (0 done | 1 launch again | 2 dt underflow | 3 non-finite error ratio), accepted steps in this launch.
 *   Reverse sweep (optional, ckpt != NULL): ckpt [cap][7][B][12] (fetode_smallnet_dp5_ckpt_floats), dts [cap],
 *   rec_step [T] / rec_w [T][9]: which accepted step produced output i and with which weights of (ya, yb, k_0..k_6).
 * _solve_bwd walks one chunk backwards: g_out [T,B,D] adjoints of the outputs, *cursor (device int, start T-1) the highest
 * output not yet consumed, carry_y / carry_k [B,D] in/out (zero before the last chunk; after the first chunk they are
 * the adjoints of y(t0) -- without g_out[0] -- and of f(y0)), gfold as fetode_smallnet_rk_bwd. */
FETODE_API int32_t fetode_smallnet_dp5_max_rows(void);
FETODE_API int64_t fetode_smallnet_dp5_ckpt_floats(const fetode_smallnet_t* net, int64_t B, int32_t cap);
FETODE_API int fetode_smallnet_dp5_solve_fwd(const fetode_smallnet_t* net, const float* packed, float* y, float* fy,
                         float* state, int state_is_zero, int64_t B, const double* times, int32_t T, float* out,
                         double* ctrl, double rtol, double atol, float* ckpt, float* dts, int32_t cap,
                         int32_t* rec_step, float* rec_w, int32_t max_attempts, void* stream);
FETODE_API int fetode_smallnet_dp5_solve_bwd(const fetode_smallnet_t* net, const float* packed, const float* ckpt,
                         const float* dts, int32_t n_steps, int32_t step_base, const int32_t* rec_step,
                         const float* rec_w, const float* g_out, int32_t* cursor, float* carry_y, float* carry_k,
                         float* gfold, int64_t B, void* stream);
FETODE_API int fetode_smallnet_rhs(const fetode_smallnet_t* net, const float* packed, const float* y,
                        float* dy, float* state, int state_is_zero, int64_t B, void* stream);

/* ---- Runge-Kutta stage algebra for right-hand sides that are not fused --------------------
 * out = y + sum_{i<ns} coef[i] * ks[i]   (y may be NULL == 0; ks / coef are HOST arrays, ns <= 8).
 * Replaces the axpy chains of integrate_euler/rk2/rk4 train_ecg_kan_fet_nn_ode.py:670-705,
 * odeint_rk4 train_kan_fet_ett.py:51-83 and of torchdiffeq's fixed-grid / dopri5 steps.
 */
#define FETODE_RK_MAX_TERMS 8
FETODE_API int fetode_rk_combine(const float* y, const float* const* ks, const float* coef, int32_t ns,
                      float* out, int64_t n, void* stream);
/* *sumsq (device double) = sum_i (err_i / (atol + rtol * max(|y0_i|, |y1_i|)))^2 : the numerator of
 * torchdiffeq's dopri5 accept/reject RMS norm (one global norm over the whole state tensor). */
FETODE_API int fetode_rk_error_sumsq(const float* err, const float* y0, const float* y1, float atol, float rtol,
                          int64_t n, double* sumsq, void* stream);

/* ---- stand-alone ferroelectric neurons ------------------------------------------------
 * FULL    : FerroelectricBasis.forward ferro_class.py:368-420 (update_branch=0) and
 *           NoisyFerroelectricBasis.forward :466-519 with noise_std=0 (update_branch=1);
 *           params [I,O,K]; out [B,O].  prev [B,I]; br [B,I,O,K] or NULL (== all ones).
 * TWO_D   : TwoDimensionFerroelectricBasis.forward :547-592; params [I,K] (O must be 1);
 *           out [B,I,K] = basis*coef.
 * HARDGATE: hysteretic LogisticBasis.forward train_ecg_kan_fet_nn_ode.py:102-134; params [I,K];
 *           out [B,I,K]; prev is ONE row [I] shared by the whole batch; gate (or NULL) [B,I,K]
 *           uint8 receives the branch actually taken; prev_out receives the LAST row of x.
 */
#define FETODE_FERRO_FULL     0
#define FETODE_FERRO_TWO_D    1
#define FETODE_FERRO_HARDGATE 2
#define FETODE_FERRO_LOGISTIC 3   /* LogisticBasis.forward efficientkan.py:20-24 stand-alone: k:=a, Ec:=b [I,K];
                                     out [B,I,K] = 2/(1+exp(-a (x-b))); Ps, bias, coef unused (NULL)        */

typedef struct fetode_ferro {
    int32_t variant, in_dim, out_dim, num_basis;
    const float *k, *Ec, *Ps, *bias, *coef;
    float gate_slope, alpha;                   /* hard gate: gate_slope=5, alpha=breaking point 0.5 */
    int32_t apply_sigmoid;                     /* HARDGATE only: 1 = KANFeatureMixer's sigmoid :419 fused in;
                                                  2 = sigmoid then nan_to_num(nan=0) (MLPKANODEFunc.forward :464-465) */
    /* FULL only: the glue of KANFetODEFunc.forward (train_ecg.py:997-1013) folded into the neuron kernel */
    float pre_bound;                           /* > 0: input x_eff = pre_bound * tanh(x / pre_bound)  :1000; prev_out <- x_eff */
    int32_t post_op;                           /* 1: tanh(out) :1003 ; 2: nan_to_num(0, +-1e3) then clamp(+-50) :1008-1011 */
} fetode_ferro_t;

typedef struct fetode_ferro_grads { float *k, *Ec, *Ps, *bias, *coef; } fetode_ferro_grads_t;

FETODE_API int fetode_ferro_fwd(const fetode_ferro_t* F, const float* x, const float* prev, int prev_is_zero,
                     const float* br, float* out, float* basis_out /* or NULL */,
                     float* br_out /* or NULL */, uint8_t* gate_out /* or NULL */,
                     float* prev_out /* or NULL: FULL/TWO_D [B,I] <- x ; HARDGATE [I] <- last row of x */,
                     int64_t B, void* stream);
FETODE_API int fetode_ferro_bwd(const fetode_ferro_t* F, const float* x, const float* prev, int prev_is_zero,
                     const float* br, const float* gout, float* gx /* [B,I] overwritten, or NULL */,
                     const fetode_ferro_grads_t* grads /* accumulated */, int64_t B, void* stream);
/* prev_is_zero (both directions): 0 = read prev, 1 = prev is all zero, 2 (FULL) = prev IS the effective input (the
 * reference's first call with a batch: buffers re-created as prev = x, ferro_class.py:373-378).
 * _bwd_fused: as fetode_ferro_bwd for a forward that ran with post_op != 0; yout [B,O] is that forward's output. */
FETODE_API int fetode_ferro_bwd_fused(const fetode_ferro_t* F, const float* x, const float* prev, int prev_is_zero,
                     const float* br, const float* gout, const float* yout, float* gx,
                     const fetode_ferro_grads_t* grads, int64_t B, void* stream);

/* ---- OriginalBatchedFerroelectricBasis.forward (ferro_class.py:49-99) + autograd ------------------------------------
 * Sequential scan over the batch: sample b uses prev_x = x[b-1] (prev0 [I] for b = 0, NULL = zeros) and the branch
 * momentum left by sample b-1 (br0 [I,O,K] for b = 0, NULL = +1), which is ALSO the value used in the shifted field
 * (:73-77) and stays differentiable across samples.  br_hist [B,I,O,K] (branch_sign before each sample) is what the
 * backward replays; br_final [I,O,K] is the state after the last sample.  F: variant FULL.  I*K <= 1024. */
FETODE_API int fetode_ferro_seq_fwd(const fetode_ferro_t* F, const float* x, const float* prev0, const float* br0,
                     float* out /* [B,O] */, float* basis_out /* [B,I,O,K] or NULL */, float* br_hist /* or NULL */,
                     float* br_final /* or NULL */, int64_t B, void* stream);
FETODE_API int fetode_ferro_seq_bwd(const fetode_ferro_t* F, const float* x, const float* prev0, const float* br_hist,
                     const float* gout, float* gx /* [B,I] overwritten, or NULL */,
                     const fetode_ferro_grads_t* grads /* accumulated */, int64_t B, void* stream);

/* ---- FerroelectricBasisConv2d / MemEfficient_FerroelectricBasisConv2d (ferro_class.py:601-779, :783-944) ----------
 * The same neuron applied to the kH x kW patches of an image, WITHOUT materialising F.unfold's (B, Cin*kH*kW, L)
 * matrix (:699-703) or the (B,Cout,Cin,K,kH,kW,L) broadcast (:712-722): the kernel gathers each patch element from the
 * image (zero padding) -- an implicit im2col around fetode_ferro_fwd's FULL variant.
 * F: variant FULL, in_dim = Cin*kH*kW ordered (channel, ky, kx), out_dim = Cout, parameters laid out [in_dim, Cout, K]
 *    (the reference's [Cout, Cin, K, kH, kW] permuted by the caller).
 * x, prev, gx: images [B, Cin, H, W]; out, gout: [B, Cout, Hout*Wout]; out_bias [Cout] or NULL (:773).
 * prev_mode: 2 = stateless call, prev == x (:724-727, the default); 0 = stateful, prev = the previous image; 1 = zeros.
 * br / br_out: [B*L, in_dim, Cout, K] branch state of the stateful plain class (NULL = +1 everywhere);
 * br_momentum != 0 stores branch_momentum (:763-767) rather than target_sign. */
typedef struct fetode_conv2d {
    int32_t Cin, H, W, kH, kW, stride, padding, Hout, Wout;
} fetode_conv2d_t;
FETODE_API int fetode_ferro_conv2d_fwd(const fetode_ferro_t* F, const fetode_conv2d_t* G, const float* x,
                     const float* prev, int prev_mode, const float* br, const float* out_bias, float* out,
                     float* basis_out /* [B*L, in_dim, Cout, K] or NULL */, float* br_out, int br_momentum,
                     int64_t B, void* stream);
FETODE_API int fetode_ferro_conv2d_bwd(const fetode_ferro_t* F, const fetode_conv2d_t* G, const float* x,
                     const float* prev, int prev_mode, const float* br, const float* gout,
                     float* gx /* image, overwritten, or NULL */, const fetode_ferro_grads_t* grads /* accumulated */,
                     float* g_out_bias /* [Cout] accumulated, or NULL */, int64_t B, void* stream);

/* ---- logistic-KAN recurrent encoder ------------------------------------------------------------------------------
 * FullyNonlinearKANCell.forward (train_kan_fet_ett.py:791-796; same cell train_ecg.py:203-217) applied T times
 * (KANRNNEncoder.forward :810-815; T = 1 is the bare cell): h_t = sigmoid([LogisticBasis_in(x_t) | LogisticBasis_h(h_{t-1})])[:, :H].
 * x [B,T,F]; h0 [B,H] or NULL (zeros); a_in,b_in [F,nb]; a_h,b_h [H,nb]; h_out [B,H] = h_T.
 * hist [T+1,B,H] (h_0..h_T) is needed by the backward only when H > F*nb (some kept column reads h_{t-1}); otherwise
 * h_T depends on x[:, T-1] alone and one step is evaluated.  bwd: g_x [B,T,F] must be zero-filled by the caller;
 * g_h0 (or NULL) overwritten; parameter gradients accumulated. */
FETODE_API int fetode_kanrnn_fwd(const float* x, const float* h0, const float* a_in, const float* b_in,
                     const float* a_h, const float* b_h, float* h_out, float* hist /* or NULL */, int64_t B,
                     int32_t T, int32_t F, int32_t H, int32_t nb, void* stream);
FETODE_API int fetode_kanrnn_bwd(const float* x, const float* a_in, const float* b_in, const float* a_h,
                     const float* b_h, const float* hist, const float* g_h, float* g_x /* or NULL */,
                     float* g_h0 /* or NULL */, float* g_a_in, float* g_b_in, float* g_a_h, float* g_b_h, int64_t B,
                     int32_t T, int32_t F, int32_t H, int32_t nb, void* stream);

/* ---- Kuramoto lattice -----------------------------------------------------------------
 * Kuramoto2D.forward mnist_kuramoto_kan.py:175-204: x [B,H*W] in [0,1] -> feat [B, 2*H*W]
 * = [cos(theta), sin(theta)] after `steps` Euler steps.  K: device scalar, omega: [H*W].
 * ckpt (or NULL) [steps, B, H*W] receives theta before every step (for the backward).
 */
FETODE_API int fetode_kuramoto_fwd(const float* x, const float* K, const float* omega, int32_t H, int32_t W,
                        int32_t steps, float dt, float* feat, float* ckpt, int64_t B, void* stream);
FETODE_API int fetode_kuramoto_bwd(const float* ckpt, const float* theta_final_feat, const float* K,
                        const float* omega, int32_t H, int32_t W, int32_t steps, float dt,
                        const float* gfeat, float* gx /* or NULL */, float* gK /* accum */,
                        float* gomega /* accum */, int64_t B, void* stream);

/* ---- fused optimiser step on one flat fp32 buffer ------------------------------
 * Replaces torch.optim.Adam.step (train_kanfet_node_predprey.py:171,257) and
 * clip_grad_norm_ + torch.optim.AdamW.step (train_ecg_kan_fet_nn_ode.py:1093,1110-1111) for parameters, gradients
 * and moments that live in flat buffers of n floats (the buffer the gradient all-reduce uses).
 * fetode_sumsq_f32: *sumsq (device double) = sum_i g_i^2.
 * fetode_adam_flat: one Adam (decoupled == 0: weight_decay is L2, added to g) or AdamW (decoupled != 0) step number
 *   `step` (1-based, bias correction as torch/optim/adam.py).  With sumsq != NULL the gradient is first scaled by
 *   min(1, max_norm / (sqrt(*sumsq) + 1e-6)) and written back (clip_grad_norm_ semantics) -- no host synchronisation. */
FETODE_API int fetode_sumsq_f32(const float* g, int64_t n, double* sumsq, void* stream);
FETODE_API int fetode_adam_flat(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                     float eps, float weight_decay, int decoupled, int64_t step, float max_norm,
                     const double* sumsq /* device, or NULL */, void* stream);
/* Same update with the step counter kept on the device, so that the launch can be captured in a CUDA graph and
 * replayed: state = 16 bytes {int64 step (incremented by this call); float lr / (1 - beta1^step); float
 * sqrt(1 - beta2^step)}, zero-initialised by the caller.  lr and the betas are baked into the captured launch. */
FETODE_API int fetode_adam_flat_graph(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1,
                           float beta2, float eps, float weight_decay, int decoupled, int64_t* state /* device */,
                           float max_norm, const double* sumsq /* device, or NULL */, void* stream);

/* ---- probes used by bench.py for the roofline denominators --------------------------- */
/* runs an FFMA-only kernel and returns achieved TFLOP/s in *tflops (host) */
FETODE_API int fetode_probe_fp32_peak(double* tflops, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FETODE_H_ */
