/*
 * mjrl_b200 -- C-ABI of the B200-native policy-update hot path.
 *
 * The reference (bennevans/mjrl) is pure Python and has no FFI; every entry
 * point below replaces a Python call site of the reference, cited as
 * file:line relative to the reference checkout.  INTEGRATION.md shows the
 * ctypes stub a maintainer adds on the reference side.
 *
 * Conventions
 *   - plain pointers and sizes only; every array pointer is a DEVICE pointer
 *     unless its name starts with `h_` (host);
 *   - every call is asynchronous on the caller-supplied CUDA stream
 *     (`stream` is a cudaStream_t passed as void*; NULL = legacy default);
 *   - return value: 0 = ok, <0 = argument/shape error (MJRL_E_*), >0 = the
 *     cudaError_t that the launch raised.  Nothing throws, nothing exits.
 *     mjrl_last_error() returns a thread-local description of the last failure.
 *   - flat timestep buffer: all paths concatenated; `path_offsets` is
 *     int32[P+1] with path p occupying [path_offsets[p], path_offsets[p+1]);
 *     `terminated` is uint8[P].
 *   - obs is fp32 [N, n] row-major, act fp32 [N, m]; rewards, returns,
 *     baseline values and advantages are fp64 [N].
 *   - the flat parameter vector is fp32 [d] in the reference's order
 *     fc0.weight (H1,n) | fc0.bias | fc1.weight (H2,H1) | fc1.bias |
 *     fc2.weight (m,H2) | fc2.bias | log_std   (gaussian_mlp.py:38,51-53),
 *     or fc0.weight (m,n) | fc0.bias | log_std for the affine policy
 *     (gaussian_linear.py:36,153); weights row-major (out,in).
 */
#ifndef MJRL_B200_H
#define MJRL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MJRL_ABI_VERSION 1

#define MJRL_E_ARG (-1)         /* null pointer / negative size            */
#define MJRL_E_SHAPE (-2)       /* shape not supported by the kernels      */
#define MJRL_E_STATE (-3)       /* call order / context state              */
#define MJRL_E_NOGPU (-4)       /* no usable CUDA device                   */

typedef struct mjrl_ctx mjrl_ctx;

/* ---- context ----------------------------------------------------------- */
int mjrl_abi_version(void);
const char* mjrl_last_error(void);
/* One context per process/GPU; owns scratch (per-CTA partial sums) only. */
int mjrl_ctx_create(mjrl_ctx** out, int device);
int mjrl_ctx_destroy(mjrl_ctx* ctx);
int mjrl_ctx_sm_count(const mjrl_ctx* ctx);

/* ---- (a) returns / advantages: mjrl/utils/process_samples.py ------------ */
/* discount_sum per path, :3-5,37-43.  returns[t] = rew[t] + gamma*returns[t+1]. */
int mjrl_returns_scan(mjrl_ctx* ctx, const double* rew, const int32_t* path_offsets,
                      int32_t P, int64_t N, double gamma, double* returns, void* stream);
/* GAE, :21-29.  td = r + gamma*b1[1:] - b1[:-1] with b1 = [b, terminated ? 0 : b[-1]];
 * adv = discount_sum(td, gamma*lam). */
int mjrl_gae_scan(mjrl_ctx* ctx, const double* rew, const double* baseline,
                  const int32_t* path_offsets, const uint8_t* terminated, int32_t P,
                  int64_t N, double gamma, double lam, double* adv, void* stream);
/* plain mode, :10-13.  adv = returns - baseline. */
int mjrl_adv_plain(mjrl_ctx* ctx, const double* returns, const double* baseline, int64_t N,
                   double* adv, void* stream);
/* moments[0..3] = {N, sum x, sum (x-mean)^2, mean} in fp64, deterministic.  Used for
 * whitening (npg_cg.py:91) and normalize=True (process_samples.py:14-19,30-35). */
int mjrl_moments(mjrl_ctx* ctx, const double* x, int64_t N, double* moments4, void* stream);
/* The same in two calls for a path-sharded batch: stage 0 writes the LOCAL {N, sum} (and the local mean);
 * the caller all-reduces moments4[0..1]; stage 1 sets moments4[3] = sum/N from them and writes the local
 * centred sum of squares to moments4[2], which the caller all-reduces.  Without the all-reduces the two
 * stages give mjrl_moments' bits. */
int mjrl_moments_stage(mjrl_ctx* ctx, const double* x, int64_t N, double* moments4, int32_t stage,
                       void* stream);
/* mjrl_moments for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels, same grid and therefore the same bits as mjrl_moments, but the per-CTA partial sums live in
 * a file-scope `static __device__` array of 4,096 doubles (32 KB, held by every GPU that loads the library
 * until the process exits) instead of the context's grow-on-demand scratch: the call path contains no
 * cudaMalloc / cudaFree / allocating library routine.  Limit: N <= 8,388,608 (4,096 partial sums x 2,048
 * elements per CTA slab); larger N returns MJRL_E_SHAPE with a message naming the limit.  The context's
 * device must be the current device (MJRL_E_STATE otherwise).  All calls on a device share the one array:
 * nothing in it needs clearing between calls (every pass overwrites the entries it reads), but two calls
 * whose kernels overlap on the GPU (different streams, or concurrent graph branches) RACE on it -- serialise
 * them on one stream. */
int mjrl_moments_noalloc(mjrl_ctx* ctx, const double* x, int64_t N, double* moments4, void* stream);
/* x <- (x - mean)/(std + eps) in place (fp64), std = sqrt(moments[2]/N). */
int mjrl_normalize_inplace(mjrl_ctx* ctx, double* x, int64_t N, const double* moments4,
                           double eps, void* stream);
/* w32[t] = (float)(scale * (adv[t]-mean)/(std+eps)): the fp32 whitened advantage the
 * policy passes consume (npg_cg.py:91 then batch_reinforce.py:38).  dapg.py:70 uses
 * scale = 1e-2/(std(white)+1e-8). */
int mjrl_whiten_f32(mjrl_ctx* ctx, const double* adv, int64_t N, const double* moments4,
                    double eps, double scale, float* w32, void* stream);
/* per-path reward sums -> stats4 = {mean, std, min, max} (npg_cg.py:97-102). */
int mjrl_path_return_stats(mjrl_ctx* ctx, const double* rew, const int32_t* path_offsets,
                           int32_t P, double* stats4, void* stream);
/* fp64 host-layout staging -> fp32 device arrays (obs/act cast of gaussian_mlp.py:103-104) */
int mjrl_cast_f64_to_f32(mjrl_ctx* ctx, const double* src, int64_t count, float* dst,
                         void* stream);

/* ---- (b) LinearBaseline: mjrl/baselines/linear_baseline.py -------------- */
/* features [clip(o,-10,10), al, al^2, al^3, 1], al = t_in_path/1000 (:10-18);
 * F = n+4.  pred[t] = features . coeffs (:46-49). */
int mjrl_linbase_predict(mjrl_ctx* ctx, const float* obs, const int32_t* path_offsets,
                         int32_t P, int64_t N, int32_t n, const double* coeffs,
                         double* pred, void* stream);
/* gram = [X^T X | X^T y] row-major F x (F+1), followed by y^T y at index F*(F+1)
 * (:22-23,33-34).  fp64 accumulate, deterministic two-stage reduction. */
int mjrl_linbase_gram(mjrl_ctx* ctx, const float* obs, const double* returns,
                      const int32_t* path_offsets, int32_t P, int64_t N, int32_t n,
                      double* gram, void* stream);
/* Solve (X^T X + lam I) c = X^T y by Cholesky, lam = reg_coeff * 10^k with the
 * reference's NaN retry ladder k = 0..9 (:30-38).  info4 = {lam used, retries,
 * error_before, error_after} where the errors are sum (y-Xc)^2 / sum y^2 (:25-28,40-44)
 * evaluated from the Gram with coeffs_prev (NULL = zeros) and the new coeffs. */
int mjrl_linbase_solve(mjrl_ctx* ctx, const double* gram, int32_t F, double reg_coeff,
                       const double* coeffs_prev, double* coeffs, double* info4,
                       void* stream);

/* ---- QuadraticBaseline: mjrl/baselines/quadratic_baseline.py ------------------------------ */
/* Features per timestep (:10-37): [o (n), o_i*o_j for i <= j row-major (n(n+1)/2), 1, al, al^2,
 * al^3, al^4], o = clip(obs,-10,10)/10, al = t_in_path/1000; F = n + n(n+1)/2 + 5 (n <= 16).
 * gram: same layout as mjrl_linbase_gram with this F; solve it with mjrl_linbase_solve(F)
 * (:51-58 is the same ridge + NaN ladder).  predict (:66-68): pred[t] = features(t) . coeffs. */
int mjrl_quadbase_gram(mjrl_ctx* ctx, const float* obs, const double* returns, const int32_t* path_offsets,
                       int32_t P, int64_t N, int32_t n, double* gram, void* stream);
int mjrl_quadbase_predict(mjrl_ctx* ctx, const float* obs, const int32_t* path_offsets, int32_t P,
                          int64_t N, int32_t n, const double* coeffs, double* pred, void* stream);

/* ---- (b) MLPBaseline: mjrl/baselines/mlp_baseline.py --------------------------------- */
/* (n+4)-128-128-1 ReLU value net, fp32 (:24-28).  Flat parameter vector in torch
 * `model.parameters()` order: fc_0.weight [128][n+4] | fc_0.bias | fc_1.weight [128][128] |
 * fc_1.bias | fc_2.weight [1][128] | fc_2.bias; returns the count (<0 on a bad n). */
int64_t mjrl_mlpbase_param_count(int32_t n);
/* _features (:37-56): feat[t] = [clip(o,-10,10)/10, al, al^2, al^3, al^4], al = t_in_path/1000,
 * fp64 arithmetic cast to fp32 [N, n+4] (:63).  If returns and y32 are non-NULL also
 * y32 = (float)returns (:64). */
int mjrl_mlpbase_features(mjrl_ctx* ctx, const float* obs, const double* returns,
                          const int32_t* path_offsets, int32_t P, int64_t N, int32_t n, float* feat,
                          float* y32, void* stream);
/* predict (:107-115) over the whole flat buffer; features are rebuilt in-kernel.  pred is
 * fp64 [N] holding the fp32 network output widened (the reference mixes the fp32 prediction
 * with fp64 rewards, process_samples.py:25-29, which numpy widens the same way). */
int mjrl_mlpbase_predict(mjrl_ctx* ctx, const float* obs, const int32_t* path_offsets, int32_t P,
                         int64_t N, int32_t n, const float* params, double* pred, void* stream);
/* sums2 = { sum ((float)ret - (float)pred)^2 , sum (float)ret^2 }: the two sums of the
 * error ratio of :75-81,98-104 (the caller adds 1e-8 to the denominator). */
int mjrl_mlpbase_sqerr(mjrl_ctx* ctx, const double* pred, const double* returns, int64_t N,
                       double* sums2, void* stream);
/* fit (:83-96): `nsteps` dependent Adam steps; step s trains on rows perm[s*batch .. (s+1)*batch)
 * of (feat, y32).  The host builds `perm` exactly like the reference: per epoch one
 * np.random.permutation(N), of which the first (int(N/batch)-1)*batch entries are used (:84-86).
 * params / adam_m / adam_v are updated in place; step0 = optimiser steps taken before this call
 * (torch.optim.Adam keeps its state across fit() calls).  MSE loss (mean), weight_decay is
 * added to the gradient, bias-corrected Adam as torch.optim.Adam.  batch <= 64. */
int mjrl_mlpbase_fit(mjrl_ctx* ctx, const float* feat, const float* y32, int64_t N, int32_t n,
                     float* params, float* adam_m, float* adam_v, int64_t step0, const int32_t* perm,
                     int64_t nsteps, int32_t batch, double lr, double beta1, double beta2, double eps,
                     double weight_decay, void* stream);

/* ---- (c)(d)(e) Gaussian policy: mjrl/policies/gaussian_mlp.py, gaussian_linear.py ---- */
/* Network description.  h1 == h2 == 0 selects the affine policy (LinearModel.forward,
 * gaussian_linear.py:171-175); otherwise the two-hidden-layer tanh MuNet
 * (gaussian_mlp.py:176-182).  Transform pointers are DEVICE fp32 arrays ([n],[n],[m],[m])
 * or NULL for identity: x~ = (x - in_shift)/(in_scale + 1e-8), mu = y*out_scale + out_shift. */
typedef struct mjrl_policy_desc {
  int32_t n, m, h1, h2;
  const float* in_shift;
  const float* in_scale;
  const float* out_shift;
  const float* out_scale;
} mjrl_policy_desc;

/* number of parameters d (gaussian_mlp.py:51-53); <0 on a bad description */
int64_t mjrl_policy_param_count(const mjrl_policy_desc* desc);
/* set_param_values' clamp (gaussian_mlp.py:75-76,87-88): params[d-m..d) = max(., min_log_std) */
int mjrl_policy_clamp_log_std(mjrl_ctx* ctx, const mjrl_policy_desc* desc, float* params,
                              float min_log_std, void* stream);
/* mean_LL (gaussian_mlp.py:100-110).  mean_out [N,m] and/or ll_out [N] may be NULL; act may be
 * NULL when ll_out is NULL. */
int mjrl_policy_forward(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                        const float* obs, const float* act, int64_t N, float* mean_out,
                        float* ll_out, void* stream);
/* CPI_surrogate + kl_old_new in one pass (batch_reinforce.py:37-49, gaussian_mlp.py:124-140).
 * old_ll [N], old_mean [N,m], old_log_std [m] describe the old policy on the same samples.
 * sums2 (fp64, device) = { sum_t exp(LL_new-LL_old)*w_t , sum_t KL_t(old||new) }; the caller
 * divides by the (global) sample count.  w may be NULL (surrogate sum = 0). */
int mjrl_policy_eval(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                     const float* obs, const float* act, const float* w, const float* old_ll,
                     const float* old_mean, const float* old_log_std, int64_t N, double* sums2,
                     void* stream);
/* flat_vpg (batch_reinforce.py:51-55): grad_sum[d] (fp64, device) = sum_t w_t LR_t dLL_t/dtheta,
 * LR_t = exp(LL_t - old_ll[t]) or 1 when old_ll is NULL (new == old).  surr_sum (fp64[1]) =
 * sum_t LR_t w_t.  Optionally caches LL [N] and mean [N,m] of this parameter set for later
 * mjrl_policy_eval calls.  The caller scales by sample_coef/N after any all-reduce. */
int mjrl_policy_grad(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                     const float* obs, const float* act, const float* w, const float* old_ll,
                     int64_t N, double* grad_sum, double* surr_sum, float* ll_out,
                     float* mean_out, void* stream);
/* Analytic Hessian-of-KL (Gauss-Newton / Fisher) product at new == old, replacing the double
 * backprop of npg_cg.py:55-74:  fvp_sum[d] (fp64, device) = sum_t J_t^T diag(2/(2 sigma^2+1e-8)) J_t vec
 * on the network block; the log_std block is N_local * h(sigma) * vec (exact second
 * derivative of the KL expression of gaussian_mlp.py:137-139).  No damping, no 1/N: the CG
 * update applies both after the all-reduce.  If stop_flag (device int32) is non-NULL and
 * non-zero the kernel exits immediately (CG already converged). */
int mjrl_policy_fvp(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                    const float* obs, int64_t N, const float* vec, double* fvp_sum,
                    const int32_t* stop_flag, void* stream);
/* The same product with the forward activations of the (fixed) policy kept in HBM between calls:
 * inside one CG solve (cg_solve.py:10-20 calls the operator with the SAME parameters and obs every
 * iteration, npg_cg.py:103-106) only `vec` changes, so tanh layers need computing once.
 * cache: device fp32 [mjrl_policy_fvp_cache_count(...)] (0 = this shape has no cached variant: pass
 * cache_mode 0).  cache_mode 1 = compute and fill, 2 = use (caller guarantees params/obs/N are
 * those of the filling call), 0 = plain mjrl_policy_fvp.  Modes 0 and 1 give identical bits; mode 2 reads
 * back the same activations and is deterministic, with its own fixed summation order (~1e-7 relative). */
int64_t mjrl_policy_fvp_cache_count(mjrl_ctx* ctx, const mjrl_policy_desc* desc, int64_t N);
/* Jacobian-cached product (act_dim <= 4, csrc/fvp_jac.cu): the m rows G1_j = d mu_j / d z1 of every
 * timestep do not depend on `vec` either, so with them cached next to h1, h2 the product needs neither
 * the tangent propagation nor the backward data products (43 % fewer MMAs, m*H*4 more bytes per
 * timestep).  mjrl_policy_fvp_jac_ok: 1 when the shape has this variant (mjrl_policy_fvp_cache_count
 * then includes the rows).  mjrl_policy_fvp_jac_fill computes the rows from an ALREADY FILLED h1|h2
 * cache (cache_mode 1 product or mjrl_policy_grad_cached at the same params / obs / N); products then
 * pass cache_mode 3.  Same result as modes 0-2 up to fp32 summation order (~1e-6 relative). */
int mjrl_policy_fvp_jac_ok(mjrl_ctx* ctx, const mjrl_policy_desc* desc);
int mjrl_policy_fvp_jac_fill(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params, int64_t N,
                             float* cache, void* stream);
/* Tensor-memory product (hidden 64 x 64, obs_dim <= 15, act_dim <= 8, csrc/fvp_tc.cu): the same Fisher-vector
 * product on the 5th-generation tensor cores (tcgen05.mma kind::tf32 with hi/lo error compensation, accumulators
 * and A operands in tensor memory).  It reads h1, h2 from a copy of the cache in its own tile order:
 * mjrl_policy_fvp_tc_ok: 1 when the shape has this variant (mjrl_policy_fvp_cache_count then includes the
 * copy).  mjrl_policy_fvp_tc_fill builds the copy from an ALREADY FILLED h1|h2 cache (once per CG solve, a pure
 * relayout); products then pass cache_mode 4.  Same result as the other modes to fp32 rounding (~1e-6).
 * Return value 2: a shape the fused kernel does not hold (obs_dim + 1 <= 128, act_dim <= 24; Ant 111 / 8, Adroit
 * 45 / 24) that runs the three-pass form of csrc/fvp_tc_wide.cu instead: first-layer tangent and weight-gradient
 * products as tcgen05 passes of their own around the main kernel.  Its cache (mjrl_policy_fvp_cache_count) holds the
 * planes, an [x~ | 1] image and two first-layer images; only mjrl_policy_grad_cached_tc fills it (no
 * mjrl_policy_fvp_tc_fill for these shapes). */
int mjrl_policy_fvp_tc_ok(mjrl_ctx* ctx, const mjrl_policy_desc* desc);
int mjrl_policy_fvp_tc_fill(mjrl_ctx* ctx, const mjrl_policy_desc* desc, int64_t N, float* cache, void* stream);
/* mjrl_policy_grad that also fills that cache (flat_vpg runs at the parameters the CG solve that
 * follows it uses, npg_cg.py:96-106): fvp_cache as above or NULL. */
int mjrl_policy_grad_cached(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                            const float* obs, const float* act, const float* w, const float* old_ll,
                            int64_t N, double* grad_sum, double* surr_sum, float* ll_out,
                            float* mean_out, float* fvp_cache, void* stream);
/* mjrl_policy_grad_cached for shapes with the tensor-memory product (mjrl_policy_fvp_tc_ok): h1, h2 go straight
 * into that product's tile-ordered copy (no fragment-ordered cache, no mjrl_policy_fvp_tc_fill pass); the CG
 * solve then runs cache_mode 4 from its first product.  Modes 1-3 are NOT valid after this call. */
int mjrl_policy_grad_cached_tc(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                               const float* obs, const float* act, const float* w, const float* old_ll,
                               int64_t N, double* grad_sum, double* surr_sum, float* ll_out,
                               float* mean_out, float* fvp_cache, void* stream);
/* mjrl_policy_eval (surrogate and KL sums of batch_reinforce.py:57-75) on the tcgen05 forward pass of csrc/pol_fwd.cu, for
 * shapes mjrl_policy_fvp_tc_ok accepts.  The gradient pass mjrl_policy_grad_cached_tc runs on tcgen05 too (csrc/pol_tc.cu)
 * unless MJRL_B200_GRAD_TC=0.  Shapes of kind 2 read the [x~ | 1] image that mjrl_policy_grad_cached_tc left in
 * `fvp_cache` for the SAME obs / N / input transforms (fvp_cache may be NULL for kind 1).  Same sums as
 * mjrl_policy_eval to fp32 rounding. */
int mjrl_policy_eval_tc(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                        const float* obs, const float* act, const float* w, const float* old_ll,
                        const float* old_mean, const float* old_log_std, int64_t N, double* sums2,
                        float* fvp_cache, void* stream);
int mjrl_policy_fvp_cached(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                           const float* obs, int64_t N, const float* vec, double* fvp_sum,
                           const int32_t* stop_flag, float* cache, int32_t cache_mode, void* stream);

/* ---- (d) conjugate gradient on device: mjrl/utils/cg_solve.py:3-22 ------------------- */
/* CG state: caller-allocated DEVICE arrays.  x, r, p, z are fp64 [d] (z is the operator
 * output slot that mjrl_policy_fvp fills and the caller all-reduces); p32 is p cast to fp32
 * [d] (the operator input); scal fp64 [8] = {rdotr, p.z, v, mu, rdotr0, -, -, -};
 * flags int32 [2] = {stop, iterations done}. */
typedef struct mjrl_cg_state {
  double* x;
  double* r;
  double* p;
  double* z;
  double* scal;
  float* p32;
  int32_t* flags;
} mjrl_cg_state;
/* x = 0, r = p = b (cg_solve.py:4-7; x_0 is ignored exactly like the reference, :4-5). */
int mjrl_cg_init(mjrl_ctx* ctx, const mjrl_cg_state* state, int64_t d, const double* b, double b_scale,
                 void* stream);
/* One iteration (cg_solve.py:10-20) given z_sum = sum over samples of the operator applied to
 * p32, as left in state->z by mjrl_policy_fvp (+ all-reduce):  z = z_sum*inv_count + damping*p;
 * v = rdotr/p.z; x += v p; r -= v z; mu = r.r/rdotr; p = r + mu p; stop when r.r < tol. */
int mjrl_cg_update(mjrl_ctx* ctx, const mjrl_cg_state* state, int64_t d, double inv_count, double damping,
                   double residual_tol, void* stream);
/* Step-size rule + parameter update (npg_cg.py:130-141 / trpo.py:100-108 / dapg.py:111-118):
 * gx = g.x; alpha = const_alpha >= 0 ? const_alpha : sqrt(|delta/(gx + 1e-20)|);
 * new_params = curr + (alpha*alpha_scale)*x with log_std clamped at min_log_std.
 * scal4 (fp64, device) = {alpha, gx, n_step_size, alpha*alpha_scale}. */
int mjrl_npg_step(mjrl_ctx* ctx, int64_t d, int32_t m, const double* g, double g_scale,
                  const double* x, const float* curr, double delta, double const_alpha,
                  double alpha_scale, float min_log_std, float* new_params, double* scal4,
                  void* stream);
/* (e) batched get_action (gaussian_mlp.py:92-98): act = mean + exp(log_std)*eps, eps ~ N(0,1)
 * from Philox4x32-10 keyed by (seed, row index + offset). */
int mjrl_policy_act(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                    const float* obs, int64_t B, uint64_t seed, uint64_t offset, float* act_out,
                    float* mean_out, void* stream);

/* ---- lock-step rollout (north_star (e)) ---------------------------------
 * One simulator step of B environment slots, replacing B calls of policy.get_action inside the
 * rollout loop (mjrl/samplers/base_sampler.py:64-66, evaluation_sampler.py:57-58):
 *   mean = policy(obs); act = sample ? mean + exp(log_std)*eps : mean.
 * eps = noise[b][j] when `noise` is given (host-generated, e.g. the reference's per-episode
 * np.random stream of base_sampler.py:47-50), else Philox4x32-10 keyed by seeds[b] (the
 * episode's seed) with counter (tsteps[b], j/4): a function of (episode, step, component) only.
 * When dst_rows[b] >= 0 the slot's observation and action rows are also written to
 * obs_store[dst_rows[b]] / act_store[dst_rows[b]] (device-resident timestep store; either may
 * be NULL).  All pointers are device pointers; seeds/tsteps may be NULL when noise is given or
 * sample == 0. */
int mjrl_rollout_act(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                     const float* obs, int32_t B, const uint64_t* seeds, const int32_t* tsteps,
                     const float* noise, int32_t sample, const int64_t* dst_rows,
                     float* obs_store, float* act_store, float* act_out, float* mean_out,
                     void* stream);
/* dst[r][0:width] = src[src_rows[r]][0:width], r < rows: squeezes the padded [episode][T] store
 * into the flat path-major buffer the update kernels take (the np.concatenate of
 * npg_cg.py:87-89 over paths of unequal length). */
int mjrl_rows_gather_f32(mjrl_ctx* ctx, const float* src, const int64_t* src_rows, int64_t rows,
                         int32_t width, float* dst, void* stream);

/* ---- PPO minibatch epochs: mjrl/algos/ppo_clip.py:47-54,85-101 ------------------------- */
/* dst[b] = src[idx[b]] for rows of `width` floats (a with-replacement minibatch, :89-92). */
int mjrl_gather_rows_f32(mjrl_ctx* ctx, const float* src, const int32_t* idx, int32_t B, int32_t width,
                         float* dst, void* stream);
/* Gradient weights of PPO_surrogate (:47-54): w_t = adv_t where LR_t*adv_t <= clamp(LR_t)*adv_t,
 * else 0, so that the gradient of mean(min(.,.)) is (1/B) sum_t w_t LR_t dLL_t, which
 * mjrl_policy_grad computes from (w, old_ll).  surr_sum (fp64[1], may be NULL) = sum_t min(.,.). */
int mjrl_ppo_clip_weights(mjrl_ctx* ctx, const float* ll_new, const float* ll_old, const float* adv,
                          int32_t B, float clip_coef, float* w, double* surr_sum, void* stream);
/* One torch.optim.Adam step (:34,96) on a flat fp32 vector from an fp64 gradient SUM:
 * g = grad_scale*grad_sum (+ weight_decay*p); `step` counts from 1. */
int mjrl_adam_step(mjrl_ctx* ctx, int64_t d, float* params, const double* grad_sum, double grad_scale,
                   float* adam_m, float* adam_v, int64_t step, double lr, double beta1, double beta2,
                   double eps, double weight_decay, void* stream);

/* ---- Behaviour cloning: mjrl/algos/behavior_cloning.py:43-68, behavior_cloning_2.py:56-77 -- */
/* `steps` serial Adam steps on the flat policy vector.  Step s gathers rows idx[s*B .. s*B+B) of
 * obs/act (np.random.choice minibatch, :56 / :66), forms grad_sum[d] = sum_t w_t dLL_t/dtheta at the
 * CURRENT params (LR = 1) and applies mjrl_adam_step with g = grad_scale*grad_sum to the first d_opt
 * parameters (d for behavior_cloning.py, loss -mean LL: grad_scale = -1/B; d - m for
 * behavior_cloning_2.py, whose MSE loss is the log_std = 0 likelihood: grad_scale = -2/(B*m) on a
 * working vector with a zeroed log_std block).  Adam step numbers are step0+1 .. step0+steps.
 * Scratch (device): mb_obs [B,n], mb_act [B,m], mb_w [B] (the weights, ones for BC), grad_sum fp64 [d].
 * No host synchronisation inside. */
int mjrl_policy_adam_chain(mjrl_ctx* ctx, const mjrl_policy_desc* desc, float* params, const float* obs,
                           const float* act, const int32_t* idx, int32_t steps, int32_t B, int64_t d_opt,
                           double grad_scale, float* adam_m, float* adam_v, int64_t step0, double lr,
                           double beta1, double beta2, double eps, double weight_decay, float* mb_obs,
                           float* mb_act, const float* mb_w, double* grad_sum, void* stream);

/* ---- (e) multi-GPU: small fp64 sum all-reduces over NVLink peer memory ---------------------
 * Replaces the collectives SURVEY 8(e) lists (gradient, each Fisher-vector product, Gram, moments,
 * surrogate/KL sums; <= 107 KB each, latency bound) by ONE single-CTA kernel per exchange that
 * signals the peers, pulls their contributions over NVLink and adds them in rank order (all ranks
 * get bit-identical sums); mjrl_cg_update_peer fuses that exchange into the CG update
 * (cg_solve.py:10-20), so a CG iteration is: product kernel -> partial reduce -> this kernel.
 * One process per GPU: every rank creates its block, the 64-byte handles are exchanged by the host
 * (torch.distributed all_gather), every rank connects.  `epoch` is a host counter starting at 1 that
 * ALL ranks advance by one per exchange, in the same order.  A peer that does not arrive within the
 * timeout (default 20 s) sets the error word read by mjrl_peer_status instead of hanging the GPU. */
typedef struct mjrl_peer mjrl_peer;
int mjrl_peer_handle_bytes(void);
int mjrl_peer_create(mjrl_ctx* ctx, int32_t rank, int32_t world, int64_t max_count, mjrl_peer** out,
                     void* handle_out);
/* handles: world * mjrl_peer_handle_bytes() bytes in rank order (own entry ignored) */
int mjrl_peer_connect(mjrl_peer* peer, const void* handles);
/* same-process variant (ranks that share an address space, e.g. tests): bases[r] = mjrl_peer_base of rank r */
int mjrl_peer_connect_ptrs(mjrl_peer* peer, void* const* bases);
void* mjrl_peer_base(mjrl_peer* peer);
/* device pointer of the local contribution slot of `epoch`: producers may write there directly */
double* mjrl_peer_slot(mjrl_peer* peer, int64_t epoch);
int mjrl_peer_set_timeout(mjrl_peer* peer, double seconds);
/* dst[0..count) = sum over ranks of their src[0..count); src == NULL: already in mjrl_peer_slot(epoch).
 * dst may alias src. */
int mjrl_peer_allreduce_f64(mjrl_ctx* ctx, mjrl_peer* peer, const double* src, double* dst, int64_t count,
                            int64_t epoch, void* stream);
/* mjrl_cg_update with z_sum = sum over ranks of mjrl_peer_slot(epoch)[0..d) */
int mjrl_cg_update_peer(mjrl_ctx* ctx, mjrl_peer* peer, const mjrl_cg_state* state, int64_t d,
                        double inv_count, double damping, double residual_tol, int64_t epoch, void* stream);
/* cg_solve.py:3-22 in one call: mjrl_cg_init(b, b_scale), then `iters` x ( mjrl_policy_fvp_cached(vec =
 * state->p32) -> mjrl_cg_update ), enqueued from C with no host work in between.  The first product uses
 * cache_mode `mode_first`, the others `mode_rest`; jac_fill_after_first != 0 runs mjrl_policy_fvp_jac_fill
 * after the first product (mode_first 1, mode_rest 3); jac_fill_after_first == 2 runs mjrl_policy_fvp_tc_fill
 * there instead (mode_rest 4).  peer != NULL: the products write into the exchange
 * slots of epochs epoch0+1 .. epoch0+iters and the updates are mjrl_cg_update_peer (the caller advances its
 * epoch counter by `iters`); peer == NULL: single rank. */
int mjrl_cg_solve(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params, const float* obs, int64_t N,
                  const mjrl_cg_state* state, const double* b, double b_scale, int32_t iters, double inv_count,
                  double damping, double residual_tol, float* cache, int32_t mode_first, int32_t mode_rest,
                  int32_t jac_fill_after_first, mjrl_peer* peer, int64_t epoch0, void* stream);
/* synchronous: *err_out = 1 when some wait timed out since creation */
int mjrl_peer_status(mjrl_peer* peer, int32_t* err_out);
/* Measurement aid, no reference counterpart: with MJRL_B200_PEER_STAMPS=1 in the environment when the block is created, every
 * exchange kernel records four %globaltimer values (ns) for its epoch: 0 kernel entry, 1 every peer's flag seen, 2 sums pulled
 * over NVLink (+ the first dot product in the fused CG variant), 3 kernel end.  Copies the stamps of epochs first_epoch ..
 * first_epoch + count - 1 (count <= 4096, ring) to the host array out[count][4].  bench.py prints their medians at N > 1. */
int mjrl_peer_stamps(mjrl_peer* peer, uint64_t* out, int64_t first_epoch, int64_t count);
int mjrl_peer_destroy(mjrl_peer* peer);

#ifdef __cplusplus
}
#endif
#endif /* MJRL_B200_H */
