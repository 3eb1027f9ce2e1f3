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

#ifdef __cplusplus
}
#endif
#endif /* MJRL_B200_H */
