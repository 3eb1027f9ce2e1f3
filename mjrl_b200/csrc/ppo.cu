// PPO's minibatch epochs (mjrl/algos/ppo_clip.py:47-54,85-101) on the device: row gather for a
// with-replacement minibatch, the clipped-surrogate gradient weights, and the Adam update of
// the flat policy vector.  The per-sample forward / gradient passes are the policy kernels of
// policy.cu; the chain of steps itself is serial (same shape as MLPBaseline.fit).
#include "common.cuh"

namespace ppo {

__global__ void gather_rows_kernel(const float* __restrict__ src, const int32_t* __restrict__ idx, int B,
                                   int width, float* __restrict__ dst) {
  const int total = B * width;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / width, k = i - b * width;
    dst[i] = src[(long long)idx[b] * width + k];
  }
}

// PPO_surrogate = mean(min(LR*adv, clamp(LR, 1-c, 1+c)*adv)) (ppo_clip.py:47-54).  Its gradient is
// (1/B) sum_t w_t LR_t dLL_t with w_t = adv_t where the unclipped term is the minimum (ties
// included: torch.min splits a tie's gradient between both arguments and the clamp passes its
// half through), 0 where the clipped term is strictly smaller (clamp saturated).
__global__ void clip_weights_kernel(const float* __restrict__ ll_new, const float* __restrict__ ll_old,
                                    const float* __restrict__ adv, int B, float clip, float* __restrict__ w,
                                    double* __restrict__ surr_sum) {
  __shared__ double red[32];
  double acc = 0.0;
  for (int t = threadIdx.x; t < B; t += blockDim.x) {
    const float lr = expf(ll_new[t] - ll_old[t]);
    const float lrc = fminf(fmaxf(lr, 1.f - clip), 1.f + clip);
    const float a = lr * adv[t], b = lrc * adv[t];
    w[t] = a <= b ? adv[t] : 0.f;
    acc += (double)fminf(a, b);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0 && surr_sum) surr_sum[0] = acc;
}

// torch.optim.Adam (no amsgrad): g = grad_scale * grad_sum (+ wd * p); m, v, bias-corrected step.
__global__ void adam_kernel(long long d, float* __restrict__ p, const double* __restrict__ grad_sum,
                            double grad_scale, float* __restrict__ m, float* __restrict__ v, float lr_over_bc1,
                            float bc2_sqrt, float one_m_b1, float b2, float one_m_b2, float eps, float wd) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < d; i += (long long)gridDim.x * blockDim.x) {
    float g = (float)(grad_scale * grad_sum[i]);
    const float w = p[i];
    g = fmaf(wd, w, g);
    const float mi = m[i] + one_m_b1 * (g - m[i]);
    const float vi = fmaf(one_m_b2 * g, g, v[i] * b2);
    m[i] = mi; v[i] = vi;
    p[i] = w - lr_over_bc1 * (mi / (sqrtf(vi) / bc2_sqrt + eps));
  }
}

}  // namespace ppo

extern "C" {

int mjrl_gather_rows_f32(mjrl_ctx* ctx, const float* src, const int32_t* idx, int32_t B, int32_t width,
                         float* dst, void* stream) {
  MJRL_REQUIRE(ctx && src && idx && dst && B >= 0 && width > 0, MJRL_E_ARG, "gather_rows: bad argument");
  if (B == 0) return 0;
  const int total = B * width;
  ppo::gather_rows_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(src, idx, B, width, dst);
  return mjrl_check_launch("gather_rows_kernel");
}

int mjrl_ppo_clip_weights(mjrl_ctx* ctx, const float* ll_new, const float* ll_old, const float* adv,
                          int32_t B, float clip_coef, float* w, double* surr_sum, void* stream) {
  MJRL_REQUIRE(ctx && ll_new && ll_old && adv && w && B > 0, MJRL_E_ARG, "ppo_clip_weights: bad argument");
  ppo::clip_weights_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(ll_new, ll_old, adv, B, clip_coef, w, surr_sum);
  return mjrl_check_launch("ppo_clip_weights_kernel");
}

int mjrl_adam_step(mjrl_ctx* ctx, int64_t d, float* params, const double* grad_sum, double grad_scale,
                   float* adam_m, float* adam_v, int64_t step, double lr, double beta1, double beta2,
                   double eps, double weight_decay, void* stream) {
  MJRL_REQUIRE(ctx && params && grad_sum && adam_m && adam_v && d > 0 && step >= 1, MJRL_E_ARG,
               "adam_step: bad argument");
  const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
  const int grid = (int)((d + 255) / 256);
  ppo::adam_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      d, params, grad_sum, grad_scale, adam_m, adam_v, (float)(lr / bc1), (float)sqrt(bc2),
      (float)(1.0 - beta1), (float)beta2, (float)(1.0 - beta2), (float)eps, (float)weight_decay);
  return mjrl_check_launch("adam_kernel");
}

// Behaviour cloning (mjrl/algos/behavior_cloning.py:52-61, behavior_cloning_2.py:62-71): a serial chain of
// Adam steps on with-replacement minibatches.  Host loop over the step kernels: no host work, no
// synchronisation between steps, so the chain runs at launch rate.
int mjrl_policy_adam_chain(mjrl_ctx* ctx, const mjrl_policy_desc* desc, float* params, const float* obs,
                           const float* act, const int32_t* idx, int32_t steps, int32_t B, int64_t d_opt,
                           double grad_scale, float* adam_m, float* adam_v, int64_t step0, double lr,
                           double beta1, double beta2, double eps, double weight_decay, float* mb_obs,
                           float* mb_act, const float* mb_w, double* grad_sum, void* stream) {
  MJRL_REQUIRE(ctx && desc && params && obs && act && idx && adam_m && adam_v && mb_obs && mb_act && mb_w &&
                   grad_sum && steps >= 0 && B > 0 && d_opt > 0 && step0 >= 0,
               MJRL_E_ARG, "policy_adam_chain: bad argument");
  const int64_t d = mjrl_policy_param_count(desc);
  MJRL_REQUIRE(d > 0 && d_opt <= d, MJRL_E_SHAPE, "policy_adam_chain: d_opt exceeds the parameter count");
  for (int32_t s = 0; s < steps; ++s) {
    const int32_t* rows = idx + (int64_t)s * B;
    int rc = mjrl_gather_rows_f32(ctx, obs, rows, B, desc->n, mb_obs, stream);
    if (rc) return rc;
    rc = mjrl_gather_rows_f32(ctx, act, rows, B, desc->m, mb_act, stream);
    if (rc) return rc;
    rc = mjrl_policy_grad(ctx, desc, params, mb_obs, mb_act, mb_w, nullptr, B, grad_sum, nullptr, nullptr,
                          nullptr, stream);
    if (rc) return rc;
    rc = mjrl_adam_step(ctx, d_opt, params, grad_sum, grad_scale, adam_m, adam_v, step0 + s + 1, lr, beta1,
                        beta2, eps, weight_decay, stream);
    if (rc) return rc;
  }
  return 0;
}

}  // extern "C"
