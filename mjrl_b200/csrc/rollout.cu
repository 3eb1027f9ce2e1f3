// Lock-step rollout support (north_star (e)): one launch pair per simulator step for ALL
// environment slots, replacing the per-environment `policy.get_action(o)` of the reference's hot loop
// (mjrl/samplers/base_sampler.py:64-66 -> mjrl/policies/gaussian_mlp.py:92-98).
//
// The step kernel also files the step's observations and actions into the device-resident timestep
// store (row = episode * T + t), so the update that follows never uploads them again: the only H2D the
// update needs after a lock-step rollout is the reward column.
#include "common.cuh"
#include "philox.cuh"

namespace rollout {

// act[b][j] = mean[b][j] + exp(log_std_j) * eps[b][j] (sample != 0) or mean[b][j] (evaluation rollouts,
// evaluation_sampler.py:57-58).  eps comes from `noise` when given (the host's MT19937 stream of the
// reference, trajectory by trajectory), else from Philox4x32-10 with key = episode seed and counter =
// (t, j/4): a function of (episode, step, component) only, so the sampled batch does not depend on how
// many slots step together nor on which slot ran which episode.
__global__ void act_store_kernel(const float* __restrict__ obs, const float* __restrict__ mean,
                                 const float* __restrict__ log_std, int n, int m, int B,
                                 const unsigned long long* __restrict__ seeds, const int* __restrict__ tsteps,
                                 const float* __restrict__ noise, int sample,
                                 const long long* __restrict__ dst_rows, float* __restrict__ obs_store,
                                 float* __restrict__ act_store, float* __restrict__ act_out) {
  const int nblk = (m + 3) / 4;
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * nblk; i += stride) {
    const int b = i / nblk, jb = i - b * nblk;
    float e[4] = {0.f, 0.f, 0.f, 0.f};
    if (sample && !noise) {
      const unsigned long long key = seeds[b];
      uint32_t c[4] = {(uint32_t)tsteps[b], 0u, (uint32_t)jb, 0x726f6c6cu};
      pol::philox4x32_10(c, (uint32_t)key, (uint32_t)(key >> 32));
      pol::philox_normal4(c, e);
    }
    const long long dst = dst_rows ? dst_rows[b] : -1;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int j = jb * 4 + q;
      if (j >= m) break;
      float a = mean[b * m + j];
      if (sample) a += expf(log_std[j]) * (noise ? noise[b * m + j] : e[q]);
      act_out[b * m + j] = a;
      if (dst >= 0 && act_store) act_store[dst * m + j] = a;
    }
  }
  if (obs_store && dst_rows) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * n; i += stride) {
      const int b = i / n, k = i - b * n;
      const long long dst = dst_rows[b];
      if (dst >= 0) obs_store[dst * n + k] = obs[i];
    }
  }
}

// dst[r][:] = src[src_rows[r]][:]: squeezes the padded [episode][T] store into the flat path-major
// timestep buffer once the path lengths are known (ragged batches only).
__global__ void gather_rows_kernel(const float* __restrict__ src, const long long* __restrict__ src_rows,
                                   long long rows, int width, float* __restrict__ dst) {
  const long long total = rows * width, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long r = i / width;
    const int k = (int)(i - r * width);
    dst[i] = ldg_stream(src + src_rows[r] * width + k);
  }
}

}  // namespace rollout

extern "C" {

int mjrl_rollout_act(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params, const float* obs,
                     int32_t B, const uint64_t* seeds, const int32_t* tsteps, const float* noise,
                     int32_t sample, const int64_t* dst_rows, float* obs_store, float* act_store,
                     float* act_out, float* mean_out, void* stream) {
  MJRL_REQUIRE(ctx && desc && params && obs && act_out && mean_out && B > 0, MJRL_E_ARG,
               "rollout_act: bad argument");
  MJRL_REQUIRE(!sample || noise || (seeds && tsteps), MJRL_E_ARG,
               "rollout_act: sampling needs either a noise block or (seeds, tsteps)");
  MJRL_REQUIRE((!obs_store && !act_store) || dst_rows, MJRL_E_ARG, "rollout_act: store without dst_rows");
  int rc = mjrl_policy_forward(ctx, desc, params, obs, nullptr, B, mean_out, nullptr, stream);
  if (rc) return rc;
  const int n = desc->n, m = desc->m;
  // log_std is the last m entries of the flat vector (gaussian_mlp.py:51-53; affine policy: h1 == 0)
  const long long d = desc->h1 > 0 ? (long long)n * desc->h1 + desc->h1 + (long long)desc->h1 * desc->h2 + desc->h2 +
                                         (long long)desc->h2 * m + m + m
                                   : (long long)n * m + m + m;
  const float* log_std = params + (d - m);
  const long long work = (long long)B * (n > m ? n : m);
  int grid = (int)((work + 255) / 256);
  if (grid > ctx->sm_count * 4) grid = ctx->sm_count * 4;
  if (grid < 1) grid = 1;
  rollout::act_store_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      obs, mean_out, log_std, n, m, B, (const unsigned long long*)seeds, tsteps, noise, sample,
      (const long long*)dst_rows, obs_store, act_store, act_out);
  return mjrl_check_launch("rollout_act");
}

int mjrl_rows_gather_f32(mjrl_ctx* ctx, const float* src, const int64_t* src_rows, int64_t rows, int32_t width,
                         float* dst, void* stream) {
  MJRL_REQUIRE(ctx && src && src_rows && dst && rows >= 0 && width > 0, MJRL_E_ARG, "rows_gather: bad argument");
  if (rows == 0) return 0;
  long long blocks = (rows * width + 255) / 256;
  const long long cap = (long long)ctx->sm_count * 8;
  rollout::gather_rows_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(
      src, (const long long*)src_rows, rows, width, dst);
  return mjrl_check_launch("rows_gather");
}

}  // extern "C"
