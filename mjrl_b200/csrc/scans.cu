// (a) Discounted returns and GAE as segmented reverse scans over the flat timestep
// buffer, plus the small fp64 reductions around them.
//
// Replaces mjrl/utils/process_samples.py:3-5,21-29,37-43 (python for-loop per timestep)
// and the whitening statistics of mjrl/algos/npg_cg.py:91.
//
// Layout: one CTA per path (a path is a segment [off[p], off[p+1]) of the flat buffer,
// so no carry ever crosses a CTA).  A path is walked from its END in chunks of
// CHUNK = 128 threads x 8 elements.  Each chunk is loaded with fully coalesced 8-byte
// accesses into padded shared memory, every thread scans its 8 consecutive elements
// in registers, and the per-thread carries are combined with a warp-shuffle scan of the
// linear recurrence  C_t = A_{t+1} + g^8 C_{t+1}  (then 4 warp aggregates through smem).
// HBM traffic = algorithmic bytes: returns R8+W8, GAE R8+R8+W8 per timestep.
#include "common.cuh"

namespace {

constexpr int SCAN_THREADS = 128;
constexpr int SCAN_E = 8;
constexpr int SCAN_CHUNK = SCAN_THREADS * SCAN_E;  // 1024
__device__ __forceinline__ int padi(int i) { return i + (i >> 3); }  // conflict-free blocked reads

template <bool GAE>
__global__ void __launch_bounds__(SCAN_THREADS, 8)
segscan_kernel(const double* __restrict__ rew, const double* __restrict__ base,
               const int32_t* __restrict__ off, const uint8_t* __restrict__ term, int P,
               double g_td /*gamma*/, double g /*scan discount*/, double* __restrict__ out) {
  __shared__ double sx[SCAN_CHUNK + SCAN_CHUNK / 8 + 8];
  __shared__ double swarp[SCAN_THREADS / 32];
  __shared__ double scarry;

  const int p = blockIdx.x;
  if (p >= P) return;
  const int64_t s0 = off[p], s1 = off[p + 1];
  const int T = (int)(s1 - s0);
  if (T <= 0) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // powers g^1..g^8 and the shuffle-scan multipliers (g^8)^(2^k)
  double gp[SCAN_E + 1];
  gp[0] = 1.0;
#pragma unroll
  for (int k = 1; k <= SCAN_E; ++k) gp[k] = gp[k - 1] * g;
  const double G8 = gp[SCAN_E];
  double G256 = G8;  // (g^8)^32 = g^256 : one warp's span
#pragma unroll
  for (int k = 0; k < 5; ++k) G256 *= G256;

  double b_tail = 0.0;  // bootstrap value b1[T]
  if (GAE) b_tail = term[p] ? 0.0 : base[s1 - 1];

  double carry = 0.0;  // y at the first element of the previously processed (later) chunk
  const int nchunks = (T + SCAN_CHUNK - 1) / SCAN_CHUNK;
  for (int c = nchunks - 1; c >= 0; --c) {
    const int c0 = c * SCAN_CHUNK;
    const int L = min(SCAN_CHUNK, T - c0);
    const int64_t gbase = s0 + c0;
    // ---- coalesced load: every request of the chunk is issued before the first use.  GAE forms
    // the TD residual here, in the coalesced domain: b[t+1] is a second, one-element-shifted
    // coalesced read of the same sectors (an L2 hit), so only the residual goes through smem.
    {
      double r[SCAN_E], b[GAE ? SCAN_E : 1], bn[GAE ? SCAN_E : 1];
#pragma unroll
      for (int k = 0; k < SCAN_E; ++k) {
        const int i = k * SCAN_THREADS + tid;
        r[k] = (i < L) ? ldg_stream(rew + gbase + i) : 0.0;
      }
      if (GAE) {
#pragma unroll
        for (int k = 0; k < SCAN_E; ++k) {
          const int i = k * SCAN_THREADS + tid;
          b[k] = (i < L) ? ldg_stream(base + gbase + i) : 0.0;
          bn[k] = (i < L && c0 + i + 1 < T) ? __ldg(base + gbase + i + 1) : b_tail;
        }
      }
#pragma unroll
      for (int k = 0; k < SCAN_E; ++k) {
        const int i = k * SCAN_THREADS + tid;
        double x = r[k];
        // (r + gamma*b[t+1]) - b[t], same association as process_samples.py:28
        if (GAE) x = (i < L) ? __dsub_rn(__dadd_rn(x, __dmul_rn(g_td, bn[k])), b[k]) : 0.0;
        sx[padi(i)] = x;
      }
    }
    __syncthreads();
    // ---- blocked read, per-thread reverse scan -------------------------------
    double y[SCAN_E];
    const int e0 = tid * SCAN_E;
#pragma unroll
    for (int j = 0; j < SCAN_E; ++j) y[j] = sx[padi(e0 + j)];
    // inject the carry of the later chunk as a virtual addend on the last valid element
    if (c != nchunks - 1 && tid == SCAN_THREADS - 1)  // L == SCAN_CHUNK here
      y[SCAN_E - 1] += g * carry;
#pragma unroll
    for (int j = SCAN_E - 2; j >= 0; --j) y[j] = fma(g, y[j + 1], y[j]);
    // ---- reverse scan of thread aggregates: S_t = A_t + G8 * S_{t+1} --------
    double S = y[0];
    double mult = G8;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      double up = __shfl_down_sync(0xffffffffu, S, o);
      if (lane + o < 32) S = fma(mult, up, S);
      mult *= mult;
    }
    if (lane == 0) swarp[warp] = S;
    __syncthreads();
    // carry entering this warp from the warps to its right
    double wc = 0.0;
    for (int w = SCAN_THREADS / 32 - 1; w > warp; --w) wc = fma(G256, wc, swarp[w]);
    // carry entering this thread: S_{t+1} within the warp + its share of the warp carry
    double C = __shfl_down_sync(0xffffffffu, S, 1);
    if (lane == 31) C = 0.0;
    // warp carry enters at the warp's right edge, (31-lane) threads away: G8^(31-lane)
    {
      double pw = 1.0, b8 = G8;
      int ex = 31 - lane;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        if (ex & (1 << k)) pw *= b8;
        b8 *= b8;
      }
      C = fma(pw, wc, C);
    }
    // ---- finalize and store through smem (coalesced) ------------------------
#pragma unroll
    for (int j = 0; j < SCAN_E; ++j) sx[padi(e0 + j)] = fma(gp[SCAN_E - j], C, y[j]);
    __syncthreads();
    if (tid == 0) scarry = sx[0];
#pragma unroll
    for (int k = 0; k < SCAN_E; ++k) {
      int i = k * SCAN_THREADS + tid;
      if (i < L) out[gbase + i] = sx[padi(i)];
    }
    __syncthreads();
    carry = scarry;
  }
}

__global__ void adv_plain_kernel(const double* __restrict__ ret, const double* __restrict__ base,
                                 int64_t N, double* __restrict__ adv) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < N; i += stride) adv[i] = ret[i] - base[i];
}

// ---- moments: two passes (sum -> mean, then centred sum of squares), fixed order --
template <bool CENTERED>
__global__ void __launch_bounds__(256)
moment_partial_kernel(const double* __restrict__ x, int64_t N, const double* __restrict__ mom,
                      double* __restrict__ partial) {
  __shared__ double red[32];
  const double mean = CENTERED ? mom[3] : 0.0;
  // contiguous slab per CTA, fixed assignment -> deterministic
  int64_t per = (N + gridDim.x - 1) / gridDim.x;
  int64_t b = (int64_t)blockIdx.x * per, e = min(N, b + per);
  double acc = 0.0;
  for (int64_t i = b + threadIdx.x; i < e; i += blockDim.x) {
    double v = ld_stream(x + i);
    if (CENTERED) { v -= mean; acc = fma(v, v, acc); } else acc += v;
  }
  double r = block_sum(acc, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = r;
}

template <bool CENTERED>
__global__ void moment_final_kernel(const double* __restrict__ partial, int nparts, int64_t N,
                                    double* __restrict__ mom) {
  __shared__ double red[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += partial[i];
  double r = block_sum(acc, red);
  if (threadIdx.x == 0) {
    if (CENTERED) {
      mom[2] = r;
    } else {
      mom[0] = (double)N;
      mom[1] = r;
      mom[3] = N > 0 ? r / (double)N : 0.0;
    }
  }
}

__global__ void normalize_kernel(double* __restrict__ x, int64_t N, const double* __restrict__ mom,
                                 double eps) {
  const double mean = mom[3], sd = sqrt(mom[2] / mom[0]);
  const double den = sd + eps;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < N; i += stride) x[i] = (x[i] - mean) / den;
}

__global__ void whiten_f32_kernel(const double* __restrict__ adv, int64_t N,
                                  const double* __restrict__ mom, double eps, double scale,
                                  float* __restrict__ w) {
  const double mean = mom[3], sd = sqrt(mom[2] / mom[0]);
  const double den = sd + eps;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < N; i += stride) {
    double v = (ld_stream(adv + i) - mean) / den;
    if (scale != 1.0) v *= scale;
    w[i] = (float)v;
  }
}

// per-path reward sums (sequential like python's sum(), one warp per path with a fixed
// lane-strided order), then {mean, std, min, max} over paths
__global__ void __launch_bounds__(256)
path_sum_kernel(const double* __restrict__ rew, const int32_t* __restrict__ off, int P,
                double* __restrict__ sums) {
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= P) return;
  int64_t s0 = off[warp], s1 = off[warp + 1];
  double acc = 0.0;
  for (int64_t i = s0 + lane; i < s1; i += 32) acc += ld_stream(rew + i);
  acc = warp_sum(acc);
  if (lane == 0) sums[warp] = acc;
}

__global__ void path_stats_kernel(const double* __restrict__ sums, int P, double* __restrict__ st) {
  __shared__ double red[32];
  __shared__ double smean;
  double acc = 0.0, mn = INFINITY, mx = -INFINITY;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    double v = sums[i];
    acc += v; mn = fmin(mn, v); mx = fmax(mx, v);
  }
  double tot = block_sum(acc, red);
  if (threadIdx.x == 0) smean = tot / P;
  __syncthreads();
  double mean = smean, a2 = 0.0;
  for (int i = threadIdx.x; i < P; i += blockDim.x) { double d = sums[i] - mean; a2 = fma(d, d, a2); }
  double ss = block_sum(a2, red);
  // min / max
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  __shared__ double smn[32], smx[32];
  int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  if (l == 0) { smn[w] = mn; smx[w] = mx; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < nw; ++i) { mn = fmin(mn, smn[i]); mx = fmax(mx, smx[i]); }
    st[0] = mean; st[1] = sqrt(ss / P); st[2] = mn; st[3] = mx;
  }
}

__global__ void cast_kernel(const double* __restrict__ src, int64_t n, float* __restrict__ dst) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) dst[i] = (float)ld_stream(src + i);
}

inline int ew_grid(const mjrl_ctx* ctx, int64_t N, int threads) {
  int64_t want = (N + threads - 1) / threads;
  int64_t cap = (int64_t)ctx->sm_count * 16;
  return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

}  // namespace

extern "C" {

int mjrl_returns_scan(mjrl_ctx* ctx, const double* rew, const int32_t* off, int32_t P, int64_t N,
                      double gamma, double* returns, void* stream) {
  MJRL_REQUIRE(ctx && P >= 0 && N >= 0, MJRL_E_ARG, "returns_scan: bad ctx/P/N");
  if (P == 0 || N == 0) return 0;
  MJRL_REQUIRE(rew && off && returns, MJRL_E_ARG, "returns_scan: null pointer");
  segscan_kernel<false><<<P, SCAN_THREADS, 0, (cudaStream_t)stream>>>(
      rew, nullptr, off, nullptr, P, gamma, gamma, returns);
  return mjrl_check_launch("returns_scan");
}

int mjrl_gae_scan(mjrl_ctx* ctx, const double* rew, const double* baseline, const int32_t* off,
                  const uint8_t* terminated, int32_t P, int64_t N, double gamma, double lam,
                  double* adv, void* stream) {
  MJRL_REQUIRE(ctx && P >= 0 && N >= 0, MJRL_E_ARG, "gae_scan: bad ctx/P/N");
  if (P == 0 || N == 0) return 0;
  MJRL_REQUIRE(rew && baseline && off && terminated && adv, MJRL_E_ARG, "gae_scan: null pointer");
  segscan_kernel<true><<<P, SCAN_THREADS, 0, (cudaStream_t)stream>>>(
      rew, baseline, off, terminated, P, gamma, gamma * lam, adv);
  return mjrl_check_launch("gae_scan");
}

int mjrl_adv_plain(mjrl_ctx* ctx, const double* returns, const double* baseline, int64_t N,
                   double* adv, void* stream) {
  MJRL_REQUIRE(ctx && N >= 0, MJRL_E_ARG, "adv_plain: bad ctx/N");
  if (N == 0) return 0;
  MJRL_REQUIRE(returns && baseline && adv, MJRL_E_ARG, "adv_plain: null pointer");
  adv_plain_kernel<<<ew_grid(ctx, N, 256), 256, 0, (cudaStream_t)stream>>>(returns, baseline, N, adv);
  return mjrl_check_launch("adv_plain");
}

int mjrl_moments(mjrl_ctx* ctx, const double* x, int64_t N, double* mom, void* stream) {
  MJRL_REQUIRE(ctx && x && mom && N > 0, MJRL_E_ARG, "moments: bad argument");
  int grid = ew_grid(ctx, N, 256 * 8);
  double* part = (double*)mjrl_scratch(ctx, sizeof(double) * grid);
  if (!part) return MJRL_E_STATE;
  cudaStream_t s = (cudaStream_t)stream;
  moment_partial_kernel<false><<<grid, 256, 0, s>>>(x, N, mom, part);
  moment_final_kernel<false><<<1, 256, 0, s>>>(part, grid, N, mom);
  moment_partial_kernel<true><<<grid, 256, 0, s>>>(x, N, mom, part);
  moment_final_kernel<true><<<1, 256, 0, s>>>(part, grid, N, mom);
  return mjrl_check_launch("moments");
}

__global__ void moment_mean_kernel(double* __restrict__ mom) {
  mom[3] = mom[0] > 0.0 ? mom[1] / mom[0] : 0.0;
}

// The two passes of mjrl_moments as separate calls, so that a path-sharded batch can all-reduce
// {N, sum} between them and the centred squares afterwards (global moments of npg_cg.py:91).
int mjrl_moments_stage(mjrl_ctx* ctx, const double* x, int64_t N, double* mom, int32_t stage, void* stream) {
  MJRL_REQUIRE(ctx && x && mom && N > 0 && (stage == 0 || stage == 1), MJRL_E_ARG, "moments_stage: bad argument");
  int grid = ew_grid(ctx, N, 256 * 8);
  double* part = (double*)mjrl_scratch(ctx, sizeof(double) * grid);
  if (!part) return MJRL_E_STATE;
  cudaStream_t s = (cudaStream_t)stream;
  if (stage == 0) {
    moment_partial_kernel<false><<<grid, 256, 0, s>>>(x, N, mom, part);
    moment_final_kernel<false><<<1, 256, 0, s>>>(part, grid, N, mom);
  } else {
    moment_mean_kernel<<<1, 1, 0, s>>>(mom);   // mean of the (all-reduced) {N, sum}
    moment_partial_kernel<true><<<grid, 256, 0, s>>>(x, N, mom, part);
    moment_final_kernel<true><<<1, 256, 0, s>>>(part, grid, N, mom);
  }
  return mjrl_check_launch("moments_stage");
}

// ---- mjrl_moments without any allocation -------------------------------------------------------
// Scratch of the allocation-free variant: the per-CTA partial sums of the two passes live in a
// file-scope device array instead of the context's grow-on-demand block (no cudaMalloc on the
// call path).  Every GPU that loads the library holds these 32 KB until the process exits.
#define MJRL_NOALLOC_MAX_PARTS 4096
#define MJRL_NOALLOC_SLAB 2048       /* elements per CTA slab below the grid cap (ew_grid(N, 256 * 8)) */
#define MJRL_NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) double g_moments_partials[MJRL_NOALLOC_MAX_PARTS];
// Device address of g_moments_partials per device.  In host code the array's NAME is only a host-side
// stand-in for the symbol (passing it to a kernel hands over a host address), so the address is
// looked up once per device with cudaGetSymbolAddress while that device is current.
static double* g_moments_partials_dev[MJRL_NOALLOC_MAX_DEVICES];

int mjrl_moments_noalloc(mjrl_ctx* ctx, const double* x, int64_t N, double* mom, void* stream) {
  MJRL_REQUIRE(ctx && x && mom && N > 0, MJRL_E_ARG, "moments_noalloc: bad argument");
  const int64_t max_n = (int64_t)MJRL_NOALLOC_MAX_PARTS * MJRL_NOALLOC_SLAB;
  MJRL_REQUIRE(N <= max_n, MJRL_E_SHAPE,
               "moments_noalloc: N = %lld exceeds the static scratch (at most %lld elements = %d partial sums "
               "x %d elements per CTA); use mjrl_moments", (long long)N, (long long)max_n,
               MJRL_NOALLOC_MAX_PARTS, MJRL_NOALLOC_SLAB);
  const int grid = ew_grid(ctx, N, 256 * 8);   // the grid of mjrl_moments: same partials, same bits
  MJRL_REQUIRE(grid <= MJRL_NOALLOC_MAX_PARTS, MJRL_E_SHAPE,
               "moments_noalloc: %d CTAs exceed the %d static partial sums", grid, MJRL_NOALLOC_MAX_PARTS);
  int cur = -1;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) { mjrl_set_error("moments_noalloc: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
  MJRL_REQUIRE(cur == ctx->device, MJRL_E_STATE,
               "moments_noalloc: device %d is current, the context belongs to device %d", cur, ctx->device);
  MJRL_REQUIRE(cur >= 0 && cur < MJRL_NOALLOC_MAX_DEVICES, MJRL_E_STATE, "moments_noalloc: device index %d", cur);
  double* part = g_moments_partials_dev[cur];
  if (!part) {   // set-up, once per device, with that device current
    e = cudaGetSymbolAddress((void**)&part, g_moments_partials);
    if (e != cudaSuccess || !part) {
      mjrl_set_error("moments_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : MJRL_E_STATE;
    }
    g_moments_partials_dev[cur] = part;
  }
  // Nothing to clear: each pass writes partials [0, grid) before its final kernel reads exactly those.
  cudaStream_t s = (cudaStream_t)stream;
  moment_partial_kernel<false><<<grid, 256, 0, s>>>(x, N, mom, part);
  moment_final_kernel<false><<<1, 256, 0, s>>>(part, grid, N, mom);
  moment_partial_kernel<true><<<grid, 256, 0, s>>>(x, N, mom, part);
  moment_final_kernel<true><<<1, 256, 0, s>>>(part, grid, N, mom);
  return mjrl_check_launch("moments_noalloc");
}

int mjrl_normalize_inplace(mjrl_ctx* ctx, double* x, int64_t N, const double* mom, double eps,
                           void* stream) {
  MJRL_REQUIRE(ctx && x && mom && N > 0, MJRL_E_ARG, "normalize: bad argument");
  normalize_kernel<<<ew_grid(ctx, N, 256), 256, 0, (cudaStream_t)stream>>>(x, N, mom, eps);
  return mjrl_check_launch("normalize");
}

int mjrl_whiten_f32(mjrl_ctx* ctx, const double* adv, int64_t N, const double* mom, double eps,
                    double scale, float* w32, void* stream) {
  MJRL_REQUIRE(ctx && adv && mom && w32 && N > 0, MJRL_E_ARG, "whiten: bad argument");
  whiten_f32_kernel<<<ew_grid(ctx, N, 256), 256, 0, (cudaStream_t)stream>>>(adv, N, mom, eps, scale, w32);
  return mjrl_check_launch("whiten");
}

int mjrl_path_return_stats(mjrl_ctx* ctx, const double* rew, const int32_t* off, int32_t P,
                           double* stats4, void* stream) {
  MJRL_REQUIRE(ctx && rew && off && stats4 && P > 0, MJRL_E_ARG, "path_return_stats: bad argument");
  double* sums = (double*)mjrl_scratch(ctx, sizeof(double) * P);
  if (!sums) return MJRL_E_STATE;
  cudaStream_t s = (cudaStream_t)stream;
  path_sum_kernel<<<(P * 32 + 255) / 256, 256, 0, s>>>(rew, off, P, sums);
  path_stats_kernel<<<1, 256, 0, s>>>(sums, P, stats4);
  return mjrl_check_launch("path_return_stats");
}

int mjrl_cast_f64_to_f32(mjrl_ctx* ctx, const double* src, int64_t count, float* dst, void* stream) {
  MJRL_REQUIRE(ctx && count >= 0, MJRL_E_ARG, "cast: bad argument");
  if (count == 0) return 0;
  MJRL_REQUIRE(src && dst, MJRL_E_ARG, "cast: null pointer");
  cast_kernel<<<ew_grid(ctx, count, 256), 256, 0, (cudaStream_t)stream>>>(src, count, dst);
  return mjrl_check_launch("cast");
}

}  // extern "C"
