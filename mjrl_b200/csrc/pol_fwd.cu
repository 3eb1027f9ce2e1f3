// Forward pass of the policy on tcgen05 with TWO tiles in flight per SM: mean, log-likelihood, likelihood ratio and
// the surrogate / KL sums of the line search (mjrl/algos/batch_reinforce.py:57-75 over mjrl/policies/gaussian_mlp.py:
// 107-139,176-182), hidden 64 x 64.
//
// The fused kernels of fvp_tc.cu / pol_tc.cu keep ONE 96-timestep tile in flight because their weight-gradient operands
// fill shared memory; a tile then pays three tensor-core <-> row-thread round trips back to back (pol_tc_kernel<EVAL>:
// 8.7 k cycles per tile of which < 1 k are tensor-pipe time).  A forward pass needs no shared-memory planes at all:
// every A operand lives in tensor memory.  So here a tile has 128 timesteps (all four TMEM lane quadrants, one thread
// per row and all 64 hidden columns, no exchange between threads), and the CTA runs two such tiles at once on two sets
// of four warps with their own tensor-memory columns; while one set waits for its MMA the other computes tanh and the
// Gaussian head.  One elected lane of a ninth warp issues the MMAs of both sets, whichever is ready first.
//
//   z1 = [x~ 1] [W0 | b0]^T   MMA-1 (narrow; wide shapes read the z1 image of l0_fwd_kernel, fvp_tc_wide.cu)
//   h1 = tanh z1               -> tensor memory hi | lo (A of MMA-2)
//   z2 = h1 W1^T               MMA-2
//   h2 = tanh(z2 + b1); mu = (h2 W2^T + b2) out_scale + out_shift; LL; c = exp(LL - LL_old) advantage; KL
//
// (A gradient pass split into this forward kernel + the product kernel's backward half was measured and is NOT faster than
// the fused pol_tc_kernel<GRAD>: profiles/probes/r02_grad_split_negative_result.patch.)
#include <stdlib.h>

#include "common.cuh"
#include "fvp_tc.cuh"
#include "policy_tile.cuh"

namespace pol {
namespace pfw {

using namespace tc;

constexpr int TROWS = 128;                 // timesteps per tile
constexpr int NROW_THREADS = 256;          // two sets of four warps
constexpr int NTHREADS = 288;              // + the issuer warp
constexpr int KX = 16;
constexpr uint32_t WPLANE = H * 256, W0PLANE = H * 64;
constexpr uint32_t OFF_W1 = 0, OFF_W0 = 2 * WPLANE, OFF_VEC = OFF_W0 + 2 * W0PLANE;
template <int MJT>
struct Map {
  static constexpr int W2 = 0, B1 = MJT * H, B2 = B1 + H, SIG = B2 + MJT, LS = SIG + MJT, OS = LS + MJT, OSH = OS + MJT,
                       OLS = OSH + MJT, SOLD = OLS + MJT, ISHIFT = SOLD + MJT, IRCP = ISHIFT + KX;
  static constexpr uint32_t VEC_FLOATS = IRCP + KX;
  static constexpr uint32_t OFF_RED = (OFF_VEC + 4 * VEC_FLOATS + 15) & ~15u;     // end-of-kernel sums: [256][2] doubles
  static constexpr uint32_t OFF_BAR = OFF_RED + NROW_THREADS * 16;
  static constexpr uint32_t BYTES = OFF_BAR + 8 * 8;
  static constexpr uint32_t ALLOC = BYTES + 1024;
};
enum { B_E0 = 0, B_M1 = 2, B_E1 = 4, B_M2 = 6 };                          // + set index
constexpr uint32_t T_SET = 256, T_X = 0, T_D = 32, T_A1 = 96;            // per set: x~ hi | lo (32), z1 -> z2 (64), h1 hi | lo (128)

struct FwdArgs {
  Args a;
  const float* z1;        // wide: z1 image
};

__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

template <int MJT, bool WIDE>
__global__ void __launch_bounds__(NTHREADS, 1) pol_fwd_kernel(const FwdArgs fa) {
  using M = Map<MJT>;
  const Args& a = fa.a;
  extern __shared__ unsigned char smraw[];
  __shared__ uint32_t tslot;
  const uint32_t sraw = smem_u32(smraw);
  const uint32_t sb = (sraw + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sb - sraw);
  float* vecs = reinterpret_cast<float*>(sm + OFF_VEC);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + M::OFF_BAR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = a.n, m = a.m;
  const Flat F(n, m, H, H);
  const int ntiles = (int)((a.N + TROWS - 1) / TROWS);
  // tiles of this CTA: blockIdx.x, + gridDim.x, ...; the k-th of them belongs to set k & 1
  int ncta = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) ++ncta;

  for (uint32_t i = tid; i < OFF_VEC / 16; i += NTHREADS) reinterpret_cast<float4*>(sm)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (uint32_t i = tid; i < M::VEC_FLOATS; i += NTHREADS) vecs[i] = 0.f;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      tma::mbar_init(&bar[B_E0 + s], TROWS); tma::mbar_init(&bar[B_E1 + s], TROWS);
      tma::mbar_init(&bar[B_M1 + s], 1); tma::mbar_init(&bar[B_M2 + s], 1);
    }
    tma::mbar_init_fence();
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tslot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  __syncthreads();
  {
    const float* P = a.params;
    auto put = [&](uint32_t plane_hi, uint32_t plane_bytes, uint32_t off, float w) {
      *reinterpret_cast<float*>(sm + plane_hi + off) = w;
      *reinterpret_cast<float*>(sm + plane_hi + plane_bytes + off) = lo_of(w);
    };
    for (int i = tid; i < H * H; i += NTHREADS) put(OFF_W1, WPLANE, k128_off(H, i >> 6, i & 63), __ldg(P + F.oW1 + i));
    if (!WIDE) {
      for (int i = tid; i < H * n; i += NTHREADS) {
        const int o = i / n, k = i - o * n;
        put(OFF_W0, W0PLANE, k64_off(o, k), __ldg(P + F.oW0 + i));
      }
      for (int o = tid; o < H; o += NTHREADS) put(OFF_W0, W0PLANE, k64_off(o, n), __ldg(P + F.ob0 + o));
      for (int k = tid; k < n; k += NTHREADS) {
        vecs[M::ISHIFT + k] = a.in_shift ? __ldg(a.in_shift + k) : 0.f;
        vecs[M::IRCP + k] = 1.f / ((a.in_scale ? __ldg(a.in_scale + k) : 1.f) + 1e-8f);
      }
    }
    for (int o = tid; o < H; o += NTHREADS) vecs[M::B1 + o] = __ldg(P + F.ob1 + o);
    for (int i = tid; i < m * H; i += NTHREADS) vecs[M::W2 + i] = __ldg(P + F.oW2 + i);
    for (int j = tid; j < MJT; j += NTHREADS) {
      const bool in = j < m;
      const float ls = in ? __ldg(P + F.ols + j) : 0.f;
      vecs[M::B2 + j] = in ? __ldg(P + F.ob2 + j) : 0.f;
      vecs[M::LS + j] = ls;
      vecs[M::SIG + j] = expf(ls);
      vecs[M::OS + j] = (in && a.out_scale) ? __ldg(a.out_scale + j) : 1.f;
      vecs[M::OSH + j] = (in && a.out_shift) ? __ldg(a.out_shift + j) : 0.f;
      const float ols = (in && a.old_log_std) ? __ldg(a.old_log_std + j) : 0.f;
      vecs[M::OLS + j] = ols;
      vecs[M::SOLD + j] = expf(ols);
    }
  }
  fence_async();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tb = tslot;

  double surr_acc = 0.0, kl_acc = 0.0;

  if (warp < 8) {
    // ============ row threads: set = warp / 4, TMEM lane quadrant = warp % 4, one thread per timestep ==============
    const int set = warp >> 2, q = warp & 3;
    const int r = 32 * q + lane;
    const uint32_t tl = tb + ((uint32_t)(32 * q) << 16) + T_SET * set;
    const float llc = (float)(-0.5 * (double)m * 1.8378770664093453);
    float ox[KX];
    auto load_obs = [&](int k) {       // raw observation row of this set's k-th tile
      const int tile = blockIdx.x + (2 * k + set) * (int)gridDim.x;
      const long long t = (long long)tile * TROWS + r;
#pragma unroll
      for (int i = 0; i < KX; ++i) ox[i] = (!WIDE && tile < ntiles && t < a.N && i < n) ? __ldg(a.obs + t * n + i) : 0.f;
    };
    if (!WIDE) load_obs(0);
    for (int k = 0; 2 * k + set < ncta; ++k) {
      const int tile = blockIdx.x + (2 * k + set) * (int)gridDim.x;
      const uint32_t ph = k & 1;
      const long long t = (long long)tile * TROWS + r;
      const bool valid = t < a.N;
      float h[64];
      // ---- P0 ----
      if (!WIDE) {
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int i = 0; i < KX; ++i) {
          float v = 0.f;
          if (valid && i < n) v = (ox[i] - vecs[M::ISHIFT + i]) * vecs[M::IRCP + i];
          if (valid && i == n) v = 1.f;
          hi[i] = __float_as_uint(v); lo[i] = __float_as_uint(lo_of(v));
        }
        tst16(tl + T_X, hi);
        tst16(tl + T_X + 16, lo);
        wait_st();
        fence_before();
        mbar_arrive(&bar[B_E0 + set]);
        load_obs(k + 1);
        tma::mbar_wait(&bar[B_M1 + set], ph);
        fence_after();
        uint32_t d[32];
        tld32w(tl + T_D, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = __uint_as_float(d[i]);
        tld32w(tl + T_D + 32, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[32 + i] = __uint_as_float(d[i]);
      } else {
        // z1 image: per 96-timestep tile [16 column chunks][96 rows][4 floats] (l0_fwd_kernel)
        const long long t96 = t / ROWS;
        const int r96 = (int)(t - t96 * ROWS);
        const float4* zp = reinterpret_cast<const float4*>(fa.z1) + (size_t)t96 * (ROWS * H / 4) + r96;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const float4 v = valid ? __ldg(zp + (size_t)c * ROWS) : make_float4(0.f, 0.f, 0.f, 0.f);
          h[4 * c] = v.x; h[4 * c + 1] = v.y; h[4 * c + 2] = v.z; h[4 * c + 3] = v.w;
        }
      }
      // ---- P1: h1 = tanh z1 -> A of MMA-2 ----
      {
        uint32_t u32[32];
#pragma unroll
        for (int i = 0; i < 64; ++i) h[i] = tanh_fast(h[i]);
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 32) {
#pragma unroll
          for (int i = 0; i < 32; ++i) u32[i] = __float_as_uint(h[c0 + i]);
          tst32(tl + T_A1 + c0, u32);
#pragma unroll
          for (int i = 0; i < 32; ++i) u32[i] = __float_as_uint(lo_of(h[c0 + i]));
          tst32(tl + T_A1 + 64 + c0, u32);
        }
        wait_st();
        fence_before();
        mbar_arrive(&bar[B_E1 + set]);
      }
      // ... while MMA-2 runs: this row's scalars
      float av[MJT], om[MJT];
      float wv = 0.f, oll = 0.f;
      if (valid) {
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          av[j] = j < m ? __ldg(a.act + t * m + j) : 0.f;
          om[j] = j < m ? __ldg(a.old_mean + t * m + j) : 0.f;
        }
        wv = a.w ? __ldg(a.w + t) : 0.f;
        oll = a.old_ll ? __ldg(a.old_ll + t) : 0.f;
      }
      // ---- P2: h2 = tanh(z2 + b1); head ----
      tma::mbar_wait(&bar[B_M2 + set], ph);
      fence_after();
      {
        uint32_t d[32];
        tld32w(tl + T_D, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = tanh_fast(__uint_as_float(d[i]) + vecs[M::B1 + i]);
        tld32w(tl + T_D + 32, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[32 + i] = tanh_fast(__uint_as_float(d[i]) + vecs[M::B1 + 32 + i]);
      }
      fence_before();
      float ss = 0.f, sum_ls = 0.f, kl = 0.f;
#pragma unroll
      for (int j = 0; j < MJT; ++j) {
        const float4* w2 = reinterpret_cast<const float4*>(vecs + M::W2 + j * H);
        float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
        for (int i4 = 0; i4 < 16; ++i4) {
          const float4 w = w2[i4];
          a0 = __ffma2_rn(make_float2(h[4 * i4], h[4 * i4 + 1]), make_float2(w.x, w.y), a0);
          a1 = __ffma2_rn(make_float2(h[4 * i4 + 2], h[4 * i4 + 3]), make_float2(w.z, w.w), a1);
        }
        const float dot = (a0.x + a0.y) + (a1.x + a1.y);
        const float mu = (dot + vecs[M::B2 + j]) * vecs[M::OS + j] + vecs[M::OSH + j];
        if (j < m && valid) {
          if (a.mean_out) a.mean_out[t * m + j] = mu;
          const float sg = vecs[M::SIG + j];
          const float z = (av[j] - mu) / sg;
          ss = fmaf(z, z, ss);
          sum_ls += vecs[M::LS + j];
          // gaussian_mlp.py:137-139, KL(old || new) per dimension
          const float so = vecs[M::SOLD + j];
          const float nr = (om[j] - mu) * (om[j] - mu) + so * so - sg * sg;
          const float dr = 2.f * sg * sg + 1e-8f;
          kl += nr / dr + vecs[M::LS + j] - vecs[M::OLS + j];
        }
      }
      const float ll = (-0.5f * ss + -sum_ls) + llc;
      if (a.ll_out && valid) a.ll_out[t] = ll;
      float lr = 1.f;
      if (a.old_ll && valid) lr = expf(ll - oll);
      const float c = valid ? lr * wv : 0.f;
      surr_acc += (double)c;
      kl_acc += valid ? (double)kl : 0.0;
    }
  } else {
    // ============ MMA issuer: serves whichever set is ready ==========================================================
    const uint32_t ID_FWD = idesc_tf32(128, 64, 0, 0);
    auto wdesc = [&](uint32_t off, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * (H * 128u) + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    auto w0desc = [&](uint32_t off, int ks) { return make_desc(sb + off + (uint32_t)ks * 32u, 16u, 512u, 4u); };
    int k[2] = {0, 0};                    // tiles done per set
    int stage[2] = {WIDE ? 1 : 0, WIDE ? 1 : 0};     // 0: waiting for E0 (x~ staged), 1: waiting for E1 (h1 staged)
    int cnt[2] = {(ncta + 1) / 2, ncta / 2};
    while (k[0] < cnt[0] || k[1] < cnt[1]) {
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if (k[s] >= cnt[s]) continue;
        const uint32_t ph = k[s] & 1;
        const uint32_t ts = tb + T_SET * s;
        if (stage[s] == 0) {
          if (!mbar_test(&bar[B_E0 + s], ph)) continue;
          fence_after();
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
              mma_ts(ts + T_D, ts + T_X + 8 * ks, w0desc(OFF_W0, ks), ID_FWD, ks > 0);
              mma_ts(ts + T_D, ts + T_X + 16 + 8 * ks, w0desc(OFF_W0, ks), ID_FWD, 1);
              mma_ts(ts + T_D, ts + T_X + 8 * ks, w0desc(OFF_W0 + W0PLANE, ks), ID_FWD, 1);
            }
            commit(&bar[B_M1 + s]);
          }
          __syncwarp();
          stage[s] = 1;
        } else {
          if (!mbar_test(&bar[B_E1 + s], ph)) continue;
          fence_after();
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              mma_ts(ts + T_D, ts + T_A1 + 8 * ks, wdesc(OFF_W1, ks), ID_FWD, ks > 0);
              mma_ts(ts + T_D, ts + T_A1 + 64 + 8 * ks, wdesc(OFF_W1, ks), ID_FWD, 1);
              mma_ts(ts + T_D, ts + T_A1 + 8 * ks, wdesc(OFF_W1 + WPLANE, ks), ID_FWD, 1);
            }
            commit(&bar[B_M2 + s]);
          }
          __syncwarp();
          stage[s] = WIDE ? 1 : 0;
          ++k[s];
        }
      }
    }
  }

  // ---- sums of the CTA ---------------------------------------------------------------------------------------------
  double* dred = reinterpret_cast<double*>(sm + M::OFF_RED);
  if (warp < 8) { dred[2 * tid] = surr_acc; dred[2 * tid + 1] = kl_acc; }
  fence_before();
  __syncthreads();
  if (tid == 0) {
    double s0 = 0.0, s1 = 0.0;
    for (int i = 0; i < NROW_THREADS; ++i) { s0 += dred[2 * i]; s1 += dred[2 * i + 1]; }
    a.spartials[2 * blockIdx.x] = s0;
    a.spartials[2 * blockIdx.x + 1] = s1;
  }
  __syncthreads();
  if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tb) : "memory");
}

template <int MJT, bool WIDE>
static int launch_one(const FwdArgs& fa, int grid, cudaStream_t s) {
  auto kern = pol_fwd_kernel<MJT, WIDE>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Map<MJT>::ALLOC);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute(pol_fwd): %s", cudaGetErrorString(e)); return (int)e; }
    attr = true;
  }
  kern<<<grid, NTHREADS, Map<MJT>::ALLOC, s>>>(fa);
  return mjrl_check_launch("pol_fwd_kernel");
}

}  // namespace pfw

// a.spartials[grid][2] = surrogate, KL sums of the batch under a.params against the cached old policy
int launch_pol_fwd(mjrl_ctx* ctx, Args& a, const TcwRegions* wide, int grid, cudaStream_t s) {
  a.spartials = (double*)mjrl_scratch2(ctx, sizeof(double) * 2 * grid);
  if (!a.spartials) return MJRL_E_STATE;
  pfw::FwdArgs fa;
  fa.a = a; fa.z1 = wide ? wide->z1 : nullptr;
  if (wide) {
    int rc = launch_tcw_l0_fwd(ctx, *wide, a.params, a.N, a.n, a.m, grid, nullptr, s);
    if (rc) return rc;
    if (a.m <= 8) return pfw::launch_one<8, true>(fa, grid, s);
    if (a.m <= 16) return pfw::launch_one<16, true>(fa, grid, s);
    return pfw::launch_one<24, true>(fa, grid, s);
  }
  if (a.m <= 4) return pfw::launch_one<4, false>(fa, grid, s);
  return pfw::launch_one<8, false>(fa, grid, s);
}

}  // namespace pol
