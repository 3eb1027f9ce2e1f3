// Policy gradient (flat_vpg, mjrl/algos/batch_reinforce.py:51-55) and surrogate / KL evaluation
// (batch_reinforce.py:57-75 over mjrl/policies/gaussian_mlp.py:107-139,176-182) on the 5th-generation tensor cores, for
// the two-hidden-layer tanh policy with 64 x 64 hidden units.  Same tile, planes, precision scheme (kind::tf32, hi + lo
// planes, 1e-6) and barrier pipeline as the Fisher-vector product of fvp_tc.cu / fvp_tc_wide.cu, with the primal pass in
// place of the tangent pass:
//
//   z1 = x~ W0^T + b0                  MMA-1 (narrow: A = x~ in TMEM)  |  wide: z1 image of pass A (fvp_tc_wide.cu)
//   h1 = tanh z1                        P1     TMEM -> registers -> TMEM (A of MMA-2), planes X | W, plane cache in HBM
//   z2 = h1 W1^T                        MMA-2
//   h2 = tanh(z2 + b1); mu; LL; LR      P2     head on the CUDA cores; the two threads of a row meet in the u^T plane
//   evaluation: surrogate and KL sums, done.            gradient: u = c dLL/dmu, d log_std, then as the product:
//   gz2 = (u W2)(1 - h2^2); gh1 = gz2 W1 (MMA-5); gz1 = gh1 (1 - h1^2); gW2, gW1, gW0 weight-gradient MMAs
//
// The gradient pass leaves h1 | h2 of every tile in the plane order the products of the following CG solve read.
// The evaluation started as the forward half of this kernel (one tile in flight: 0.30 ms at the Hopper shape) and moved to the
// two-tiles-in-flight forward kernel of pol_fwd.cu (0.19 ms); this file is the gradient pass only.
#include <stdlib.h>

#include "common.cuh"
#include "fvp_tc.cuh"
#include "policy_tile.cuh"

namespace pol {
namespace ptc {

using namespace tc;

constexpr int NTHREADS = 256;
constexpr int ETHREADS = 2 * ROWS;
constexpr int KX = 16;
constexpr uint32_t WPLANE = H * 256, W0PLANE = H * 64;
constexpr uint32_t OFF_X = 0, OFF_W = PLANE, OFF_Y = 2 * PLANE, OFF_Z = 3 * PLANE;
constexpr uint32_t OFF_W1 = 4 * PLANE, OFF_W1T = OFF_W1 + 2 * WPLANE, OFF_W0 = OFF_W1T + 2 * WPLANE;
constexpr uint32_t OFF_XT = OFF_W0 + 2 * W0PLANE;     // [x~ hi (16) | x~ lo (16)][96 t] K-major
constexpr uint32_t OFF_UT = OFF_XT + 3 * 32 * 128;
template <int MJT>
struct Map {
  static constexpr int UR = MJT <= 8 ? 16 : 2 * MJT;     // rows of the u^T plane: u hi at row j, u lo at row UR / 2 + j
  static constexpr int LOROW = UR / 2;
  static constexpr uint32_t OFF_VEC = OFF_UT + 3 * UR * 128;
  static constexpr int W2 = 0, B1 = MJT * H, B2 = B1 + H, SIG = B2 + MJT, LS = SIG + MJT, OS = LS + MJT, OSH = OS + MJT,
                       ISHIFT = OSH + MJT, IRCP = ISHIFT + KX;
  static constexpr uint32_t VEC_FLOATS = IRCP + KX;
  static constexpr uint32_t OFF_BAR = (OFF_VEC + 4 * VEC_FLOATS + 15) & ~15u;
  static constexpr uint32_t BYTES = OFF_BAR + 8 * 32;
  static constexpr uint32_t ALLOC = BYTES + 1024;
};
enum { B_E0 = 0, B_E1, B_E2A, B_E2B, B_M1, B_M2B, B_M5, B_GW0, B_GW2 = 8, B_E2C = 11, B_GW1 = 14, B_E3 = 17, B_W3 = 20, B_COUNT = 23 };
constexpr uint32_t T_D1 = 0, T_LO1 = 64, T_D2 = 128, T_A1 = 192, T_ACC1 = 320, T_ACC0 = 448;

struct PolArgs {
  Args a;
  float* planes;      // gradient: tile-ordered h1 | h2 cache to fill (may be null)
  const float* z1;    // wide: z1 image ([x~ 1] [W0 | b0]^T, pass A)
  float* gz1;         // wide gradient: GZ1 image for pass C
  long long* dbg;     // debug: per-phase cycle sums of CTA 0, thread 0 (MJRL_B200_TC_PHASES=1), else null
};

// one row's 32 columns -> a plane in HBM (32-byte atoms in natural order)
__device__ __forceinline__ void gstore_row32(float* plane, int r, int blk, const float (&v)[32]) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float4* p = reinterpret_cast<float4*>(plane + mn_atom(r, blk * 4 + q) / 4);
    p[0] = make_float4(v[8 * q + 0], v[8 * q + 1], v[8 * q + 2], v[8 * q + 3]);
    p[1] = make_float4(v[8 * q + 4], v[8 * q + 5], v[8 * q + 6], v[8 * q + 7]);
  }
}

template <int MJT, bool WIDE>
__global__ void __launch_bounds__(NTHREADS, 1) pol_tc_kernel(const PolArgs pa) {
  using M = Map<MJT>;
  constexpr int UR = M::UR, LOROW = M::LOROW;
  constexpr uint32_t T_ACC2 = WIDE ? 448u : 480u;
  const Args& a = pa.a;
  extern __shared__ unsigned char smraw[];
  __shared__ uint32_t tslot;
  const uint32_t sraw = smem_u32(smraw);
  const uint32_t sb = (sraw + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sb - sraw);
  float* vecs = reinterpret_cast<float*>(sm + M::OFF_VEC);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + M::OFF_BAR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = a.n, m = a.m;
  const Flat F(n, m, H, H);
  const int ntiles = (int)((a.N + ROWS - 1) / ROWS);

  for (uint32_t i = tid; i < (M::OFF_VEC - OFF_W1) / 16; i += NTHREADS)
    reinterpret_cast<float4*>(sm + OFF_W1)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (uint32_t i = tid; i < M::VEC_FLOATS; i += NTHREADS) vecs[i] = 0.f;
  if (tid == 0) {
    for (int b = B_E0; b <= B_E2B; ++b) tma::mbar_init(&bar[b], ETHREADS);
    for (int b = B_M1; b <= B_GW0; ++b) tma::mbar_init(&bar[b], 1);
    for (int qq = 0; qq < 3; ++qq) {
      tma::mbar_init(&bar[B_GW2 + qq], 1); tma::mbar_init(&bar[B_GW1 + qq], 1);
      tma::mbar_init(&bar[B_E2C + qq], 64); tma::mbar_init(&bar[B_E3 + qq], 64);
      tma::mbar_init(&bar[B_W3 + qq], 32);
    }
    tma::mbar_init_fence();
  }
  if (warp == 7) {
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
    for (int i = tid; i < H * H; i += NTHREADS) {
      const int o = i >> 6, k = i & 63;
      const float w = __ldg(P + F.oW1 + i);
      put(OFF_W1, WPLANE, k128_off(H, o, k), w);                 // B[N = o][K = i]: h1 W1^T
      put(OFF_W1T, WPLANE, k128_off(H, k, o), w);                // B[N = i][K = o]: gz2 W1
    }
    if (!WIDE) {
      for (int i = tid; i < H * n; i += NTHREADS) {
        const int o = i / n, k = i - o * n;
        put(OFF_W0, W0PLANE, k64_off(o, k), __ldg(P + F.oW0 + i));
      }
      for (int o = tid; o < H; o += NTHREADS) put(OFF_W0, W0PLANE, k64_off(o, n), __ldg(P + F.ob0 + o));   // b0 on the ones column
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
    }
  }
  fence_async();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tb = tslot;

  unsigned char* pX = sm + OFF_X; unsigned char* pW = sm + OFF_W;
  unsigned char* pY = sm + OFF_Y; unsigned char* pZ = sm + OFF_Z;
  unsigned char* pU = sm + OFF_UT;
  double surr_acc = 0.0;
  float gls[MJT];
#pragma unroll
  for (int j = 0; j < MJT; ++j) gls[j] = 0.f;

  if ((warp & 3) != 3) {
    // ============ row threads: warps q and q + 4 share TMEM lane quadrant q, 32 of the 64 hidden columns each =========
    const int q = warp & 3, half = warp >> 2, blk = half;
    const int r = 32 * q + lane;
    const uint32_t tl = tb + ((uint32_t)(32 * q) << 16);
    const uint32_t pair_bar = 1 + q;
    auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory"); };
    float ox[KX];
    auto prefetch_obs = [&](int tile) {
      const long long t = (long long)tile * ROWS + r;
#pragma unroll
      for (int k = 0; k < KX; ++k) ox[k] = (!WIDE && half == 1 && tile < ntiles && t < a.N && k < n) ? __ldg(a.obs + t * n + k) : 0.f;
    };
    if (!WIDE) prefetch_obs(blockIdx.x);
    const float llc = (float)(-0.5 * (double)m * 1.8378770664093453);
    long long ph_sum[12], tprev = 0;
#pragma unroll
    for (int i = 0; i < 12; ++i) ph_sum[i] = 0;
    const bool stamp_on = pa.dbg != nullptr && tid == 0 && blockIdx.x == 0;
    auto stamp = [&](int i) { if (stamp_on) { const long long c = clock64(); ph_sum[i] += c - tprev; tprev = c; } };
    if (stamp_on) tprev = clock64();
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const long long t = (long long)tile * ROWS + r;
      const bool valid = t < a.N;
      // ---- P0: narrow: x~ -> tensor memory (A of MMA-1); wide: z1 of this thread's 32 columns is requested --------
      float xs[KX];
      float4 zr[8];
      if (!WIDE) {
        if (half == 1) {
          uint32_t hi[16], lo[16];
#pragma unroll
          for (int k = 0; k < KX; ++k) {
            float v = 0.f;
            if (valid && k < n) v = (ox[k] - vecs[M::ISHIFT + k]) * vecs[M::IRCP + k];
            if (valid && k == n) v = 1.f;
            xs[k] = v;
            hi[k] = __float_as_uint(v); lo[k] = __float_as_uint(lo_of(v));
          }
          tst16(tl + T_LO1, hi);
          tst16(tl + T_LO1 + 16, lo);
        }
        wait_st();
        fence_before();
        mbar_arrive(&bar[B_E0]);
        stamp(0);
      } else {
        const float4* zp = reinterpret_cast<const float4*>(pa.z1) + (size_t)tile * (ROWS * H / 4) + (size_t)(8 * blk) * ROWS + r;
#pragma unroll
        for (int j = 0; j < 8; ++j) zr[j] = __ldg(zp + (size_t)j * ROWS);
      }
      // ---- P1: h1 = tanh z1 -> tensor memory (A of MMA-2), planes X | W, plane cache ----------------------------
      float s1[32];
      {
        float h[32];
        uint32_t u32[32];
        if (!WIDE) {
          tma::mbar_wait(&bar[B_M1], ph);
          stamp(1);
          fence_after();
          tld32w(tl + T_D1 + 32 * blk, u32);
#pragma unroll
          for (int i = 0; i < 32; ++i) h[i] = tanh_fast(__uint_as_float(u32[i]));
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            h[4 * j] = tanh_fast(zr[j].x); h[4 * j + 1] = tanh_fast(zr[j].y);
            h[4 * j + 2] = tanh_fast(zr[j].z); h[4 * j + 3] = tanh_fast(zr[j].w);
          }
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) { u32[i] = __float_as_uint(h[i]); s1[i] = fmaf(-h[i], h[i], 1.f); }
        tst32(tl + T_A1 + 32 * blk, u32);
        uint32_t l32[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) l32[i] = __float_as_uint(lo_of(h[i]));
        tst32(tl + T_A1 + 64 + 32 * blk, l32);
        wait_st();
        fence_before();
        mbar_arrive(&bar[B_E1]);
        stamp(2);
        // ... while MMA-2 runs: the planes gW1 reads much later (every thread fences before its B_E2A / B_E2C arrival)
        store_row32(pX, r, blk, h);
        if (pa.planes) gstore_row32(pa.planes + (size_t)tile * (2 * PLANE / 4), r, blk, h);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = __uint_as_float(l32[i]);
        store_row32(pW, r, blk, h);
      }

      // ---- P2: h2 = tanh(z2 + b1); head ------------------------------------------------------------------------------
      float av[MJT];
      float wv = 0.f, oll = 0.f;
      if (half == 0 && valid) {      // this row's scalars, requested while MMA-2 runs
#pragma unroll
        for (int j = 0; j < MJT; ++j) av[j] = j < m ? __ldg(a.act + t * m + j) : 0.f;
        wv = a.w ? __ldg(a.w + t) : 0.f;
        oll = a.old_ll ? __ldg(a.old_ll + t) : 0.f;
      }
      float s2[32];
      float u[MJT];
      {
        float h[32];
        uint32_t d[32];
        stamp(3);
        tma::mbar_wait(&bar[B_M2B], ph);
        stamp(4);
        fence_after();
        tld32w(tl + T_D2 + 32 * blk, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = tanh_fast(__uint_as_float(d[i]) + vecs[M::B1 + 32 * blk + i]);
        float2 hv[16];
#pragma unroll
        for (int i2 = 0; i2 < 16; ++i2) hv[i2] = make_float2(h[2 * i2], h[2 * i2 + 1]);
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          const float4* w2 = reinterpret_cast<const float4*>(vecs + M::W2 + j * H + 32 * blk);
          float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 w = w2[i4];
            a0 = __ffma2_rn(hv[2 * i4], make_float2(w.x, w.y), a0);
            a1 = __ffma2_rn(hv[2 * i4 + 1], make_float2(w.z, w.w), a1);
          }
          u[j] = (a0.x + a0.y) + (a1.x + a1.y);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) s2[i] = fmaf(-h[i], h[i], 1.f);
        if (it > 0) tma::mbar_wait(&bar[B_GW0], (it - 1) & 1);     // the previous tile's gz1 planes have been read
        store_row32(pY, r, blk, h);
        if (pa.planes) gstore_row32(pa.planes + (size_t)tile * (2 * PLANE / 4) + PLANE / 4, r, blk, h);
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = lo_of(h[i]);
        store_row32(pZ, r, blk, h);
      }
      stamp(5);
      // the two threads of a row meet in the u^T plane (row j, column = timestep): half 1 -> half 0 (-> half 1)
      if (half == 1) {
#pragma unroll
        for (int j = 0; j < MJT; ++j) *reinterpret_cast<float*>(pU + k128_off(UR, j, r)) = u[j];
      }
      pair_sync();
      if (half == 0) {
        float ss = 0.f, sum_ls = 0.f;
        float zz[MJT];
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          const float dot = u[j] + *reinterpret_cast<const float*>(pU + k128_off(UR, j, r));
          const float mu = (dot + vecs[M::B2 + j]) * vecs[M::OS + j] + vecs[M::OSH + j];
          if (j < m && valid) {
            if (a.mean_out) a.mean_out[t * m + j] = mu;
            const float sg = vecs[M::SIG + j];
            const float z = (av[j] - mu) / sg;
            ss = fmaf(z, z, ss);
            sum_ls += vecs[M::LS + j];
            zz[j] = z;
          } else {
            zz[j] = 0.f;
          }
        }
        // gaussian_mlp.py:107-109
        const float ll = (-0.5f * ss + -sum_ls) + llc;
        if (a.ll_out && valid) a.ll_out[t] = ll;
        float lr = 1.f;
        if (a.old_ll && valid) lr = expf(ll - oll);
        const float c = valid ? lr * wv : 0.f;
        surr_acc += (double)c;
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          const bool in = j < m && valid;
          u[j] = in ? c * (zz[j] / vecs[M::SIG + j]) * vecs[M::OS + j] : 0.f;     // dL / dy_j
          gls[j] += in ? c * (zz[j] * zz[j] - 1.f) : 0.f;                         // dL / dlog_std_j
          *reinterpret_cast<float*>(pU + k128_off(UR, j, r)) = u[j];
        }
      }
      pair_sync();
      if (half == 1) {
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          u[j] = *reinterpret_cast<const float*>(pU + k128_off(UR, j, r));
          *reinterpret_cast<float*>(pU + k128_off(UR, LOROW + j, r)) = lo_of(u[j]);
        }
      }
      fence_async();
      mbar_arrive(&bar[B_E2A]);
      stamp(6);
      float gz[32];
      {
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          float2 g0 = make_float2(0.f, 0.f), g1 = g0;
#pragma unroll
          for (int j = 0; j < MJT; ++j) {
            const float4 w = *reinterpret_cast<const float4*>(vecs + M::W2 + j * H + 32 * blk + 4 * i4);
            const float2 uu = make_float2(u[j], u[j]);
            g0 = __ffma2_rn(uu, make_float2(w.x, w.y), g0);
            g1 = __ffma2_rn(uu, make_float2(w.z, w.w), g1);
          }
          gz[4 * i4 + 0] = g0.x * s2[4 * i4 + 0]; gz[4 * i4 + 1] = g0.y * s2[4 * i4 + 1];
          gz[4 * i4 + 2] = g1.x * s2[4 * i4 + 2]; gz[4 * i4 + 3] = g1.y * s2[4 * i4 + 3];
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) { hi[i] = __float_as_uint(gz[i]); lo[i] = __float_as_uint(lo_of(gz[i])); }
        tst32(tl + T_A1 + 32 * blk, hi);
        tst32(tl + T_A1 + 64 + 32 * blk, lo);
      }
      wait_st();
      fence_before();
      mbar_arrive(&bar[B_E2B]);
      stamp(7);
      tma::mbar_wait(&bar[B_GW2 + q], ph);              // gW2 has read this quadrant's rows of h2 hi | lo: they become gz2
      {
        store_row32(pY, r, blk, gz);
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = lo_of(gz[i]);
        store_row32(pZ, r, blk, v);
      }
      fence_async();
      mbar_arrive(&bar[B_E2C + q]);
      stamp(8);
      // ---- P3: gz1 = gh1 (1 - h1^2) ------------------------------------------------------------------------------------
      {
        uint32_t d[32];
        tma::mbar_wait(&bar[B_M5], ph);
        stamp(9);
        fence_after();
        tld32w(tl + T_D1 + 32 * blk, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) gz[i] = valid ? __uint_as_float(d[i]) * s1[i] : 0.f;
      }
      fence_before();
      if (!WIDE && half == 1) {     // the x~^T planes were last read by gW0 of the previous tile (B_GW0, waited in P2)
#pragma unroll
        for (int k = 0; k < KX; ++k) {
          *reinterpret_cast<float*>(sm + OFF_XT + k128_off(32, k, r)) = xs[k];
          *reinterpret_cast<float*>(sm + OFF_XT + k128_off(32, 16 + k, r)) = lo_of(xs[k]);
        }
      }
      tma::mbar_wait(&bar[B_GW1 + q], ph);              // gW1 has read this quadrant's rows of gz2
      tma::mbar_wait(&bar[B_W3 + q], ph);               // warp 3 has summed them for gb1
      stamp(10);
      store_row32(pY, r, blk, gz);
      if (!WIDE) {
        float gl[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) gl[i] = lo_of(gz[i]);
        store_row32(pZ, r, blk, gl);
      }
      fence_async();
      mbar_arrive(&bar[B_E3 + q]);
      if (!WIDE) prefetch_obs(tile + gridDim.x);
      stamp(11);
    }
    if (stamp_on) { for (int i = 0; i < 12; ++i) pa.dbg[i] = ph_sum[i]; pa.dbg[12] = it; }
  } else if (warp == 3) {
    // ============ bias gradients of layers 1 and 2 ================================================================
    double gb1a = 0.0, gb1b = 0.0, gb2 = 0.0;
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      tma::mbar_wait(&bar[B_E2A], ph);            // u of the tile is complete; rewritten only after this warp's B_W3 arrivals
      if (lane < MJT) {
        float f[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
        for (int rr = 0; rr < ROWS; rr += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) f[k] += *reinterpret_cast<const float*>(pU + k128_off(UR, lane, rr + k));
        }
        gb2 += (double)((f[0] + f[1]) + (f[2] + f[3]));
      }
      const int c0 = 2 * lane, qa = c0 >> 3;
      float sx = 0.f, sy = 0.f;
#pragma unroll 1
      for (int qq = 0; qq < 3; ++qq) {
        tma::mbar_wait(&bar[B_E2C + qq], ph);
        float2 acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = make_float2(0.f, 0.f);
#pragma unroll
        for (int rr = 0; rr < 32; rr += 8) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float2 v = *reinterpret_cast<const float2*>(pY + mn_atom(32 * qq + rr + k, qa) + (c0 & 7) * 4);
            acc[k].x += v.x; acc[k].y += v.y;
          }
        }
        mbar_arrive(&bar[B_W3 + qq]);
        sx += ((acc[0].x + acc[1].x) + (acc[2].x + acc[3].x)) + ((acc[4].x + acc[5].x) + (acc[6].x + acc[7].x));
        sy += ((acc[0].y + acc[1].y) + (acc[2].y + acc[3].y)) + ((acc[4].y + acc[5].y) + (acc[6].y + acc[7].y));
      }
      gb1a += (double)sx; gb1b += (double)sy;
    }
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    out[F.ob1 + 2 * lane] = (float)gb1a;
    out[F.ob1 + 2 * lane + 1] = (float)gb1b;
    if (lane < m) out[F.ob2 + lane] = (float)gb2;
  } else {
    // ============ MMA issuer (one elected lane of warp 7) ==========================================================
    const uint32_t ID_FWD = idesc_tf32(128, 64, 0, 0);
    const uint32_t ID_GW1 = idesc_tf32(128, 128, 1, 1);
    const uint32_t ID_GW0 = idesc_tf32(128, 32, 1, 0);
    const uint32_t ID_GW2 = idesc_tf32(128, UR, 1, 0);
    auto wdesc = [&](uint32_t off, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * (H * 128u) + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    auto w0desc = [&](uint32_t off, int ks) { return make_desc(sb + off + (uint32_t)ks * 32u, 16u, 512u, 4u); };
    auto mndesc = [&](uint32_t off, int ks) { return make_desc(sb + off + (uint32_t)ks * 1024u, BLK, 512u, 1u); };
    auto ktdesc = [&](uint32_t off, int R, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * (uint32_t)R * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const uint32_t first = it == 0 ? 0u : 1u;
      if (!WIDE) {
        tma::mbar_wait(&bar[B_E0], ph);
        fence_after();
        if (elect_one()) {
          // MMA-1: z1 = [x~ hi | lo] [W0 | b0]^T     (K = 16)
#pragma unroll
          for (int ks = 0; ks < 2; ++ks) {
            mma_ts(tb + T_D1, tb + T_LO1 + 8 * ks, w0desc(OFF_W0, ks), ID_FWD, ks > 0);
            mma_ts(tb + T_D1, tb + T_LO1 + 16 + 8 * ks, w0desc(OFF_W0, ks), ID_FWD, 1);
            mma_ts(tb + T_D1, tb + T_LO1 + 8 * ks, w0desc(OFF_W0 + W0PLANE, ks), ID_FWD, 1);
          }
          commit(&bar[B_M1]);
        }
        __syncwarp();
      }
      tma::mbar_wait(&bar[B_E1], ph);
      fence_after();
      if (elect_one()) {
        // MMA-2: z2 = h1 W1^T
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          mma_ts(tb + T_D2, tb + T_A1 + 8 * ks, wdesc(OFF_W1, ks), ID_FWD, ks > 0);
          mma_ts(tb + T_D2, tb + T_A1 + 64 + 8 * ks, wdesc(OFF_W1, ks), ID_FWD, 1);
          mma_ts(tb + T_D2, tb + T_A1 + 8 * ks, wdesc(OFF_W1 + WPLANE, ks), ID_FWD, 1);
        }
        commit(&bar[B_M2B]);
      }
      __syncwarp();
      tma::mbar_wait(&bar[B_E2A], ph);
      fence_after();
      if (elect_one()) {
        // gW2 (+)= [h2 hi | lo]^T [u hi | lo], one part per lane quadrant
#pragma unroll
        for (int qq = 0; qq < 3; ++qq) {
#pragma unroll
          for (int ks = 4 * qq; ks < 4 * qq + 4; ++ks)
            mma_ss(tb + T_ACC2, mndesc(OFF_Y, ks), ktdesc(OFF_UT, UR, ks), ID_GW2, first | (ks > 0));
          commit(&bar[B_GW2 + qq]);
        }
      }
      __syncwarp();
      tma::mbar_wait(&bar[B_E2B], ph);
      fence_after();
      if (elect_one()) {
        // MMA-5: gh1 = gz2 W1
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          mma_ts(tb + T_D1, tb + T_A1 + 8 * ks, wdesc(OFF_W1T, ks), ID_FWD, ks > 0);
          mma_ts(tb + T_D1, tb + T_A1 + 64 + 8 * ks, wdesc(OFF_W1T, ks), ID_FWD, 1);
          mma_ts(tb + T_D1, tb + T_A1 + 8 * ks, wdesc(OFF_W1T + WPLANE, ks), ID_FWD, 1);
        }
        commit(&bar[B_M5]);
      }
      __syncwarp();
#pragma unroll 1
      for (int qq = 0; qq < 3; ++qq) {
        tma::mbar_wait(&bar[B_E2C + qq], ph);
        fence_after();
        if (elect_one()) {
          // gW1 (+)= [gz2 hi | lo]^T [h1 hi | lo]
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            mma_ss(tb + T_ACC1, mndesc(OFF_Y, 4 * qq + ks), mndesc(OFF_X, 4 * qq + ks), ID_GW1, first | ((qq | ks) > 0));
          commit(&bar[B_GW1 + qq]);
        }
        __syncwarp();
      }
      if (!WIDE) {
#pragma unroll 1
        for (int qq = 0; qq < 3; ++qq) {
          tma::mbar_wait(&bar[B_E3 + qq], ph);
          fence_after();
          if (elect_one()) {
            // gW0 (+)= [gz1 hi | lo]^T [x~ hi | lo]
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              mma_ss(tb + T_ACC0, mndesc(OFF_Y, 4 * qq + ks), ktdesc(OFF_XT, 32, 4 * qq + ks), ID_GW0, first | ((qq | ks) > 0));
            if (qq == 2) commit(&bar[B_GW0]);
          }
          __syncwarp();
        }
      } else {
        for (int qq = 0; qq < 3; ++qq) tma::mbar_wait(&bar[B_E3 + qq], ph);
        if (lane == 0) {      // the gz1 plane leaves for pass C as it lies; reusable once the copy has READ it
          bulk_s2g(pa.gz1 + (size_t)tile * (PLANE / 4), sm + OFF_Y, PLANE);
          bulk_commit();
          bulk_wait_read();
          mbar_arrive(&bar[B_GW0]);
        }
        __syncwarp();
      }
    }
    if (WIDE && lane == 0) bulk_wait_all();
  }

  // ---- all tiles done: sums and accumulators -> per-CTA partials ------------------------------------------------------
  {
    int nit = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) ++nit;
    if (nit > 0) tma::mbar_wait(&bar[B_GW0], (nit - 1) & 1);
    fence_after();
    __syncthreads();
    constexpr int SW = 160 + UR + 1;
    float* stage = reinterpret_cast<float*>(sm);          // [128 lanes][SW]: the plane and weight regions are free now
    float* gst = stage + 128 * SW;                        // [96 rows][MJT] d log_std terms of the half-0 threads
    double* dst = reinterpret_cast<double*>(gst + ROWS * MJT);   // [96 rows][2] surrogate, KL
    if (warp < 4 && nit > 0) {
      const uint32_t tl = tb + ((uint32_t)(32 * warp) << 16);
      float* row = stage + (size_t)tid * SW;
#pragma unroll 1
      for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t v[32];
        tld32w(tl + T_ACC1 + c0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) row[c0 + i] = __uint_as_float(v[i]);
      }
      if (!WIDE) {
        uint32_t v[32];
        tld32w(tl + T_ACC0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) row[128 + i] = __uint_as_float(v[i]);
      }
#pragma unroll 1
      for (int c0 = 0; c0 < UR; c0 += 16) {
        uint32_t w[16];
        tld16(tl + T_ACC2 + c0, w);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) row[160 + c0 + i] = __uint_as_float(w[i]);
      }
    }
    if ((warp & 3) != 3 && (warp >> 2) == 0) {
      const int r = 32 * (warp & 3) + lane;
      dst[2 * r] = surr_acc; dst[2 * r + 1] = 0.0;
#pragma unroll
      for (int j = 0; j < MJT; ++j) gst[r * MJT + j] = gls[j];
    }
    fence_before();
    __syncthreads();
    if (tid == 0) {
      double s0 = 0.0, s1 = 0.0;
      for (int r = 0; r < ROWS; ++r) { s0 += dst[2 * r]; s1 += dst[2 * r + 1]; }
      a.spartials[2 * blockIdx.x] = s0;
      a.spartials[2 * blockIdx.x + 1] = s1;
    }
    {
      float* out = a.partials + (size_t)blockIdx.x * a.dstride;
      if (tid < m) {
        double s = 0.0;
        for (int r = 0; r < ROWS; ++r) s += (double)gst[r * MJT + tid];
        out[F.ols + tid] = (float)s;
      }
      if (nit == 0) {
        for (int i = tid; i < F.ols; i += NTHREADS) out[i] = 0.f;
      } else {
        auto S = [&](int lane_, int col) { return stage[(size_t)lane_ * SW + col]; };
        for (int i = tid; i < H * H; i += NTHREADS) {        // gW1[o][i]: lanes o | 64 + o, columns i | 64 + i
          const int o = i >> 6, k = i & 63;
          out[F.oW1 + i] = (S(o, k) + S(o, 64 + k)) + (S(64 + o, k) + S(64 + o, 64 + k));
        }
        if (!WIDE) {
          for (int i = tid; i < H * (n + 1); i += NTHREADS) {  // gW0[o][k], gb0[o] = column n
            const int o = i / (n + 1), k = i - o * (n + 1);
            const float v = (S(o, 128 + k) + S(o, 144 + k)) + (S(64 + o, 128 + k) + S(64 + o, 144 + k));
            if (k < n) out[F.oW0 + o * n + k] = v; else out[F.ob0 + o] = v;
          }
        }
        for (int i = tid; i < m * H; i += NTHREADS) {        // gW2[j][h]: lanes h | 64 + h, columns j | LOROW + j
          const int j = i >> 6, h = i & 63;
          out[F.oW2 + i] = (S(h, 160 + j) + S(h, 160 + LOROW + j)) + (S(64 + h, 160 + j) + S(64 + h, 160 + LOROW + j));
        }
      }
    }
    __syncthreads();
    if (warp == 7) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tb) : "memory");
  }
}

template <int MJT, bool WIDE>
static int launch_one(const PolArgs& pa, int grid, cudaStream_t s) {
  auto kern = pol_tc_kernel<MJT, WIDE>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Map<MJT>::ALLOC);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute(pol_tc): %s", cudaGetErrorString(e)); return (int)e; }
    attr = true;
  }
  kern<<<grid, NTHREADS, Map<MJT>::ALLOC, s>>>(pa);
  return mjrl_check_launch("pol_tc_kernel");
}

}  // namespace ptc

template <int MODE>
int launch_pol_tc(mjrl_ctx* ctx, Args& a, float* planes, const TcwRegions* wide, int* grid_out, cudaStream_t s) {
  const Flat F(a.n, a.m, a.h1, a.h2);
  const long long ntiles = (a.N + tc::ROWS - 1) / tc::ROWS;
  int grid = (int)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  static_assert(MODE == M_GRAD, "pol_tc.cu holds the gradient pass; the evaluation is launch_pol_fwd (pol_fwd.cu)");
  a.dstride = (F.d + 31) & ~31;
  a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
  if (!a.partials) return MJRL_E_STATE;
  a.spartials = (double*)mjrl_scratch2(ctx, sizeof(double) * 2 * grid);
  if (!a.spartials) return MJRL_E_STATE;
  ptc::PolArgs pa;
  pa.a = a; pa.planes = planes; pa.z1 = wide ? wide->z1 : nullptr; pa.gz1 = wide ? wide->gz1 : nullptr;
  static const bool phases = getenv("MJRL_B200_TC_PHASES") != nullptr;
  static long long* dbg = nullptr;
  if (phases && !dbg) cudaMalloc(&dbg, 16 * sizeof(long long));
  pa.dbg = phases ? dbg : nullptr;
  struct Report {      // debug only: synchronous print of the per-phase cycle stamps when the launcher returns
    long long* d; cudaStream_t st;
    ~Report() {
      if (!d) return;
      long long h[16];
      cudaStreamSynchronize(st);
      cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
      static const char* nm[12] = {"x~ + st", "wait M1", "P1 tanh->tmem", "P1 planes, loads", "wait M2", "P2 tanh, head, planes", "P2 exchange, LL, u^T",
                                   "gz2->tmem", "wait gW2 + planes", "wait M5", "P3 compute + waits", "P3 planes"};
      long long tot = 0;
      for (int i = 0; i < 12; ++i) tot += h[i];
      fprintf(stderr, "pol_tc phases (CTA 0, %lld tiles, %lld cycles per tile):", h[12], h[12] ? tot / h[12] : 0);
      for (int i = 0; i < 12; ++i) fprintf(stderr, " [%s %lld]", nm[i], h[12] ? h[i] / h[12] : 0);
      fprintf(stderr, "\n");
    }
  } report{pa.dbg, s};
  int rc;
  if (wide) {
    rc = launch_tcw_l0_fwd(ctx, *wide, a.params, a.N, a.n, a.m, grid, nullptr, s);
    if (rc) return rc;
    if (a.m <= 8) rc = ptc::launch_one<8, true>(pa, grid, s);
    else if (a.m <= 16) rc = ptc::launch_one<16, true>(pa, grid, s);
    else rc = ptc::launch_one<24, true>(pa, grid, s);
    if (rc) return rc;
    rc = launch_tcw_l0_grad(ctx, *wide, a.partials, a.dstride, a.N, a.n, a.m, grid, nullptr, s);
    return rc;
  }
  if (a.m <= 4) return ptc::launch_one<4, false>(pa, grid, s);
  return ptc::launch_one<8, false>(pa, grid, s);
}

template int launch_pol_tc<M_GRAD>(mjrl_ctx*, Args&, float*, const TcwRegions*, int*, cudaStream_t);

}  // namespace pol
