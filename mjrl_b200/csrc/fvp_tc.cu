// Fisher-vector product on the 5th-generation tensor cores (tcgen05.mma, accumulators and A operands in
// tensor memory) -- replaces the double backprop of mjrl/algos/npg_cg.py:55-74 inside a CG solve, for the
// two-hidden-layer tanh policy of mjrl/policies/gaussian_mlp.py:143-182 with 64 x 64 hidden units.
//
//   F v = sum_t J_t^T M J_t v,   J_t = d mu_t / d theta,   M = diag(out_scale^2 * 2 / (2 sigma^2 + 1e-8))
//
// computed per tile of 96 timesteps as a tangent pass followed by a backward pass, with h1, h2 of the (fixed)
// policy read from the activation cache:
//
//   dz1 = x~ dW0^T (+db0)                      MMA-1   A = x~      (TMEM)  B = dW0  (smem, K-major)
//   dh1 = (1 - h1^2) dz1                       E1      TMEM -> registers -> TMEM
//   dz2 = h1 dW1^T + dh1 W1^T (+db1)           MMA-2   A = h1, dh1 (TMEM)  B = dW1, W1
//   dh2 = (1 - h2^2) dz2;  dmu = h2 dW2^T + dh2 W2^T + db2;  u = M dmu;  gz2 = (u W2)(1 - h2^2)     E2 (CUDA cores)
//   gh1 = gz2 W1                               MMA-5   A = gz2     (TMEM)  B = W1^T
//   gz1 = gh1 (1 - h1^2)                       E3
//   gW2 += h2^T u,  gW1 += gz2^T h1,  gW0 += gz1^T [x~ 1]          weight-gradient MMAs, A and B from smem,
//                                                                   accumulators stay in TMEM for the whole kernel
//
// Precision: kind::tf32 with error compensation.  The tensor core truncates an fp32 operand to tf32, so the
// "hi" plane of every operand is the RAW fp32 value and the "lo" plane is x - trunc(x): a*b = hi*hi + lo*hi +
// hi*lo (+ lo*lo where it is free), 1e-6 relative (profiles/probes/umma2_probe.cu), i.e. fp32-class like the
// reference.  Forward-type products issue the three terms into one accumulator; the weight gradients stack
// [hi | lo] along M and along N, so ONE instruction per 8 timesteps yields all four blocks, summed at the end.
//
// Layouts (all measured in the probe): K-major operands use the 128-byte swizzle (16-byte atoms); MN-major
// tf32 operands -- everything contracted over the timestep -- only exist with the 128-byte swizzle on 32-byte
// atoms (SWIZZLE_128B_BASE32B), so the activation cache is stored in HBM in exactly that order and a tile
// lands in shared memory by ONE linear bulk copy, ready to be a weight-gradient operand.
//
// Roles: warps 0-2 and 4-6 own the 96 rows (TMEM lanes 0-95; warps q and q + 4 share a lane quadrant and split the
// 64 hidden columns of a row), warp 3 feeds the bulk copies and sums the bias gradients, one elected lane of
// warp 7 issues every MMA.  All hand-offs are mbarriers.
#include <stdlib.h>

#include "common.cuh"
#include "fvp_tc.cuh"
#include "policy_tile.cuh"

namespace pol {
namespace tc {

constexpr int KX = 16;              // padded width of [x~, 1]
constexpr int MJ = 8;               // largest action dimension
constexpr int NTHREADS = 256;
constexpr int ETHREADS = 2 * ROWS;  // epilogue threads (warps 0-2 and 4-6: two per row)
constexpr int KSTEPS = ROWS / 8;    // weight-gradient MMAs per product and tile

// ---- shared memory map (bytes) -------------------------------------------------------------------------
constexpr uint32_t OFF_X = 0;                       // h1 (raw = hi)              } adjacent: [h1_hi | h1_lo] is one
constexpr uint32_t OFF_W = OFF_X + PLANE;           // h1 lo                      } N = 128 MN-major operand
constexpr uint32_t OFF_Y = OFF_W + PLANE;           // h2 -> gz2 hi -> gz1 hi     } adjacent: one M = 128 operand
constexpr uint32_t OFF_Z = OFF_Y + PLANE;           // h2 lo -> gz2 lo -> gz1 lo  }
constexpr uint32_t WPLANE = H * 256;                // [64 rows][64 K] K-major, 128 B swizzle: 2 blocks of [64][128 B]
constexpr uint32_t OFF_DW1 = OFF_Z + PLANE;         // hi, lo
constexpr uint32_t OFF_W1 = OFF_DW1 + 2 * WPLANE;
constexpr uint32_t OFF_W1T = OFF_W1 + 2 * WPLANE;
constexpr uint32_t W0PLANE = H * 64;                // [64 rows][16 K] K-major, 64 B swizzle
constexpr uint32_t OFF_DW0 = OFF_W1T + 2 * WPLANE;  // hi, lo
constexpr uint32_t XTPLANE = 3 * 32 * 128;          // [32 rows = x~ hi (16) | x~ lo (16)][96 t] K-major: 3 blocks of [32][128 B]
constexpr uint32_t OFF_XT = OFF_DW0 + 2 * W0PLANE;
constexpr uint32_t UTPLANE = 3 * 16 * 128;          // [16 rows = u hi (8) | u lo (8)][96 t]
constexpr uint32_t OFF_UT = OFF_XT + XTPLANE;
constexpr uint32_t OFF_VEC = OFF_UT + UTPLANE;      // fp32 vectors, see Vec
constexpr uint32_t VEC_FLOATS = 2 * MJ * H + H + 3 * MJ + 2 * KX + ROWS * MJ;
constexpr uint32_t OFF_BAR = OFF_VEC + 4 * VEC_FLOATS;
constexpr int NBAR = 29;
constexpr uint32_t SMEM_BYTES = OFF_BAR + 8 * NBAR + 16;
constexpr uint32_t SMEM_ALLOC = SMEM_BYTES + 1024;  // the base is aligned up to 1024 in the kernel
struct Vec {   // float offsets inside the vector block
  static constexpr int W2 = 0;                 // [MJ][64]
  static constexpr int DW2 = MJ * H;           // [MJ][64]
  static constexpr int DB1 = 2 * MJ * H;       // [64]
  static constexpr int DB2 = DB1 + H;          // [MJ]
  static constexpr int FISHER = DB2 + MJ;      // [MJ]
  static constexpr int PADV = FISHER + MJ;     // [MJ] spare
  static constexpr int ISHIFT = PADV + MJ;     // [KX]
  static constexpr int IRCP = ISHIFT + KX;     // [KX] 1/(in_scale + 1e-8)
  static constexpr int UARR = IRCP + KX;       // [96][MJ] u of the tile (bias gradient gb2)
};
// barriers
// the tail of a tile is pipelined per TMEM lane quadrant q (32 rows): gW2 / gW1 / gW0 are issued in three parts of
// four k-steps, and the row threads of quadrant q only wait for (and signal) part q
enum { B_H1 = 0, B_H2, B_E0, B_E1, B_E2A, B_E2B, B_M1, B_M2A, B_M2B, B_M5, B_GW0,
       B_GW2 = 11, B_E2C = 14, B_GW1 = 17, B_E3 = 20, B_W3 = 23, B_H1Q = 26, B_COUNT = 29 };

// ---- tensor memory map (columns) ------------------------------------------------------------------------
constexpr uint32_t T_D1 = 0;        // dz1 -> dh1 hi (in place) -> gh1
constexpr uint32_t T_LO1 = 64;      // x~ hi (64..79), x~ lo (80..95) -> dh1 lo (64..127)
constexpr uint32_t T_D2 = 128;      // dz2
constexpr uint32_t T_A1 = 192;      // h1 hi (192..255), h1 lo (256..319) -> gz2 hi, gz2 lo
constexpr uint32_t T_ACC1 = 320;    // gW1: 128 lanes [gz2 hi | gz2 lo] x 128 columns [h1 hi | h1 lo]
constexpr uint32_t T_ACC0 = 448;    // gW0: x 32 columns [x~ hi | x~ lo]
constexpr uint32_t T_ACC2 = 480;    // gW2: 128 lanes [h2 hi | h2 lo] x 16 columns [u hi | u lo]
constexpr uint32_t T_COLS = 512;

struct TcArgs {
  const float* params;
  const float* vec;
  const float* obs;
  const float* cache;     // tc region: per tile [h1 plane | h2 plane] in shared-memory order
  const float* in_shift;
  const float* in_scale;
  const float* out_scale;
  float* partials;        // [grid][dstride]
  const int32_t* stop_flag;
  long long* dbg;         // debug: per-phase cycle sums of CTA 0 (MJRL_B200_TC_PHASES=1), else null
  long long N;
  int n, m, dstride;
};

template <int MJT>   // action dimension padded to 3, 4 or 8: the head loops are branch-free (rows j >= m of W2, dW2 are zero)
__global__ void __launch_bounds__(NTHREADS, 1) fvp_tc_kernel(const TcArgs a) {
  extern __shared__ unsigned char smraw[];
  __shared__ uint32_t tslot;
  if (a.stop_flag && *a.stop_flag) return;
  const uint32_t sraw = smem_u32(smraw);
  const uint32_t sb = (sraw + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sb - sraw);
  float* vecs = reinterpret_cast<float*>(sm + OFF_VEC);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + OFF_BAR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = a.n, m = a.m;
  const Flat F(n, m, H, H);
  const int ntiles = (int)((a.N + ROWS - 1) / ROWS);

  // ---- one-time staging ---------------------------------------------------------------------------------
  for (uint32_t i = tid; i < (OFF_VEC - OFF_DW1) / 16; i += NTHREADS)
    reinterpret_cast<float4*>(sm + OFF_DW1)[i] = make_float4(0.f, 0.f, 0.f, 0.f);   // weights, x~^T, u^T planes
  for (uint32_t i = tid; i < VEC_FLOATS; i += NTHREADS) vecs[i] = 0.f;
  if (tid == 0) {
    tma::mbar_init(&bar[B_H1], 1); tma::mbar_init(&bar[B_H2], 1);
    for (int b = B_E0; b <= B_E2B; ++b) tma::mbar_init(&bar[b], ETHREADS);
    for (int b = B_M1; b <= B_GW0; ++b) tma::mbar_init(&bar[b], 1);
    for (int qq = 0; qq < 3; ++qq) {
      tma::mbar_init(&bar[B_GW2 + qq], 1); tma::mbar_init(&bar[B_GW1 + qq], 1);
      tma::mbar_init(&bar[B_E2C + qq], 64); tma::mbar_init(&bar[B_E3 + qq], 64);
      tma::mbar_init(&bar[B_W3 + qq], 32);
      tma::mbar_init(&bar[B_H1Q + qq], 1);
    }
    tma::mbar_init_fence();
  }
  if (warp == 7) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tslot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  __syncthreads();
  {
    const float* P = a.params; const float* V = a.vec;
    auto put = [&](uint32_t plane_hi, uint32_t plane_bytes, uint32_t off, float w) {
      *reinterpret_cast<float*>(sm + plane_hi + off) = w;                      // raw: the tensor core truncates
      *reinterpret_cast<float*>(sm + plane_hi + plane_bytes + off) = lo_of(w);
    };
    for (int i = tid; i < H * H; i += NTHREADS) {
      const int o = i >> 6, k = i & 63;
      const float w = __ldg(P + F.oW1 + i);
      put(OFF_W1, WPLANE, k128_off(H, o, k), w);            // B[N = o][K = i]: dh1 W1^T
      put(OFF_W1T, WPLANE, k128_off(H, k, o), w);           // B[N = i][K = o]: gz2 W1
      put(OFF_DW1, WPLANE, k128_off(H, o, k), __ldg(V + F.oW1 + i));
    }
    for (int i = tid; i < H * n; i += NTHREADS) {
      const int o = i / n, k = i - o * n;
      put(OFF_DW0, W0PLANE, k64_off(o, k), __ldg(V + F.oW0 + i));
    }
    for (int o = tid; o < H; o += NTHREADS) {
      put(OFF_DW0, W0PLANE, k64_off(o, n), __ldg(V + F.ob0 + o));   // db0 rides on the ones column of x~
      vecs[Vec::DB1 + o] = __ldg(V + F.ob1 + o);
    }
    for (int i = tid; i < m * H; i += NTHREADS) {
      vecs[Vec::W2 + i] = __ldg(P + F.oW2 + i);
      vecs[Vec::DW2 + i] = __ldg(V + F.oW2 + i);
    }
    for (int j = tid; j < m; j += NTHREADS) {
      vecs[Vec::DB2 + j] = __ldg(V + F.ob2 + j);
      const float sg = expf(__ldg(P + F.ols + j));
      const float os = a.out_scale ? __ldg(a.out_scale + j) : 1.f;
      // d^2 KL / d mu^2 = 2/(2 sigma^2 + 1e-8) (gaussian_mlp.py:137-138), chained through out_scale twice
      vecs[Vec::FISHER + j] = os * os * (2.f / (2.f * sg * sg + 1e-8f));
    }
    for (int k = tid; k < n; k += NTHREADS) {
      vecs[Vec::ISHIFT + k] = a.in_shift ? __ldg(a.in_shift + k) : 0.f;
      vecs[Vec::IRCP + k] = 1.f / ((a.in_scale ? __ldg(a.in_scale + k) : 1.f) + 1e-8f);
    }
  }
  fence_async();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tb = tslot;

  unsigned char* pW = sm + OFF_W;
  unsigned char* pY = sm + OFF_Y; unsigned char* pZ = sm + OFF_Z;
  const float* tile_base = a.cache;     // + tile * 2 * PLANE / 4

  if ((warp & 3) != 3) {
    // ============ epilogue warps 0-2 and 4-6: thread = (timestep row, 32-column half of the hidden layer) =========
    // Warps q and q + 4 share TMEM lane quadrant q (tcgen05.ld/st reach lanes 32 (warp % 4) ..): two warps per
    // scheduler, each thread 32 of the 64 columns of its row.
    const int q = warp & 3, half = warp >> 2, blk = half;
    const int r = 32 * q + lane;
    const uint32_t tl = tb + ((uint32_t)(32 * q) << 16);
    const uint32_t pair_bar = 1 + q;
    auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory"); };
    long long ph_sum[14], tprev = 0;
#pragma unroll
    for (int i = 0; i < 14; ++i) ph_sum[i] = 0;
    const bool stamp_on = a.dbg != nullptr && tid == 0 && blockIdx.x == 0;
    auto stamp = [&](int i) { if (stamp_on) { const long long c = clock64(); ph_sum[i] += c - tprev; tprev = c; } };
    if (stamp_on) tprev = clock64();
    // raw observation row of the next tile, requested one tile ahead (half 1 owns x~)
    float ox[KX];
    auto prefetch_obs = [&](int tile) {
      const long long t = (long long)tile * ROWS + r;
#pragma unroll
      for (int k = 0; k < KX; ++k) ox[k] = (half == 1 && tile < ntiles && t < a.N && k < n) ? __ldg(a.obs + t * n + k) : 0.f;
    };
    prefetch_obs(blockIdx.x);
    // this thread's 32 columns of h1 come straight from the (L2-resident) plane cache, requested one tile ahead: the bulk copy
    // into the X plane is only awaited by the MMA issuer (gW1), not by the row threads (measured: 1,284 of 8,964 cycles per
    // tile were spent waiting for it here; 18 now, tile 8,850 cycles -- most of the wait moved to the MMA waits behind it)
    float4 hp[8];
    auto prefetch_h1 = [&](int tile) {
      if (tile < ntiles) {
        const float* base = tile_base + (size_t)tile * (2 * PLANE / 4);
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const float4* p4 = reinterpret_cast<const float4*>(base + mn_atom(r, 4 * blk + qq) / 4);
          hp[2 * qq] = __ldg(p4); hp[2 * qq + 1] = __ldg(p4 + 1);
        }
      }
    };
    prefetch_h1(blockIdx.x);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const long long t = (long long)tile * ROWS + r;
      const bool valid = t < a.N;
      // ---- E0: x~ and h1 as tensor-memory A operands, h1 lo as a weight-gradient operand -----------------
      float xs[KX];
      if (half == 1) {
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int k = 0; k < KX; ++k) {
          float v = 0.f;
          if (valid && k < n) v = (ox[k] - vecs[Vec::ISHIFT + k]) * vecs[Vec::IRCP + k];
          if (valid && k == n) v = 1.f;
          xs[k] = v;
          hi[k] = __float_as_uint(v); lo[k] = __float_as_uint(lo_of(v));
        }
        tst16(tl + T_LO1, hi);
        tst16(tl + T_LO1 + 16, lo);
      }
      stamp(0);
      stamp(1);
      float s1[32];      // 1 - h1^2 of this thread's 32 columns: used by E1 and E3
      {
        float h[32];
        uint32_t u32[32];
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          h[8 * qq + 0] = hp[2 * qq].x; h[8 * qq + 1] = hp[2 * qq].y; h[8 * qq + 2] = hp[2 * qq].z; h[8 * qq + 3] = hp[2 * qq].w;
          h[8 * qq + 4] = hp[2 * qq + 1].x; h[8 * qq + 5] = hp[2 * qq + 1].y; h[8 * qq + 6] = hp[2 * qq + 1].z; h[8 * qq + 7] = hp[2 * qq + 1].w;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) { u32[i] = __float_as_uint(h[i]); s1[i] = fmaf(-h[i], h[i], 1.f); }
        tst32(tl + T_A1 + 32 * blk, u32);
#pragma unroll
        for (int i = 0; i < 32; ++i) { h[i] = lo_of(h[i]); u32[i] = __float_as_uint(h[i]); }
        tst32(tl + T_A1 + 64 + 32 * blk, u32);
        wait_st();
        fence_before();
        mbar_arrive(&bar[B_E0]);
        // after the hand-off: the h1 lo plane is an operand of gW1 only (every thread fences before its B_E2A / B_E2C arrival)
        store_row32(pW, r, blk, h);
      }
      stamp(2);

      // ---- E1: dh1 = (1 - h1^2) dz1, in place over dz1 (hi) and next to it (lo) ---------------------------
      {
        uint32_t d[32], l[32];
        tma::mbar_wait(&bar[B_M1], ph);
        fence_after();
        stamp(3);
        tld32w(tl + T_D1 + 32 * blk, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float v = __uint_as_float(d[i]) * s1[i];
          d[i] = __float_as_uint(v); l[i] = __float_as_uint(lo_of(v));
        }
        tst32(tl + T_D1 + 32 * blk, d);
        tst32(tl + T_LO1 + 32 * blk, l);
      }
      wait_st();
      fence_before();
      mbar_arrive(&bar[B_E1]);
      stamp(4);

      // ---- E2: head on the CUDA cores: dh2, dmu, u, gz2 ---------------------------------------------------
      float s2[32];      // 1 - h2^2 of this thread's 32 columns, kept for gz2
      float u[MJT];
      {
        float h[32];
        uint32_t d[32];
        tma::mbar_wait(&bar[B_H2], ph);
        load_row32(pY, r, blk, h);
        float2 hv[16], dhv[16];      // packed pairs: the head runs on FFMA2 (two fp32 FMAs per instruction)
        float2 acc[MJT][2];
        // ... while MMA-2 is running: 1 - h2^2, the h2 dW2^T half of dmu, h2 lo
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
#pragma unroll
          for (int e = 0; e < 4; ++e) s2[4 * i4 + e] = fmaf(-h[4 * i4 + e], h[4 * i4 + e], 1.f);
          hv[2 * i4] = make_float2(h[4 * i4], h[4 * i4 + 1]);
          hv[2 * i4 + 1] = make_float2(h[4 * i4 + 2], h[4 * i4 + 3]);
        }
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          const float4* v2 = reinterpret_cast<const float4*>(vecs + Vec::DW2 + j * H + 32 * blk);
          float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 v = v2[i4];
            a0 = __ffma2_rn(hv[2 * i4], make_float2(v.x, v.y), a0);
            a1 = __ffma2_rn(hv[2 * i4 + 1], make_float2(v.z, v.w), a1);
          }
          acc[j][0] = a0; acc[j][1] = a1;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = lo_of(h[i]);
        store_row32(pZ, r, blk, h);                      // h2 lo: weight-gradient operand of gW2
        tma::mbar_wait(&bar[B_M2B], ph);
        fence_after();
        stamp(5);
        tld32w(tl + T_D2 + 32 * blk, d);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 b = *reinterpret_cast<const float4*>(vecs + Vec::DB1 + 32 * blk + 4 * i4);
          dhv[2 * i4] = __fmul2_rn(__fadd2_rn(make_float2(__uint_as_float(d[4 * i4]), __uint_as_float(d[4 * i4 + 1])), make_float2(b.x, b.y)),
                                   make_float2(s2[4 * i4], s2[4 * i4 + 1]));
          dhv[2 * i4 + 1] = __fmul2_rn(__fadd2_rn(make_float2(__uint_as_float(d[4 * i4 + 2]), __uint_as_float(d[4 * i4 + 3])), make_float2(b.z, b.w)),
                                       make_float2(s2[4 * i4 + 2], s2[4 * i4 + 3]));
        }
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          const float4* w2 = reinterpret_cast<const float4*>(vecs + Vec::W2 + j * H + 32 * blk);
          float2 a0 = acc[j][0], a1 = acc[j][1];
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 w = w2[i4];
            a0 = __ffma2_rn(dhv[2 * i4], make_float2(w.x, w.y), a0);
            a1 = __ffma2_rn(dhv[2 * i4 + 1], make_float2(w.z, w.w), a1);
          }
          u[j] = (a0.x + a0.y) + (a1.x + a1.y);
        }
      }
      // the two halves of a row swap their partial sums (half 0's through the lo rows of the u^T plane, half 1's through the u
      // array), ONE barrier, and both form the same u (the sum is commutative); then half 0 files u (u^T hi rows, u array for
      // the gb2 sums) and half 1 its lo part over the slot it has just read
      if (half == 0) {
#pragma unroll
        for (int j = 0; j < MJT; ++j) *reinterpret_cast<float*>(sm + OFF_UT + k128_off(16, 8 + j, r)) = u[j];
      } else {
#pragma unroll
        for (int j = 0; j < MJT; ++j) vecs[Vec::UARR + r * MJ + j] = u[j];
      }
      pair_sync();
#pragma unroll
      for (int j = 0; j < MJT; ++j) {
        const float other = half == 0 ? vecs[Vec::UARR + r * MJ + j]
                                      : *reinterpret_cast<const float*>(sm + OFF_UT + k128_off(16, 8 + j, r));
        const float dmu = u[j] + other;
        u[j] = (valid && j < m) ? vecs[Vec::FISHER + j] * (dmu + vecs[Vec::DB2 + j]) : 0.f;
      }
      if (half == 0) {
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          vecs[Vec::UARR + r * MJ + j] = u[j];
          // u^T planes, K-major over the timestep: rows j (hi) and 8 + j (lo), column r
          *reinterpret_cast<float*>(sm + OFF_UT + k128_off(16, j, r)) = u[j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < MJT; ++j) *reinterpret_cast<float*>(sm + OFF_UT + k128_off(16, 8 + j, r)) = lo_of(u[j]);
      }
      fence_async();
      mbar_arrive(&bar[B_E2A]);
      stamp(6);
      // gz2 = (u W2) (1 - h2^2): A operand of MMA-5 now, weight-gradient operand once gW2 has read h2
      float gz[32];
      {
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          float2 g0 = make_float2(0.f, 0.f), g1 = g0;
#pragma unroll
          for (int j = 0; j < MJT; ++j) {
            const float4 w = *reinterpret_cast<const float4*>(vecs + Vec::W2 + j * H + 32 * blk + 4 * i4);
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
      stamp(8);
      {
        store_row32(pY, r, blk, gz);
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = lo_of(gz[i]);
        store_row32(pZ, r, blk, v);
      }
      fence_async();
      mbar_arrive(&bar[B_E2C + q]);
      stamp(9);

      // ---- E3: gz1 = gh1 (1 - h1^2); x~^T planes for gW0 -------------------------------------------------
      {
        uint32_t d[32];
        tma::mbar_wait(&bar[B_M5], ph);
        fence_after();
        stamp(10);
        tld32w(tl + T_D1 + 32 * blk, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) gz[i] = valid ? __uint_as_float(d[i]) * s1[i] : 0.f;
      }
      fence_before();
      if (half == 1) {     // the x~^T planes were last read by gW0 of the previous tile, long complete
#pragma unroll
        for (int k = 0; k < KX; ++k) {
          *reinterpret_cast<float*>(sm + OFF_XT + k128_off(32, k, r)) = xs[k];
          *reinterpret_cast<float*>(sm + OFF_XT + k128_off(32, 16 + k, r)) = lo_of(xs[k]);
        }
      }
      float gl[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) gl[i] = lo_of(gz[i]);
      prefetch_h1(tile + gridDim.x);
      stamp(11);
      tma::mbar_wait(&bar[B_GW1 + q], ph);              // gW1 has read this quadrant's rows of gz2 (and every earlier MMA is done)
      tma::mbar_wait(&bar[B_W3 + q], ph);               // warp 3 has summed them for gb1
      stamp(12);
      store_row32(pY, r, blk, gz);
      store_row32(pZ, r, blk, gl);
      fence_async();
      mbar_arrive(&bar[B_E3 + q]);
      prefetch_obs(tile + gridDim.x);
      stamp(13);
    }
    if (stamp_on) { for (int i = 0; i < 14; ++i) a.dbg[i] = ph_sum[i]; a.dbg[14] = it; }
  } else if (warp == 3) {
    // =========================== loader + bias gradients ==================================================
    double gb1a = 0.0, gb1b = 0.0, gb2 = 0.0;
    auto load_h2 = [&](int tile) {                 // h2 -> Y, one copy
      if (lane == 0) {
        tma::mbar_expect(&bar[B_H2], PLANE);
        tma::bulk_g2s(sm + OFF_Y, tile_base + (size_t)tile * (2 * PLANE / 4) + PLANE / 4, PLANE, &bar[B_H2]);
      }
    };
    auto load_h1q = [&](int tile, int qq) {        // rows 32 qq .. of h1 -> X: 4 KB of each of the two column blocks
      if (lane == 0) {
        uint64_t* b = &bar[B_H1Q + qq];
        tma::mbar_expect(b, 2 * 32 * 128);
        const float* src = tile_base + (size_t)tile * (2 * PLANE / 4) + qq * (32 * 128 / 4);
        tma::bulk_g2s(sm + OFF_X + qq * (32 * 128), src, 32 * 128, b);
        tma::bulk_g2s(sm + OFF_X + BLK + qq * (32 * 128), src + BLK / 4, 32 * 128, b);
      }
    };
    if ((int)blockIdx.x < ntiles) {
      for (int qq = 0; qq < 3; ++qq) load_h1q(blockIdx.x, qq);
      load_h2(blockIdx.x);
    }
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      if (lane == 0 && tile + (int)gridDim.x < ntiles)   // the next tile's planes start their way to L2 now
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                     :: "l"(tile_base + (size_t)(tile + gridDim.x) * (2 * PLANE / 4)), "r"(2 * PLANE) : "memory");
      {
        const int c0 = 2 * lane, qa = c0 >> 3;
        float sx = 0.f, sy = 0.f;
#pragma unroll 1
        for (int qq = 0; qq < 3; ++qq) {
          tma::mbar_wait(&bar[B_E2C + qq], ph);       // rows of quadrant qq of gz2 (hi = the full fp32 value) are in Y
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
        if (lane < MJ) {    // u of the tile (written before B_E2A, which precedes every B_E2C)
          float f[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
          for (int rr = 0; rr < ROWS; rr += 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k) f[k] += vecs[Vec::UARR + (rr + k) * MJ + lane];
          }
          gb2 += (double)((f[0] + f[1]) + (f[2] + f[3]));
        }
      }
      const int next = tile + gridDim.x;
#pragma unroll 1
      for (int qq = 0; qq < 3; ++qq) {
        // gW1 part qq has read rows 32 qq .. of h1 hi | lo (the row threads read X only in E0: s1 stays in registers)
        tma::mbar_wait(&bar[B_GW1 + qq], ph);
        if (next < ntiles) load_h1q(next, qq);
      }
      tma::mbar_wait(&bar[B_GW0], ph);       // gz1 planes (Y, Z) are free
      if (next < ntiles) load_h2(next);
    }
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    out[F.ob1 + 2 * lane] = (float)gb1a;
    out[F.ob1 + 2 * lane + 1] = (float)gb1b;
    if (lane < m) out[F.ob2 + lane] = (float)gb2;
  } else {
    // =========================== MMA issuer (one elected lane of warp 7) ==================================
    const uint32_t ID_FWD = idesc_tf32(128, 64, 0, 0);
    const uint32_t ID_GW1 = idesc_tf32(128, 128, 1, 1);
    const uint32_t ID_GW0 = idesc_tf32(128, 32, 1, 0);
    const uint32_t ID_GW2 = idesc_tf32(128, 16, 1, 0);
    // K-major weights: k-step ks covers columns 8 ks..: block ks/4, 32 bytes per step inside the 128-byte row
    auto wdesc = [&](uint32_t off, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * (H * 128u) + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    auto w0desc = [&](uint32_t off, int ks) { return make_desc(sb + off + (uint32_t)ks * 32u, 16u, 512u, 4u); };
    // MN-major planes: k-step = 8 rows = 1024 bytes; 32-column blocks BLK apart; 4-row atoms 512 bytes apart
    auto mndesc = [&](uint32_t off, int ks) { return make_desc(sb + off + (uint32_t)ks * 1024u, BLK, 512u, 1u); };
    // K-major planes over the timestep ([R rows][96 t]): block = 32 timesteps
    auto ktdesc = [&](uint32_t off, int R, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * (uint32_t)R * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const uint32_t first = it == 0 ? 0u : 1u;
      tma::mbar_wait(&bar[B_E0], ph);
      fence_after();
      if (elect_one()) {
        // MMA-1: dz1 = [x~ hi | lo] dW0^T     (K = 16)
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
          mma_ts(tb + T_D1, tb + T_LO1 + 8 * ks, w0desc(OFF_DW0, ks), ID_FWD, ks > 0);
          mma_ts(tb + T_D1, tb + T_LO1 + 16 + 8 * ks, w0desc(OFF_DW0, ks), ID_FWD, 1);
          mma_ts(tb + T_D1, tb + T_LO1 + 8 * ks, w0desc(OFF_DW0 + W0PLANE, ks), ID_FWD, 1);
        }
        commit(&bar[B_M1]);
        // MMA-2a: dz2 = h1 dW1^T
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          mma_ts(tb + T_D2, tb + T_A1 + 8 * ks, wdesc(OFF_DW1, ks), ID_FWD, ks > 0);
          mma_ts(tb + T_D2, tb + T_A1 + 64 + 8 * ks, wdesc(OFF_DW1, ks), ID_FWD, 1);
          mma_ts(tb + T_D2, tb + T_A1 + 8 * ks, wdesc(OFF_DW1 + WPLANE, ks), ID_FWD, 1);
        }
        commit(&bar[B_M2A]);
      }
      __syncwarp();
      tma::mbar_wait(&bar[B_E1], ph);
      fence_after();
      if (elect_one()) {
        // MMA-2b: dz2 += dh1 W1^T
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          mma_ts(tb + T_D2, tb + T_D1 + 8 * ks, wdesc(OFF_W1, ks), ID_FWD, 1);
          mma_ts(tb + T_D2, tb + T_LO1 + 8 * ks, wdesc(OFF_W1, ks), ID_FWD, 1);
          mma_ts(tb + T_D2, tb + T_D1 + 8 * ks, wdesc(OFF_W1 + WPLANE, ks), ID_FWD, 1);
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
            mma_ss(tb + T_ACC2, mndesc(OFF_Y, ks), ktdesc(OFF_UT, 16, ks), ID_GW2, first | (ks > 0));
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
        tma::mbar_wait(&bar[B_H1Q + qq], ph);       // the bulk copy of this quadrant's h1 rows into X has landed
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
    }
  }

  // ---- all tiles done: accumulators -> per-CTA partial sums in flat parameter order -------------------------
  {
    int nit = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) ++nit;
    if (nit > 0 && warp != 3) tma::mbar_wait(&bar[B_GW0], (nit - 1) & 1);   // warp 3 has waited for it already
    fence_after();
    __syncthreads();
    float* stage = reinterpret_cast<float*>(sm);         // [128 lanes][176]: the plane region is free now
    constexpr int SW = 176;
    if (warp < 4) {
      const uint32_t tl = tb + ((uint32_t)(32 * warp) << 16);
      float* row = stage + (size_t)tid * SW;
#pragma unroll 1
      for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t v[32];
        tld32(tl + T_ACC1 + c0, v);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) row[c0 + i] = __uint_as_float(v[i]);
      }
      {
        uint32_t v[32];
        tld32(tl + T_ACC0, v);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) row[128 + i] = __uint_as_float(v[i]);
        uint32_t w[16];
        tld16(tl + T_ACC2, w);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) row[160 + i] = __uint_as_float(w[i]);
      }
    }
    fence_before();
    __syncthreads();
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    if (nit == 0) {
      for (int i = tid; i < F.ols; i += NTHREADS) out[i] = 0.f;
    } else {
      auto S = [&](int lane_, int col) { return stage[(size_t)lane_ * SW + col]; };
      for (int i = tid; i < H * H; i += NTHREADS) {        // gW1[o][i]: lanes o | 64 + o, columns i | 64 + i
        const int o = i >> 6, k = i & 63;
        out[F.oW1 + i] = (S(o, k) + S(o, 64 + k)) + (S(64 + o, k) + S(64 + o, 64 + k));
      }
      for (int i = tid; i < H * (n + 1); i += NTHREADS) {  // gW0[o][k], gb0[o] = column n
        const int o = i / (n + 1), k = i - o * (n + 1);
        const float v = (S(o, 128 + k) + S(o, 144 + k)) + (S(64 + o, 128 + k) + S(64 + o, 144 + k));
        if (k < n) out[F.oW0 + o * n + k] = v; else out[F.ob0 + o] = v;
      }
      for (int i = tid; i < m * H; i += NTHREADS) {        // gW2[j][h]: lanes h | 64 + h, columns j | 8 + j
        const int j = i >> 6, h = i & 63;
        out[F.oW2 + i] = (S(h, 160 + j) + S(h, 168 + j)) + (S(64 + h, 160 + j) + S(64 + h, 168 + j));
      }
    }
    __syncthreads();
    if (warp == 7) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tb) : "memory");
  }
}

// Activation cache of the rows-in-registers kernels (MMA-fragment order, pol_rows.cuh) -> tile planes in the
// shared-memory order of this kernel.  One thread moves one 32-byte atom of h1 and of h2; rows past N are zero.
__global__ void __launch_bounds__(256) tc_fill_kernel(const float* __restrict__ frag, float* __restrict__ out,
                                                      long long N, long long natoms) {
  constexpr int NF = H / 8;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < natoms; i += stride) {
    const int q = (int)(i & 7);
    const long long row = i >> 3;
    const long long tile = row / ROWS;
    const int r = (int)(row - tile * ROWS);
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, b0 = a0, b1 = a0;
    if (row < N) {
      const long long t64 = row >> 6;
      const int r64 = (int)(row & 63), hw = r64 >> 4, rr = r64 & 15, g = rr & 7, isB = rr >> 3;
      const float4* src = reinterpret_cast<const float4*>(
          reinterpret_cast<const float2*>(frag) + ((size_t)(t64 * 4 + hw) * (4 * NF * 32) + (size_t)(2 * q + isB) * 32 + g * 4));
      a0 = src[0]; a1 = src[1];
      const float4* src2 = src + (2 * NF * 32) / 2;     // h2 slots follow the 2 NF h1 slots (float2 units)
      b0 = src2[0]; b1 = src2[1];
    }
    float* dst = out + (size_t)tile * (2 * PLANE / 4) + mn_atom(r, q) / 4;
    reinterpret_cast<float4*>(dst)[0] = a0; reinterpret_cast<float4*>(dst)[1] = a1;
    reinterpret_cast<float4*>(dst + PLANE / 4)[0] = b0; reinterpret_cast<float4*>(dst + PLANE / 4)[1] = b1;
  }
}

}  // namespace tc

// ---- host side (declared in policy_tile.cuh) -------------------------------------------------------------
bool tc_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2) {
  static const bool off = getenv("MJRL_B200_FVP_TC") && getenv("MJRL_B200_FVP_TC")[0] == '0';
  if (off || !ctx) return false;
  if (h1 != tc::H || h2 != tc::H || n + 1 > tc::KX || m < 1 || m > tc::MJ) return false;
  return (size_t)tc::SMEM_ALLOC <= (size_t)ctx->max_smem_optin;
}

long long tc_cache_floats(long long N) {
  const long long ntiles = (N + tc::ROWS - 1) / tc::ROWS;
  return ntiles * (2 * tc::PLANE / 4);
}

int launch_tc_fill(mjrl_ctx* ctx, const float* frag_cache, float* tc_cache, long long N, cudaStream_t s) {
  const long long ntiles = (N + tc::ROWS - 1) / tc::ROWS;
  const long long natoms = ntiles * tc::ROWS * 8;
  long long blocks = (natoms + 255) / 256;
  const long long cap = (long long)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  tc::tc_fill_kernel<<<(int)blocks, 256, 0, s>>>(frag_cache, tc_cache, N, natoms);
  return mjrl_check_launch("tc_fill_kernel");
}

int launch_fvp_tc(mjrl_ctx* ctx, Args& a, const float* tc_cache, int* grid_out, cudaStream_t s) {
  if (!tc_cache || !tc_fits(ctx, a.n, a.m, a.h1, a.h2)) return -1;
  const Flat F(a.n, a.m, a.h1, a.h2);
  const long long ntiles = (a.N + tc::ROWS - 1) / tc::ROWS;
  int grid = (int)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  a.dstride = (F.d + 31) & ~31;
  a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
  if (!a.partials) return MJRL_E_STATE;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(tc::fvp_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM_ALLOC);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(tc::fvp_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM_ALLOC);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(tc::fvp_tc_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM_ALLOC);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute(fvp_tc): %s", cudaGetErrorString(e)); return (int)e; }
    attr = true;
  }
  tc::TcArgs t;
  t.params = a.params; t.vec = a.vec; t.obs = a.obs; t.cache = tc_cache;
  t.in_shift = a.in_shift; t.in_scale = a.in_scale; t.out_scale = a.out_scale;
  t.partials = a.partials; t.stop_flag = a.stop_flag; t.N = a.N; t.n = a.n; t.m = a.m; t.dstride = a.dstride;
  static const bool phases = getenv("MJRL_B200_TC_PHASES") != nullptr;
  static long long* dbg = nullptr;
  if (phases && !dbg) cudaMalloc(&dbg, 16 * sizeof(long long));
  t.dbg = phases ? dbg : nullptr;
  if (a.m <= 3) tc::fvp_tc_kernel<3><<<grid, tc::NTHREADS, tc::SMEM_ALLOC, s>>>(t);
  else if (a.m <= 4) tc::fvp_tc_kernel<4><<<grid, tc::NTHREADS, tc::SMEM_ALLOC, s>>>(t);
  else tc::fvp_tc_kernel<8><<<grid, tc::NTHREADS, tc::SMEM_ALLOC, s>>>(t);
  if (phases) {   // debug only: synchronous
    long long h[16];
    cudaStreamSynchronize(s);
    cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
    static const char* nm[14] = {"x~ + st", "wait h1", "E0 h1->tmem", "wait M1", "E1", "wait M2b (+h2)", "E2a dmu,u", "E2b gz2->tmem",
                                 "wait gW2", "E2c planes", "wait M5", "E3 compute", "wait gW1,w3", "E3 planes"};
    long long tot = 0;
    for (int i = 0; i < 14; ++i) tot += h[i];
    fprintf(stderr, "fvp_tc phases (CTA 0, %lld tiles, %lld cycles per tile):", h[14], h[14] ? tot / h[14] : 0);
    for (int i = 0; i < 14; ++i) fprintf(stderr, " [%s %lld]", nm[i], h[14] ? h[i] / h[14] : 0);
    fprintf(stderr, "\n");
  }
  return mjrl_check_launch("fvp_tc_kernel");
}

}  // namespace pol
