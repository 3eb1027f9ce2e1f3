// Fisher-vector product on tcgen05 for WIDE observation / action vectors (Ant 111 / 8, Adroit 45 / 24: obs_dim + 1
// <= 128, act_dim <= 24, 64 x 64 hidden units) -- same mathematics and precision scheme as fvp_tc.cu (which holds the
// whole product of a narrow policy in one kernel), same reference lines (mjrl/algos/npg_cg.py:55-74 over
// mjrl/policies/gaussian_mlp.py:143-182).
//
// A [x~ | 1] tile of 112 columns with its two fp32 split planes and the 112 x 64 first-layer operand do not fit next
// to the four activation planes and three 64 x 64 weight operands of the fused kernel (shared memory and tensor
// memory are both full there), so the first layer runs as two passes of its own, streaming two images through HBM:
//
//   pass A  l0_fwd_kernel     dz1 = [x~ 1] [dW0 | db0]^T                     -> DZ1 image   (256 B / timestep)
//   main    fvp_wide_kernel   the fused kernel without its first-layer MMAs:    DZ1 -> ... -> GZ1 image, gW1, gW2, gb1, gb2
//   pass C  l0_grad_kernel    [gW0 | gb0] = gz1^T [x~ 1]                     <- GZ1 image   (256 B / timestep)
//
// The observation side of both passes is ONE image built once per gradient pass (ximg_kernel): [x~ | 1 | 0..] of
// every 96-timestep tile in exactly the K-major 128-byte-swizzled order the MMA reads, so a tile lands in shared memory
// by one bulk copy and IS the hi plane (the tensor core truncates fp32 to tf32); the lo plane is an elementwise pass
// over it.  The GZ1 image is the main kernel's shared-memory gz1 plane stored as it lies (MN-major, 32-byte atoms), so
// pass C gets its B operand the same way; its A operand is x~ transposed into tensor memory (lane = feature, column =
// timestep), which the K-major swizzle makes a conflict-free column read.
//
// Algorithmic HBM bytes per timestep and product: 512 (h1 | h2) + 2 x 128 KB (x~ image) + 4 x 256 (DZ1, GZ1 written and
// read) = 1,536 + 256 KB with KB = ceil((obs_dim + 1) / 32) 32-column blocks.
#include <stdlib.h>

#include "common.cuh"
#include "fvp_tc.cuh"
#include "policy_tile.cuh"

namespace pol {
namespace tcw {

using namespace tc;

constexpr int MAXKB = 4;                           // obs_dim + 1 <= 128
constexpr uint32_t XBLK = ROWS * 128;              // one 32-column block of the x~ image: [96 rows][128 B]
constexpr int AUX_THREADS = 192;                   // passes A and C: 4 row / lane warps, loader, MMA issuer

// ---- x~ image ----------------------------------------------------------------------------------------------
// per tile: KB blocks of [96 rows][32 columns], 128-byte rows, 16-byte chunks XOR-swizzled by the row index mod 8;
// column obs_dim holds the 1 of the bias, columns past it and rows past N are zero.  One CTA per (tile, block): no index
// divisions, the transform of the block's 32 columns staged once (the first version decomposed a flat 64-bit index and
// divided per element: 1.23 ms at the Ant shape, issue bound).
__global__ void __launch_bounds__(256) ximg_kernel(const float* __restrict__ obs, const float* __restrict__ in_shift,
                                                   const float* __restrict__ in_scale, float* __restrict__ img,
                                                   long long N, int n, int KB) {
  __shared__ float sh[32], rc[32];
  const int tid = threadIdx.x, kb = blockIdx.y;
  const long long tile = blockIdx.x;
  if (tid < 32) {
    const int k = kb * 32 + tid;
    sh[tid] = (k < n && in_shift) ? __ldg(in_shift + k) : 0.f;
    rc[tid] = k < n ? 1.f / ((in_scale ? __ldg(in_scale + k) : 1.f) + 1e-8f) : 0.f;
  }
  __syncthreads();
  float4* out = reinterpret_cast<float4*>(img) + ((size_t)tile * KB + kb) * (ROWS * 8);
  for (int c = tid; c < ROWS * 8; c += 256) {
    const int r = c >> 3, ch = c & 7;
    const long long t = tile * ROWS + r;
    const int lc = (ch ^ (r & 7)) << 2, k0 = kb * 32 + lc;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = k0 + e;
      float x = 0.f;
      if (t < N) {
        if (k < n) x = (__ldg(obs + t * n + k) - sh[lc + e]) * rc[lc + e];
        else if (k == n) x = 1.f;
      }
      v[e] = x;
    }
    out[c] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

struct AuxArgs {
  const float* ximg;      // x~ image
  const float* vec;       // pass A: the CG direction (dW0, db0 rows)
  float* dz1;             // pass A out: per tile [16 column chunks][96 rows][4 floats]
  const float* gz1;       // pass C in: per tile the gz1 plane in shared-memory order
  float* partials;        // pass C out: [grid][dstride], the W0 and b0 slices
  const int32_t* stop_flag;
  long long N;
  int n, m, dstride, KB;
};

// ---- pass A: dz1 = [x~ 1] [dW0 | db0]^T -----------------------------------------------------------------------
// shared memory: XHI[2] (bulk copies, double buffered) | XLO | W operand: KB blocks of [128 rows = dW0 hi (64) | lo (64)]
// [32 K] K-major.  Tensor memory: two accumulators of 128 columns [x hi . W hi + x lo . W hi | x hi . W lo].
template <int KB>
struct FwdMap {
  static constexpr uint32_t XB = KB * XBLK;
  static constexpr uint32_t OFF_XHI = 0;
  static constexpr uint32_t OFF_XLO = 2 * XB;
  static constexpr uint32_t OFF_WB = 3 * XB;
  static constexpr uint32_t WBLK = 128 * 128;
  static constexpr uint32_t OFF_BAR = OFF_WB + KB * WBLK + 4096;   // slack: an M = 128 operand reads 32 rows past a 96-row block
  static constexpr uint32_t BYTES = OFF_BAR + 64;
  static constexpr uint32_t ALLOC = BYTES + 1024;
};
enum { FA_FULL = 0, FA_LO = 2, FA_MMA = 3, FA_COUNT = 5 };

template <int KB>
__global__ void __launch_bounds__(AUX_THREADS, 1) l0_fwd_kernel(const AuxArgs a) {
  using M = FwdMap<KB>;
  extern __shared__ unsigned char smraw[];
  __shared__ uint32_t tslot;
  if (a.stop_flag && *a.stop_flag) return;
  const uint32_t sraw = smem_u32(smraw);
  const uint32_t sb = (sraw + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sb - sraw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + M::OFF_BAR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = a.n;
  const Flat F(n, a.m, H, H);
  const int ntiles = (int)((a.N + ROWS - 1) / ROWS);
  const int ksteps = (n + 1 + 7) >> 3;

  for (uint32_t i = tid; i < (M::OFF_BAR - M::OFF_XLO) / 16; i += AUX_THREADS)
    reinterpret_cast<float4*>(sm + M::OFF_XLO)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid == 0) {
    tma::mbar_init(&bar[FA_FULL], 1); tma::mbar_init(&bar[FA_FULL + 1], 1);
    tma::mbar_init(&bar[FA_LO], 128);
    tma::mbar_init(&bar[FA_MMA], 1); tma::mbar_init(&bar[FA_MMA + 1], 1);
    tma::mbar_init_fence();
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tslot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  __syncthreads();
  for (int i = tid; i < H * (n + 1); i += AUX_THREADS) {
    const int o = i / (n + 1), k = i - o * (n + 1);
    const float w = k < n ? __ldg(a.vec + F.oW0 + o * n + k) : __ldg(a.vec + F.ob0 + o);
    *reinterpret_cast<float*>(sm + M::OFF_WB + k128_off(128, o, k)) = w;
    *reinterpret_cast<float*>(sm + M::OFF_WB + k128_off(128, 64 + o, k)) = lo_of(w);
  }
  fence_async();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tb = tslot;

  if (warp < 4) {
    const uint32_t tl = tb + ((uint32_t)(32 * warp) << 16);
    const int r = 32 * warp + lane;
    auto epilogue = [&](int tile, int buf) {      // accumulator buf -> DZ1 image (warps 0-2: the 96 real rows)
      if (warp == 3) return;
      float4* dst = reinterpret_cast<float4*>(a.dz1) + (size_t)tile * (ROWS * H / 4) + r;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t p[32], q[32];
        tld32w(tl + 128 * buf + c0, p);
        tld32w(tl + 128 * buf + 64 + c0, q);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          dst[(size_t)(c0 / 4 + j) * ROWS] =
              make_float4(__uint_as_float(p[4 * j]) + __uint_as_float(q[4 * j]), __uint_as_float(p[4 * j + 1]) + __uint_as_float(q[4 * j + 1]),
                          __uint_as_float(p[4 * j + 2]) + __uint_as_float(q[4 * j + 2]), __uint_as_float(p[4 * j + 3]) + __uint_as_float(q[4 * j + 3]));
      }
    };
    int it = 0, prev_tile = -1;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      tma::mbar_wait(&bar[FA_FULL + buf], (it >> 1) & 1);
      if (it > 0) { tma::mbar_wait(&bar[FA_MMA + (buf ^ 1)], ((it - 1) >> 1) & 1); fence_after(); }
      const float4* src = reinterpret_cast<const float4*>(sm + M::OFF_XHI + buf * M::XB);
      float4* dst = reinterpret_cast<float4*>(sm + M::OFF_XLO);
#pragma unroll 4
      for (uint32_t i = tid; i < M::XB / 16; i += 128) {
        const float4 v = src[i];
        dst[i] = make_float4(lo_of(v.x), lo_of(v.y), lo_of(v.z), lo_of(v.w));
      }
      fence_async();
      fence_before();
      mbar_arrive(&bar[FA_LO]);
      if (it > 0) epilogue(prev_tile, buf ^ 1);
      prev_tile = tile;
    }
    if (it > 0) {
      tma::mbar_wait(&bar[FA_MMA + ((it - 1) & 1)], ((it - 1) >> 1) & 1);
      fence_after();
      epilogue(prev_tile, (it - 1) & 1);
    }
    fence_before();
  } else if (warp == 4) {
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      if (it >= 2) tma::mbar_wait(&bar[FA_MMA + buf], ((it - 2) >> 1) & 1);   // the MMAs of tile it - 2 have read XHI[buf]
      if (lane == 0) {
        tma::mbar_expect(&bar[FA_FULL + buf], M::XB);
        tma::bulk_g2s(sm + M::OFF_XHI + buf * M::XB, a.ximg + (size_t)tile * (M::XB / 4), M::XB, &bar[FA_FULL + buf]);
      }
    }
  } else {
    const uint32_t ID128 = idesc_tf32(128, 128, 0, 0), ID64 = idesc_tf32(128, 64, 0, 0);
    auto xdesc = [&](uint32_t off, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * XBLK + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    auto wdesc = [&](int ks) {
      return make_desc(sb + M::OFF_WB + (uint32_t)(ks >> 2) * M::WBLK + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      tma::mbar_wait(&bar[FA_LO], it & 1);
      fence_after();
      if (elect_one()) {
        const uint32_t d = tb + 128 * buf;
        for (int ks = 0; ks < ksteps; ++ks) {
          mma_ss(d, xdesc(M::OFF_XHI + buf * M::XB, ks), wdesc(ks), ID128, ks > 0);
          mma_ss(d, xdesc(M::OFF_XLO, ks), wdesc(ks), ID64, 1);
        }
        commit(&bar[FA_MMA + buf]);
      }
      __syncwarp();
    }
  }
  __syncthreads();
  if (warp == 5) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tb) : "memory");
}

// ---- pass C: [gW0 | gb0] = gz1^T [x~ 1] -----------------------------------------------------------------------
// shared memory: XS[2] (x~ image tiles) | G[2] = [gz1 hi (bulk copy) | gz1 lo] (one N = 128 MN-major operand each).
// Tensor memory: A[2] = x~^T hi (96 columns) | lo (96 columns), lane = feature; accumulator 128 columns
// [x hi . g hi + x lo . g hi | x hi . g lo] for the whole kernel.
template <int KB>
struct GradMap {
  static constexpr uint32_t XB = KB * XBLK;
  static constexpr uint32_t OFF_XS = 0;
  static constexpr uint32_t OFF_G = 2 * XB;
  static constexpr uint32_t OFF_BAR = OFF_G + 4 * PLANE;
  static constexpr uint32_t BYTES = OFF_BAR + 64;
  static constexpr uint32_t ALLOC = BYTES + 1024;
};
enum { GC_FULL = 0, GC_RDY = 2, GC_MMA = 4, GC_COUNT = 6 };
constexpr uint32_t TC_A = 0, TC_D = 384;

template <int KB>
__global__ void __launch_bounds__(AUX_THREADS, 1) l0_grad_kernel(const AuxArgs a) {
  using M = GradMap<KB>;
  extern __shared__ unsigned char smraw[];
  __shared__ uint32_t tslot;
  if (a.stop_flag && *a.stop_flag) return;
  const uint32_t sraw = smem_u32(smraw);
  const uint32_t sb = (sraw + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sb - sraw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + M::OFF_BAR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = a.n;
  const Flat F(n, a.m, H, H);
  const int ntiles = (int)((a.N + ROWS - 1) / ROWS);

  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      tma::mbar_init(&bar[GC_FULL + b], 1); tma::mbar_init(&bar[GC_RDY + b], 128); tma::mbar_init(&bar[GC_MMA + b], 1);
    }
    tma::mbar_init_fence();
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tslot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tb = tslot;
  int nit = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) ++nit;

  if (warp < 4) {
    const uint32_t tl = tb + ((uint32_t)(32 * warp) << 16);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      tma::mbar_wait(&bar[GC_FULL + buf], (it >> 1) & 1);
      if (it >= 2) { tma::mbar_wait(&bar[GC_MMA + buf], ((it - 2) >> 1) & 1); fence_after(); }   // A[buf], G lo[buf] are free
      if (warp < KB) {
        // feature k = 32 warp + lane: column k of the tile; for a fixed row the 32 lanes read one 128-byte line
        const unsigned char* xs = sm + M::OFF_XS + buf * M::XB + (uint32_t)warp * XBLK;
#pragma unroll 1
        for (int c0 = 0; c0 < ROWS; c0 += 32) {
          uint32_t hi[32], lo[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int r = c0 + i;
            const float v = *reinterpret_cast<const float*>(xs + r * 128 + ((((lane >> 2) ^ (r & 7))) << 4) + (lane & 3) * 4);
            hi[i] = __float_as_uint(v); lo[i] = __float_as_uint(lo_of(v));
          }
          tst32(tl + TC_A + 192 * buf + c0, hi);
          tst32(tl + TC_A + 192 * buf + 96 + c0, lo);
        }
      }
      {
        const float4* src = reinterpret_cast<const float4*>(sm + M::OFF_G + buf * 2 * PLANE);
        float4* dst = reinterpret_cast<float4*>(sm + M::OFF_G + buf * 2 * PLANE + PLANE);
#pragma unroll 4
        for (uint32_t i = tid; i < PLANE / 16; i += 128) {
          const float4 v = src[i];
          dst[i] = make_float4(lo_of(v.x), lo_of(v.y), lo_of(v.z), lo_of(v.w));
        }
      }
      wait_st();
      fence_async();
      fence_before();
      mbar_arrive(&bar[GC_RDY + buf]);
    }
    // ---- accumulator -> this CTA's W0 / b0 slice of the partial sums ----
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    if (nit > 0) {
      tma::mbar_wait(&bar[GC_MMA + ((nit - 1) & 1)], ((nit - 1) >> 1) & 1);
      fence_after();
      const int k = 32 * warp + lane;
      if (warp < KB) {
#pragma unroll 1
        for (int c0 = 0; c0 < 64; c0 += 32) {
          uint32_t p[32], q[32];
          tld32w(tl + TC_D + c0, p);
          tld32w(tl + TC_D + 64 + c0, q);
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float v = __uint_as_float(p[i]) + __uint_as_float(q[i]);
            if (k < n) out[F.oW0 + (c0 + i) * n + k] = v;
            else if (k == n) out[F.ob0 + c0 + i] = v;
          }
        }
      }
    } else {
      for (int i = tid; i < H * n; i += 128) out[F.oW0 + i] = 0.f;
      for (int i = tid; i < H; i += 128) out[F.ob0 + i] = 0.f;
    }
    fence_before();
  } else if (warp == 4) {
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      if (it >= 2) tma::mbar_wait(&bar[GC_MMA + buf], ((it - 2) >> 1) & 1);
      if (lane == 0) {
        tma::mbar_expect(&bar[GC_FULL + buf], M::XB + PLANE);
        tma::bulk_g2s(sm + M::OFF_XS + buf * M::XB, a.ximg + (size_t)tile * (M::XB / 4), M::XB, &bar[GC_FULL + buf]);
        tma::bulk_g2s(sm + M::OFF_G + buf * 2 * PLANE, a.gz1 + (size_t)tile * (PLANE / 4), PLANE, &bar[GC_FULL + buf]);
      }
    }
  } else {
    const uint32_t ID128 = idesc_tf32(128, 128, 0, 1), ID64 = idesc_tf32(128, 64, 0, 1);
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      tma::mbar_wait(&bar[GC_RDY + buf], (it >> 1) & 1);
      fence_after();
      if (elect_one()) {
        const uint32_t g = sb + M::OFF_G + buf * 2 * PLANE;
#pragma unroll
        for (int ks = 0; ks < ROWS / 8; ++ks) {
          const uint64_t gd = make_desc(g + (uint32_t)ks * 1024u, BLK, 512u, 1u);
          mma_ts(tb + TC_D, tb + TC_A + 192 * buf + 8 * ks, gd, ID128, (it > 0) | (ks > 0));
          mma_ts(tb + TC_D, tb + TC_A + 192 * buf + 96 + 8 * ks, gd, ID64, 1);
        }
        commit(&bar[GC_MMA + buf]);
      }
      __syncwarp();
    }
  }
  __syncthreads();
  if (warp == 5) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tb) : "memory");
}

// ---- main kernel ----------------------------------------------------------------------------------------------
// fvp_tc_kernel without MMA-1 and the gW0 MMAs: dz1 comes from the DZ1 image (E0 / E1), the gz1 plane leaves as one
// bulk store (after E3).  The action dimension is padded to MJT = 8, 16 or 24; the two threads of a row exchange the
// head's partial sums through the u^T operand plane itself.
constexpr int NTHREADS = 256;
constexpr int ETHREADS = 2 * ROWS;
constexpr uint32_t WPLANE = H * 256;
constexpr uint32_t OFF_X = 0, OFF_W = PLANE, OFF_Y = 2 * PLANE, OFF_Z = 3 * PLANE;
constexpr uint32_t OFF_DW1 = 4 * PLANE, OFF_W1 = OFF_DW1 + 2 * WPLANE, OFF_W1T = OFF_W1 + 2 * WPLANE;
constexpr uint32_t OFF_UT = OFF_W1T + 2 * WPLANE;
template <int MJT>
struct WideMap {
  static constexpr int UR = 2 * MJT;                               // rows of the u^T plane: u hi | u lo
  static constexpr uint32_t UTPLANE = 3 * UR * 128;                // [UR rows][96 t] K-major: 3 blocks of [UR][128 B]
  static constexpr uint32_t OFF_VEC = OFF_UT + UTPLANE;
  static constexpr int W2 = 0, DW2 = MJT * H, DB1 = 2 * MJT * H, DB2 = DB1 + H, FISHER = DB2 + MJT;
  static constexpr uint32_t VEC_FLOATS = FISHER + MJT;
  static constexpr uint32_t OFF_BAR = (OFF_VEC + 4 * VEC_FLOATS + 15) & ~15u;
  static constexpr uint32_t BYTES = OFF_BAR + 8 * 32;
  static constexpr uint32_t ALLOC = BYTES + 1024;
};
enum { B_H1 = 0, B_H2, B_E0, B_E1, B_E2A, B_E2B, B_M1, B_M2A, B_M2B, B_M5, B_GW0,
       B_GW2 = 11, B_E2C = 14, B_GW1 = 17, B_E3 = 20, B_W3 = 23, B_H1Q = 26, B_COUNT = 29 };
constexpr uint32_t T_D1 = 0, T_LO1 = 64, T_D2 = 128, T_A1 = 192, T_ACC1 = 320, T_ACC2 = 448;

struct WideArgs {
  const float* params;
  const float* vec;
  const float* cache;     // per tile [h1 plane | h2 plane] in shared-memory order
  const float* dz1;       // DZ1 image (pass A)
  float* gz1;             // GZ1 image (for pass C)
  const float* out_scale;
  float* partials;
  const int32_t* stop_flag;
  long long N;
  int n, m, dstride;
};

template <int MJT>
__global__ void __launch_bounds__(NTHREADS, 1) fvp_wide_kernel(const WideArgs a) {
  using M = WideMap<MJT>;
  constexpr int UR = M::UR;
  extern __shared__ unsigned char smraw[];
  __shared__ uint32_t tslot;
  if (a.stop_flag && *a.stop_flag) return;
  const uint32_t sraw = smem_u32(smraw);
  const uint32_t sb = (sraw + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sb - sraw);
  float* vecs = reinterpret_cast<float*>(sm + M::OFF_VEC);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + M::OFF_BAR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = a.n, m = a.m;
  const Flat F(n, m, H, H);
  const int ntiles = (int)((a.N + ROWS - 1) / ROWS);

  for (uint32_t i = tid; i < (M::OFF_VEC - OFF_DW1) / 16; i += NTHREADS)
    reinterpret_cast<float4*>(sm + OFF_DW1)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (uint32_t i = tid; i < M::VEC_FLOATS; i += NTHREADS) vecs[i] = 0.f;
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
    auto put = [&](uint32_t plane_hi, uint32_t off, float w) {
      *reinterpret_cast<float*>(sm + plane_hi + off) = w;
      *reinterpret_cast<float*>(sm + plane_hi + WPLANE + off) = lo_of(w);
    };
    for (int i = tid; i < H * H; i += NTHREADS) {
      const int o = i >> 6, k = i & 63;
      const float w = __ldg(P + F.oW1 + i);
      put(OFF_W1, k128_off(H, o, k), w);
      put(OFF_W1T, k128_off(H, k, o), w);
      put(OFF_DW1, k128_off(H, o, k), __ldg(V + F.oW1 + i));
    }
    for (int o = tid; o < H; o += NTHREADS) vecs[M::DB1 + o] = __ldg(V + F.ob1 + o);
    for (int i = tid; i < m * H; i += NTHREADS) {
      vecs[M::W2 + i] = __ldg(P + F.oW2 + i);
      vecs[M::DW2 + i] = __ldg(V + F.oW2 + i);
    }
    for (int j = tid; j < m; j += NTHREADS) {
      vecs[M::DB2 + j] = __ldg(V + F.ob2 + j);
      const float sg = expf(__ldg(P + F.ols + j));
      const float os = a.out_scale ? __ldg(a.out_scale + j) : 1.f;
      vecs[M::FISHER + j] = os * os * (2.f / (2.f * sg * sg + 1e-8f));     // gaussian_mlp.py:137-138
    }
  }
  fence_async();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tb = tslot;

  unsigned char* pW = sm + OFF_W;
  unsigned char* pY = sm + OFF_Y; unsigned char* pZ = sm + OFF_Z;
  unsigned char* pU = sm + OFF_UT;
  const float* tile_base = a.cache;

  if ((warp & 3) != 3) {
    const int q = warp & 3, half = warp >> 2, blk = half;
    const int r = 32 * q + lane;
    const uint32_t tl = tb + ((uint32_t)(32 * q) << 16);
    const uint32_t pair_bar = 1 + q;
    auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory"); };
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
      // ---- E0: dz1 of this thread's 32 columns is on its way; h1 -> tensor memory, h1 lo plane ------------
      float4 dzr[8];
      {
        const float4* dzp = reinterpret_cast<const float4*>(a.dz1) + (size_t)tile * (ROWS * H / 4) + (size_t)(8 * blk) * ROWS + r;
#pragma unroll
        for (int j = 0; j < 8; ++j) dzr[j] = __ldg(dzp + (size_t)j * ROWS);
      }
      float s1[32];
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
      // ---- E1: dh1 = (1 - h1^2) dz1 -------------------------------------------------------------------------
      {
        uint32_t d[32], l[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v0 = dzr[j].x * s1[4 * j], v1 = dzr[j].y * s1[4 * j + 1], v2 = dzr[j].z * s1[4 * j + 2], v3 = dzr[j].w * s1[4 * j + 3];
          d[4 * j] = __float_as_uint(v0); d[4 * j + 1] = __float_as_uint(v1); d[4 * j + 2] = __float_as_uint(v2); d[4 * j + 3] = __float_as_uint(v3);
          l[4 * j] = __float_as_uint(lo_of(v0)); l[4 * j + 1] = __float_as_uint(lo_of(v1));
          l[4 * j + 2] = __float_as_uint(lo_of(v2)); l[4 * j + 3] = __float_as_uint(lo_of(v3));
        }
        tst32(tl + T_D1 + 32 * blk, d);
        tst32(tl + T_LO1 + 32 * blk, l);
      }
      wait_st();
      fence_before();
      mbar_arrive(&bar[B_E1]);
      // ---- E2: head on the CUDA cores -----------------------------------------------------------------------
      float s2[32];
      float u[MJT];
      {
        float h[32];
        uint32_t d[32];
        tma::mbar_wait(&bar[B_H2], ph);
        load_row32(pY, r, blk, h);
        float2 hv[16];
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
#pragma unroll
          for (int e = 0; e < 4; ++e) s2[4 * i4 + e] = fmaf(-h[4 * i4 + e], h[4 * i4 + e], 1.f);
          hv[2 * i4] = make_float2(h[4 * i4], h[4 * i4 + 1]);
          hv[2 * i4 + 1] = make_float2(h[4 * i4 + 2], h[4 * i4 + 3]);
        }
#pragma unroll
        for (int j = 0; j < MJT; ++j) {
          const float4* v2 = reinterpret_cast<const float4*>(vecs + M::DW2 + j * H + 32 * blk);
          float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 v = v2[i4];
            a0 = __ffma2_rn(hv[2 * i4], make_float2(v.x, v.y), a0);
            a1 = __ffma2_rn(hv[2 * i4 + 1], make_float2(v.z, v.w), a1);
          }
          u[j] = (a0.x + a0.y) + (a1.x + a1.y);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) h[i] = lo_of(h[i]);
        store_row32(pZ, r, blk, h);
        tma::mbar_wait(&bar[B_M2B], ph);
        fence_after();
        tld32w(tl + T_D2 + 32 * blk, d);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 b = *reinterpret_cast<const float4*>(vecs + M::DB1 + 32 * blk + 4 * i4);
          hv[2 * i4] = __fmul2_rn(__fadd2_rn(make_float2(__uint_as_float(d[4 * i4]), __uint_as_float(d[4 * i4 + 1])), make_float2(b.x, b.y)),
                                  make_float2(s2[4 * i4], s2[4 * i4 + 1]));
          hv[2 * i4 + 1] = __fmul2_rn(__fadd2_rn(make_float2(__uint_as_float(d[4 * i4 + 2]), __uint_as_float(d[4 * i4 + 3])), make_float2(b.z, b.w)),
                                      make_float2(s2[4 * i4 + 2], s2[4 * i4 + 3]));
        }
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
          u[j] += (a0.x + a0.y) + (a1.x + a1.y);
        }
      }
      // the two halves of a row swap their partial sums through the u^T plane (half 0's in the lo rows MJT + j, half 1's in the
      // hi rows j, column = timestep), ONE barrier, and both form the same u (the sum is commutative); then half 0 files u in
      // the hi rows and half 1 its lo part in the lo rows, each over the slot it has just read.  gW2 of the previous tile has
      // read the plane (B_GW2 + q, E2c).
#pragma unroll
      for (int j = 0; j < MJT; ++j) *reinterpret_cast<float*>(pU + k128_off(UR, (half == 0 ? MJT : 0) + j, r)) = u[j];
      pair_sync();
#pragma unroll
      for (int j = 0; j < MJT; ++j) {
        const float other = *reinterpret_cast<const float*>(pU + k128_off(UR, (half == 0 ? 0 : MJT) + j, r));
        const float dmu = u[j] + other;
        u[j] = (valid && j < m) ? vecs[M::FISHER + j] * (dmu + vecs[M::DB2 + j]) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < MJT; ++j)
        *reinterpret_cast<float*>(pU + k128_off(UR, (half == 0 ? 0 : MJT) + j, r)) = half == 0 ? u[j] : lo_of(u[j]);
      fence_async();
      mbar_arrive(&bar[B_E2A]);
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
      tma::mbar_wait(&bar[B_GW2 + q], ph);
      {
        store_row32(pY, r, blk, gz);
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = lo_of(gz[i]);
        store_row32(pZ, r, blk, v);
      }
      fence_async();
      mbar_arrive(&bar[B_E2C + q]);
      // ---- E3: gz1 = gh1 (1 - h1^2) -> Y plane (it leaves for the GZ1 image from there) ---------------------
      {
        uint32_t d[32];
        tma::mbar_wait(&bar[B_M5], ph);
        fence_after();
        tld32w(tl + T_D1 + 32 * blk, d);
#pragma unroll
        for (int i = 0; i < 32; ++i) gz[i] = valid ? __uint_as_float(d[i]) * s1[i] : 0.f;
      }
      fence_before();
      prefetch_h1(tile + gridDim.x);
      tma::mbar_wait(&bar[B_GW1 + q], ph);
      tma::mbar_wait(&bar[B_W3 + q], ph);
      store_row32(pY, r, blk, gz);
      fence_async();
      mbar_arrive(&bar[B_E3 + q]);
    }
  } else if (warp == 3) {
    double gb1a = 0.0, gb1b = 0.0, gb2 = 0.0;
    auto load_h2 = [&](int tile) {
      if (lane == 0) {
        tma::mbar_expect(&bar[B_H2], PLANE);
        tma::bulk_g2s(sm + OFF_Y, tile_base + (size_t)tile * (2 * PLANE / 4) + PLANE / 4, PLANE, &bar[B_H2]);
      }
    };
    auto load_h1q = [&](int tile, int qq) {
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
      if (lane == 0 && tile + (int)gridDim.x < ntiles) {
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                     :: "l"(tile_base + (size_t)(tile + gridDim.x) * (2 * PLANE / 4)), "r"(2 * PLANE) : "memory");
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                     :: "l"(a.dz1 + (size_t)(tile + gridDim.x) * (PLANE / 4)), "r"(PLANE) : "memory");
      }
      // u of the tile is complete once every row thread has passed E2a; it is overwritten in the next tile's E2,
      // which no row thread reaches before this warp's B_W3 arrivals below
      tma::mbar_wait(&bar[B_E2A], ph);
      if (lane < MJT) {
        float f[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
        for (int rr = 0; rr < ROWS; rr += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) f[k] += *reinterpret_cast<const float*>(pU + k128_off(UR, lane, rr + k));
        }
        gb2 += (double)((f[0] + f[1]) + (f[2] + f[3]));
      }
      {
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
      const int next = tile + gridDim.x;
#pragma unroll 1
      for (int qq = 0; qq < 3; ++qq) {
        tma::mbar_wait(&bar[B_GW1 + qq], ph);
        if (next < ntiles) load_h1q(next, qq);
      }
      tma::mbar_wait(&bar[B_GW0], ph);       // the gz1 plane has left Y
      if (next < ntiles) load_h2(next);
    }
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    out[F.ob1 + 2 * lane] = (float)gb1a;
    out[F.ob1 + 2 * lane + 1] = (float)gb1b;
    if (lane < m) out[F.ob2 + lane] = (float)gb2;
  } else {
    const uint32_t ID_FWD = idesc_tf32(128, 64, 0, 0);
    const uint32_t ID_GW1 = idesc_tf32(128, 128, 1, 1);
    const uint32_t ID_GW2 = idesc_tf32(128, UR, 1, 0);
    auto wdesc = [&](uint32_t off, int ks) {
      return make_desc(sb + off + (uint32_t)(ks >> 2) * (H * 128u) + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    auto mndesc = [&](uint32_t off, int ks) { return make_desc(sb + off + (uint32_t)ks * 1024u, BLK, 512u, 1u); };
    auto utdesc = [&](int ks) {
      return make_desc(sb + OFF_UT + (uint32_t)(ks >> 2) * (uint32_t)UR * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2u);
    };
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1;
      const uint32_t first = it == 0 ? 0u : 1u;
      tma::mbar_wait(&bar[B_E0], ph);
      fence_after();
      if (elect_one()) {
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
#pragma unroll
        for (int qq = 0; qq < 3; ++qq) {
#pragma unroll
          for (int ks = 4 * qq; ks < 4 * qq + 4; ++ks)
            mma_ss(tb + T_ACC2, mndesc(OFF_Y, ks), utdesc(ks), ID_GW2, first | (ks > 0));
          commit(&bar[B_GW2 + qq]);
        }
      }
      __syncwarp();
      tma::mbar_wait(&bar[B_E2B], ph);
      fence_after();
      if (elect_one()) {
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
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            mma_ss(tb + T_ACC1, mndesc(OFF_Y, 4 * qq + ks), mndesc(OFF_X, 4 * qq + ks), ID_GW1, first | ((qq | ks) > 0));
          commit(&bar[B_GW1 + qq]);
        }
        __syncwarp();
      }
      for (int qq = 0; qq < 3; ++qq) tma::mbar_wait(&bar[B_E3 + qq], ph);
      if (lane == 0) {      // the gz1 plane leaves as it lies; the plane is reusable once the copy has READ it
        bulk_s2g(a.gz1 + (size_t)tile * (PLANE / 4), sm + OFF_Y, PLANE);
        bulk_commit();
        bulk_wait_read();
        mbar_arrive(&bar[B_GW0]);
      }
      __syncwarp();
    }
    if (lane == 0) bulk_wait_all();
  }

  // ---- accumulators -> per-CTA partial sums (W1, W2; the bias sums were written by warp 3) ---------------------
  {
    int nit = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) ++nit;
    if (nit > 0 && warp != 3) tma::mbar_wait(&bar[B_GW0], (nit - 1) & 1);
    fence_after();
    __syncthreads();
    float* stage = reinterpret_cast<float*>(sm);
    constexpr int SW = 128 + UR + 1;
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
#pragma unroll 1
      for (int c0 = 0; c0 < UR; c0 += 16) {
        uint32_t w[16];
        tld16(tl + T_ACC2 + c0, w);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) row[128 + c0 + i] = __uint_as_float(w[i]);
      }
    }
    fence_before();
    __syncthreads();
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    if (nit == 0) {
      for (int i = tid; i < H * H + H; i += NTHREADS) out[F.oW1 + i] = 0.f;
      for (int i = tid; i < m * H + m; i += NTHREADS) out[F.oW2 + i] = 0.f;
    } else {
      auto S = [&](int lane_, int col) { return stage[(size_t)lane_ * SW + col]; };
      for (int i = tid; i < H * H; i += NTHREADS) {
        const int o = i >> 6, k = i & 63;
        out[F.oW1 + i] = (S(o, k) + S(o, 64 + k)) + (S(64 + o, k) + S(64 + o, 64 + k));
      }
      for (int i = tid; i < m * H; i += NTHREADS) {
        const int j = i >> 6, h = i & 63;
        out[F.oW2 + i] = (S(h, 128 + j) + S(h, 128 + MJT + j)) + (S(64 + h, 128 + j) + S(64 + h, 128 + MJT + j));
      }
    }
    __syncthreads();
    if (warp == 7) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tb) : "memory");
  }
}

}  // namespace tcw

// ---- host side (declared in policy_tile.cuh) -----------------------------------------------------------------
bool tcw_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2) {
  static const bool off = getenv("MJRL_B200_FVP_TC") && getenv("MJRL_B200_FVP_TC")[0] == '0';
  if (off || !ctx || tc_fits(ctx, n, m, h1, h2)) return false;
  if (h1 != tc::H || h2 != tc::H || n < 1 || n + 1 > 32 * tcw::MAXKB || m < 1 || m > 24) return false;
  return (size_t)tcw::WideMap<24>::ALLOC <= (size_t)ctx->max_smem_optin &&
         (size_t)tcw::FwdMap<tcw::MAXKB>::ALLOC <= (size_t)ctx->max_smem_optin;
}

static inline int tcw_kb(int n) { return (n + 1 + 31) / 32; }

// cache regions (floats): [h1 | h2 planes][x~ image][DZ1 image][GZ1 image], each per 96-timestep tile
long long tcw_cache_floats(int n, long long N) {
  const long long ntiles = (N + tc::ROWS - 1) / tc::ROWS;
  return ntiles * (2 * (long long)tc::PLANE / 4 + (long long)tcw_kb(n) * (tcw::XBLK / 4) + 2 * (long long)tc::PLANE / 4);
}

int launch_tcw_ximg(mjrl_ctx* ctx, const Args& a, float* cache, cudaStream_t s) {
  const long long ntiles = (a.N + tc::ROWS - 1) / tc::ROWS;
  const int KB = tcw_kb(a.n);
  float* img = cache + ntiles * (2 * (long long)tc::PLANE / 4);
  dim3 grid((unsigned)ntiles, (unsigned)KB);
  tcw::ximg_kernel<<<grid, 256, 0, s>>>(a.obs, a.in_shift, a.in_scale, img, a.N, a.n, KB);
  return mjrl_check_launch("ximg_kernel");
}

template <typename K>
static int tcw_attr(K kern, uint32_t bytes, const char* what) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute(%s): %s", what, cudaGetErrorString(e)); return (int)e; }
  return 0;
}

TcwRegions tcw_regions(float* cache, int n, long long N) {
  const long long ntiles = (N + tc::ROWS - 1) / tc::ROWS;
  TcwRegions r;
  r.KB = tcw_kb(n);
  r.h12 = cache;
  r.ximg = r.h12 + ntiles * (2 * (long long)tc::PLANE / 4);
  r.z1 = r.ximg + ntiles * (long long)r.KB * (tcw::XBLK / 4);
  r.gz1 = r.z1 + ntiles * ((long long)tc::PLANE / 4);
  return r;
}

static int tcw_aux_attrs() {
  static bool attr = false;
  if (attr) return 0;
  int rc = tcw_attr(tcw::l0_fwd_kernel<1>, tcw::FwdMap<1>::ALLOC, "l0_fwd<1>");
  if (!rc) rc = tcw_attr(tcw::l0_fwd_kernel<2>, tcw::FwdMap<2>::ALLOC, "l0_fwd<2>");
  if (!rc) rc = tcw_attr(tcw::l0_fwd_kernel<3>, tcw::FwdMap<3>::ALLOC, "l0_fwd<3>");
  if (!rc) rc = tcw_attr(tcw::l0_fwd_kernel<4>, tcw::FwdMap<4>::ALLOC, "l0_fwd<4>");
  if (!rc) rc = tcw_attr(tcw::l0_grad_kernel<1>, tcw::GradMap<1>::ALLOC, "l0_grad<1>");
  if (!rc) rc = tcw_attr(tcw::l0_grad_kernel<2>, tcw::GradMap<2>::ALLOC, "l0_grad<2>");
  if (!rc) rc = tcw_attr(tcw::l0_grad_kernel<3>, tcw::GradMap<3>::ALLOC, "l0_grad<3>");
  if (!rc) rc = tcw_attr(tcw::l0_grad_kernel<4>, tcw::GradMap<4>::ALLOC, "l0_grad<4>");
  if (!rc) attr = true;
  return rc;
}

// z1 image = [x~ 1] [W0 | b0]^T with W0, b0 the first-layer rows of the flat vector `weights` (the CG direction for a
// product, the parameters for a forward pass)
int launch_tcw_l0_fwd(mjrl_ctx* ctx, const TcwRegions& r, const float* weights, long long N, int n, int m, int grid,
                      const int32_t* stop_flag, cudaStream_t s) {
  int rc = tcw_aux_attrs();
  if (rc) return rc;
  tcw::AuxArgs x = {};
  x.ximg = r.ximg; x.vec = weights; x.dz1 = r.z1; x.stop_flag = stop_flag; x.N = N; x.n = n; x.m = m; x.KB = r.KB;
  switch (r.KB) {
    case 1: tcw::l0_fwd_kernel<1><<<grid, tcw::AUX_THREADS, tcw::FwdMap<1>::ALLOC, s>>>(x); break;
    case 2: tcw::l0_fwd_kernel<2><<<grid, tcw::AUX_THREADS, tcw::FwdMap<2>::ALLOC, s>>>(x); break;
    case 3: tcw::l0_fwd_kernel<3><<<grid, tcw::AUX_THREADS, tcw::FwdMap<3>::ALLOC, s>>>(x); break;
    default: tcw::l0_fwd_kernel<4><<<grid, tcw::AUX_THREADS, tcw::FwdMap<4>::ALLOC, s>>>(x); break;
  }
  return mjrl_check_launch("l0_fwd_kernel");
}

// W0 / b0 slices of partials[grid][dstride] = gz1^T [x~ 1] from the GZ1 image
int launch_tcw_l0_grad(mjrl_ctx* ctx, const TcwRegions& r, float* partials, int dstride, long long N, int n, int m,
                       int grid, const int32_t* stop_flag, cudaStream_t s) {
  int rc = tcw_aux_attrs();
  if (rc) return rc;
  tcw::AuxArgs x = {};
  x.ximg = r.ximg; x.gz1 = r.gz1; x.partials = partials; x.stop_flag = stop_flag;
  x.N = N; x.n = n; x.m = m; x.dstride = dstride; x.KB = r.KB;
  switch (r.KB) {
    case 1: tcw::l0_grad_kernel<1><<<grid, tcw::AUX_THREADS, tcw::GradMap<1>::ALLOC, s>>>(x); break;
    case 2: tcw::l0_grad_kernel<2><<<grid, tcw::AUX_THREADS, tcw::GradMap<2>::ALLOC, s>>>(x); break;
    case 3: tcw::l0_grad_kernel<3><<<grid, tcw::AUX_THREADS, tcw::GradMap<3>::ALLOC, s>>>(x); break;
    default: tcw::l0_grad_kernel<4><<<grid, tcw::AUX_THREADS, tcw::GradMap<4>::ALLOC, s>>>(x); break;
  }
  return mjrl_check_launch("l0_grad_kernel");
}

int launch_fvp_tcw(mjrl_ctx* ctx, Args& a, float* cache, int* grid_out, cudaStream_t s) {
  if (!cache || !tcw_fits(ctx, a.n, a.m, a.h1, a.h2)) return -1;
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
    int rc = tcw_attr(tcw::fvp_wide_kernel<8>, tcw::WideMap<8>::ALLOC, "fvp_wide<8>");
    if (!rc) rc = tcw_attr(tcw::fvp_wide_kernel<16>, tcw::WideMap<16>::ALLOC, "fvp_wide<16>");
    if (!rc) rc = tcw_attr(tcw::fvp_wide_kernel<24>, tcw::WideMap<24>::ALLOC, "fvp_wide<24>");
    if (rc) return rc;
    attr = true;
  }
  const TcwRegions r = tcw_regions(cache, a.n, a.N);
  int rc = launch_tcw_l0_fwd(ctx, r, a.vec, a.N, a.n, a.m, grid, a.stop_flag, s);
  if (rc) return rc;

  tcw::WideArgs w = {};
  w.params = a.params; w.vec = a.vec; w.cache = r.h12; w.dz1 = r.z1; w.gz1 = r.gz1; w.out_scale = a.out_scale;
  w.partials = a.partials; w.stop_flag = a.stop_flag; w.N = a.N; w.n = a.n; w.m = a.m; w.dstride = a.dstride;
  if (a.m <= 8) tcw::fvp_wide_kernel<8><<<grid, tcw::NTHREADS, tcw::WideMap<8>::ALLOC, s>>>(w);
  else if (a.m <= 16) tcw::fvp_wide_kernel<16><<<grid, tcw::NTHREADS, tcw::WideMap<16>::ALLOC, s>>>(w);
  else tcw::fvp_wide_kernel<24><<<grid, tcw::NTHREADS, tcw::WideMap<24>::ALLOC, s>>>(w);
  rc = mjrl_check_launch("fvp_wide_kernel");
  if (rc) return rc;
  return launch_tcw_l0_grad(ctx, r, a.partials, a.dstride, a.N, a.n, a.m, grid, a.stop_flag, s);
}

}  // namespace pol
