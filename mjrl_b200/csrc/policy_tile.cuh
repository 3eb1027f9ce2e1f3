// Tile engine shared by the fused Gaussian-MLP kernels (forward / eval / gradient / FVP).
//
// One CTA of 256 threads works on a tile of TS timesteps.  All activations of the tile live
// in shared memory as row-major [t][feature] matrices (row stride = features + 4 floats), the
// layout obs already has in HBM.  Products are register-tiled: a thread owns RT=4 timesteps x
// 4 features of a layer output and walks k in steps of 4 with 128-bit shared loads; the
// activation operand is a warp-broadcast load (2 distinct rows per warp), the weight operand a
// contiguous 256 B row segment, so both are bank-conflict free.  Weight gradients
// (sum over timesteps of outer products) are accumulated in registers for the CTA's whole
// persistent loop as 4x4 tiles and reduced once at kernel end.
#pragma once
#include "common.cuh"

namespace pol {

constexpr int NTHREADS = 256;
constexpr int RT = 4;  // timesteps per thread in layer products

__host__ __device__ constexpr int round4(int x) { return (x + 3) & ~3; }

// Offsets into the flat parameter vector (gaussian_mlp.py:38,51-53 order).
struct Flat {
  int n, m, h1, h2;
  int oW0, ob0, oW1, ob1, oW2, ob2, ols, d;
  __host__ __device__ Flat() {}
  __host__ __device__ Flat(int n_, int m_, int h1_, int h2_) : n(n_), m(m_), h1(h1_), h2(h2_) {
    if (h1 > 0) {
      oW0 = 0; ob0 = h1 * n; oW1 = ob0 + h1; ob1 = oW1 + h2 * h1; oW2 = ob1 + h2;
      ob2 = oW2 + m * h2; ols = ob2 + m; d = ols + m;
    } else {  // affine: fc0 is (m, n)
      oW0 = 0; ob0 = m * n; oW1 = ob1 = oW2 = ob2 = 0; ols = ob0 + m; d = ols + m;
    }
  }
};

// ---------------------------------------------------------------------------------------
// acc[r][c] += sum_k A[row0 + r][k] * Wt[k][col0 + c]   (K multiple of 4)
// A: smem [TS][lda] activations; Wt: smem [K][ldw] weights stored k-major.
template <int R>
__device__ __forceinline__ void mm_acc(float (&acc)[R][4], const float* __restrict__ A, int lda,
                                       const float* __restrict__ Wt, int ldw, int K) {
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float4 w0 = *reinterpret_cast<const float4*>(Wt + (k + 0) * ldw);
    float4 w1 = *reinterpret_cast<const float4*>(Wt + (k + 1) * ldw);
    float4 w2 = *reinterpret_cast<const float4*>(Wt + (k + 2) * ldw);
    float4 w3 = *reinterpret_cast<const float4*>(Wt + (k + 3) * ldw);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float4 a = *reinterpret_cast<const float4*>(A + r * lda + k);
      acc[r][0] = fmaf(a.x, w0.x, acc[r][0]); acc[r][1] = fmaf(a.x, w0.y, acc[r][1]);
      acc[r][2] = fmaf(a.x, w0.z, acc[r][2]); acc[r][3] = fmaf(a.x, w0.w, acc[r][3]);
      acc[r][0] = fmaf(a.y, w1.x, acc[r][0]); acc[r][1] = fmaf(a.y, w1.y, acc[r][1]);
      acc[r][2] = fmaf(a.y, w1.z, acc[r][2]); acc[r][3] = fmaf(a.y, w1.w, acc[r][3]);
      acc[r][0] = fmaf(a.z, w2.x, acc[r][0]); acc[r][1] = fmaf(a.z, w2.y, acc[r][1]);
      acc[r][2] = fmaf(a.z, w2.z, acc[r][2]); acc[r][3] = fmaf(a.z, w2.w, acc[r][3]);
      acc[r][0] = fmaf(a.w, w3.x, acc[r][0]); acc[r][1] = fmaf(a.w, w3.y, acc[r][1]);
      acc[r][2] = fmaf(a.w, w3.z, acc[r][2]); acc[r][3] = fmaf(a.w, w3.w, acc[r][3]);
    }
  }
}

// Tangent product sharing operand loads with the primal:
//   accz += A*Wt ;  accd += A*Vt + D*Wt       (D may be nullptr: first layer, dx~ = 0)
template <int R, bool HAS_D>
__device__ __forceinline__ void mm_acc_dual(float (&accz)[R][4], float (&accd)[R][4],
                                            const float* __restrict__ A, const float* __restrict__ D,
                                            int lda, const float* __restrict__ Wt,
                                            const float* __restrict__ Vt, int ldw, int K) {
  for (int k = 0; k < K; k += 4) {
    float4 w[4], v[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      w[c] = *reinterpret_cast<const float4*>(Wt + (k + c) * ldw);
      v[c] = *reinterpret_cast<const float4*>(Vt + (k + c) * ldw);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float4 a4 = *reinterpret_cast<const float4*>(A + r * lda + k);
      float4 d4 = HAS_D ? *reinterpret_cast<const float4*>(D + r * lda + k) : make_float4(0, 0, 0, 0);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float dd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        accz[r][0] = fmaf(a[c], w[c].x, accz[r][0]); accz[r][1] = fmaf(a[c], w[c].y, accz[r][1]);
        accz[r][2] = fmaf(a[c], w[c].z, accz[r][2]); accz[r][3] = fmaf(a[c], w[c].w, accz[r][3]);
        accd[r][0] = fmaf(a[c], v[c].x, accd[r][0]); accd[r][1] = fmaf(a[c], v[c].y, accd[r][1]);
        accd[r][2] = fmaf(a[c], v[c].z, accd[r][2]); accd[r][3] = fmaf(a[c], v[c].w, accd[r][3]);
        if (HAS_D) {
          accd[r][0] = fmaf(dd[c], w[c].x, accd[r][0]); accd[r][1] = fmaf(dd[c], w[c].y, accd[r][1]);
          accd[r][2] = fmaf(dd[c], w[c].z, accd[r][2]); accd[r][3] = fmaf(dd[c], w[c].w, accd[r][3]);
        }
      }
    }
  }
}

// Weight-gradient tile: acc[a][b] += sum_{t = t_begin; t < TS; t += t_step} A[t][a0+a] * B[t][b0+b]
__device__ __forceinline__ void wgrad_acc(float (&acc)[16], const float* __restrict__ A, int lda,
                                          const float* __restrict__ B, int ldb, int a0, int b0,
                                          int t_begin, int t_step, int TS) {
  const float* pa = A + a0;
  const float* pb = B + b0;
#pragma unroll 4
  for (int t = t_begin; t < TS; t += t_step) {
    float4 a4 = *reinterpret_cast<const float4*>(pa + t * lda);
    float4 b4 = *reinterpret_cast<const float4*>(pb + t * ldb);
    const float a[4] = {a4.x, a4.y, a4.z, a4.w};
    const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
    for (int x = 0; x < 4; ++x)
#pragma unroll
      for (int y = 0; y < 4; ++y) acc[x * 4 + y] = fmaf(a[x], b[y], acc[x * 4 + y]);
  }
}

// Ownership of 4x4 weight-gradient tiles by threads: `ntiles` tiles over NTHREADS threads.
// ntiles <= NTHREADS: G = NTHREADS / ntiles groups split the timesteps (thread -> tile tid % ntiles,
// group tid / ntiles); ntiles > NTHREADS: a thread owns tiles tid, tid + NTHREADS, ... (G = 1).
struct TileOwner {
  int tile, grp, G;
  bool active;
  __device__ TileOwner(int ntiles, int tid, int slot) {
    if (ntiles <= NTHREADS) {
      G = NTHREADS / ntiles;
      tile = tid % ntiles; grp = tid / ntiles;
      active = slot == 0 && grp < G;
    } else {
      G = 1; grp = 0; tile = tid + slot * NTHREADS;
      active = tile < ntiles;
    }
  }
};

}  // namespace pol
