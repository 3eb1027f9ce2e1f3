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

// =========================================================================================
// Tensor-core path: error-compensated 3xTF32 on mma.sync.m16n8k8 (fp32 accumulate).
//   a*b ~= a_lo*b_hi + a_hi*b_lo + a_hi*b_hi,  x_hi = tf32(x), x_lo = tf32(x - x_hi)
// keeps ~2^-21 relative error per product (plain TF32 is 2^-11 and breaks the 1e-4 parity gate).
// Measured on B200 (profiles/probes/mma_probe.cu): 279 TFLOP/s TF32 mma.sync vs 72 TFLOP/s FFMA.
//
// Fragment ownership (PTX ISA, m16n8k8 .tf32):  g = lane>>2, c = lane&3
//   A (16x8 row):  a0 (g, c)   a1 (g+8, c)   a2 (g, c+4)   a3 (g+8, c+4)
//   B (8x8  col):  b0 (k=c, n=g)   b1 (k=c+4, n=g)
//   C (16x8):      c0 (g, 2c)  c1 (g, 2c+1)  c2 (g+8, 2c)  c3 (g+8, 2c+1)
// Shared-memory strides are chosen so every fragment load is bank-conflict free:
// activations [t][lda] with lda = 4 (mod 8) words, k-major weights [k][ldw] with ldw = 8 (mod 32).
// =========================================================================================
namespace pol {

struct Frag4 { uint32_t hi[4], lo[4]; };
struct Frag2 { uint32_t hi[2], lo[2]; };

// Split with full-rate integer ALU ops instead of cvt.rna.tf32.f32 (a conversion-pipe instruction
// whose throughput, not the tensor pipe, bounded the MMA loops): adding half a TF32 ulp to the bit
// pattern and clearing the 13 low mantissa bits rounds to nearest (ties away).  lo = x - hi is
// exact and has <= 12 significant bits; the tensor core ignores the low 13 bits of a .tf32
// operand, i.e. truncates it to 11 (one bit, symmetric around zero: ~2^-23 |x|).
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}

// tanh to ~1.5 ulp without branches: odd minimax polynomial x + x^3 Q(x^2) on |x| < 0.6
// (7.8e-8 relative, fitted offline), 1 - 2/(exp(2|x|) + 1) through ex2.approx / rcp.approx
// beyond.  tanhf() costs ~2.5x the instructions here because both of its branches execute in
// a warp whose lanes straddle its threshold.
__device__ __forceinline__ float tanh_fast(float x) {
  const float ax = fabsf(x), u = x * x;
  float q = fmaf(u, -0.005909046623855829f, 0.02080880105495453f);
  q = fmaf(q, u, -0.05378798395395279f);
  q = fmaf(q, u, 0.13331978023052216f);
  q = fmaf(q, u, -0.33333301544189453f);
  const float p = fmaf(x * u, q, x);
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(ax * 2.8853900817779268f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.f));
  const float t = copysignf(fmaf(-2.f, r, 1.f), x);
  return ax < 0.6f ? p : t;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__device__ __forceinline__ void mma3(float (&d)[4], const Frag4& a, const Frag2& b) {
  mma_tf32(d, a.lo, b.hi);  // small terms first
  mma_tf32(d, a.hi, b.lo);
  mma_tf32(d, a.hi, b.hi);
}

// A[m][k] = S[m][k], S row-major with stride ld; p = &S[m0][k0]
__device__ __forceinline__ void load_a(Frag4& f, const float* __restrict__ p, int ld, int g, int c) {
  split_tf32(p[g * ld + c], f.hi[0], f.lo[0]);
  split_tf32(p[(g + 8) * ld + c], f.hi[1], f.lo[1]);
  split_tf32(p[g * ld + c + 4], f.hi[2], f.lo[2]);
  split_tf32(p[(g + 8) * ld + c + 4], f.hi[3], f.lo[3]);
}
// Weight-gradient operands (k walks the rows = timesteps of an activation matrix whose row stride
// is = 4 mod 8 words).  The MMA's k index j is mapped to row 2j (j < 4) / 2(j-4)+1 (j >= 4) of the
// 8-row block -- any permutation of k is fine as long as A and B use the same one -- so that the
// four k values a warp touches per load sit 8 banks apart: 8c + g covers all 32 banks.
// A[m][k] = S[perm(k)][m]; p = &S[t0][m0]
__device__ __forceinline__ void load_a_t(Frag4& f, const float* __restrict__ p, int ld, int g, int c) {
  split_tf32(p[(2 * c) * ld + g], f.hi[0], f.lo[0]);
  split_tf32(p[(2 * c) * ld + g + 8], f.hi[1], f.lo[1]);
  split_tf32(p[(2 * c + 1) * ld + g], f.hi[2], f.lo[2]);
  split_tf32(p[(2 * c + 1) * ld + g + 8], f.hi[3], f.lo[3]);
}
// B[k][n] = S[perm(k)][n]; p = &S[t0][n0]
__device__ __forceinline__ void load_b_t(Frag2& f, const float* __restrict__ p, int ld, int g, int c) {
  split_tf32(p[(2 * c) * ld + g], f.hi[0], f.lo[0]);
  split_tf32(p[(2 * c + 1) * ld + g], f.hi[1], f.lo[1]);
}
// B[k][n] = S[k][n]; p = &S[k0][n0]
__device__ __forceinline__ void load_b(Frag2& f, const float* __restrict__ p, int ld, int g, int c) {
  split_tf32(p[c * ld + g], f.hi[0], f.lo[0]);
  split_tf32(p[(c + 4) * ld + g], f.hi[1], f.lo[1]);
}

// Layer product on a 32(t) x 16(f) warp tile:  z += A Wt ;  dz += A Vt (+ D Wt)
//   A, D: smem activations, pointers at the warp's first row; Wt, Vt: k-major weights at the warp's
//   first column.  K multiple of 8.
template <bool DUAL, bool HAS_D>
__device__ __forceinline__ void mma_layer(float (&z)[2][2][4], float (&dz)[2][2][4],
                                          const float* __restrict__ A, const float* __restrict__ D, int lda,
                                          const float* __restrict__ Wt, const float* __restrict__ Vt, int ldw,
                                          int K, int g, int c) {
#pragma unroll 2
  for (int k0 = 0; k0 < K; k0 += 8) {
    Frag4 a[2], d[2];
    Frag2 w[2], v[2];
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      load_a(a[mi], A + mi * 16 * lda + k0, lda, g, c);
      if (HAS_D) load_a(d[mi], D + mi * 16 * lda + k0, lda, g, c);
    }
#pragma unroll
    for (int ni = 0; ni < 2; ++ni) {
      load_b(w[ni], Wt + k0 * ldw + ni * 8, ldw, g, c);
      if (DUAL) load_b(v[ni], Vt + k0 * ldw + ni * 8, ldw, g, c);
    }
    // Issue order: one split term at a time across ALL accumulators, so that two MMAs on the same
    // accumulator are 4-12 instructions apart (a warp issues in order; back-to-back dependent
    // mma.sync stall for the full MMA latency and leave the tensor pipe half empty).
#pragma unroll
    for (int term = 0; term < 3; ++term) {
#pragma unroll
      for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int ni = 0; ni < 2; ++ni)
          mma_tf32(z[mi][ni], term == 0 ? a[mi].lo : a[mi].hi, term == 1 ? w[ni].lo : w[ni].hi);
      if (DUAL) {
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
          for (int ni = 0; ni < 2; ++ni)
            mma_tf32(dz[mi][ni], term == 0 ? a[mi].lo : a[mi].hi, term == 1 ? v[ni].lo : v[ni].hi);
      }
      if (HAS_D) {
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
          for (int ni = 0; ni < 2; ++ni)
            mma_tf32(dz[mi][ni], term == 0 ? d[mi].lo : d[mi].hi, term == 1 ? w[ni].lo : w[ni].hi);
      }
    }
  }
}

// Weight-gradient fragments: acc[j] (16 x 8 each, rows = A columns m0.., cols = B columns n0 + 8j..)
//   += sum_t A_s[t][m0 + .] * B_s[t][n0 + 8j + .]   over t in [0, TS), TS multiple of 8
template <int NFR>
__device__ __forceinline__ void mma_wgrad(float (&acc)[NFR][4], const float* __restrict__ A_s, int lda,
                                          const float* __restrict__ B_s, int ldb, int m0, int n0, int nvalid,
                                          int TS, int g, int c) {
  // Two independent accumulator sets (even / odd timestep blocks) double the number of MMAs
  // that can be in flight between two uses of the same accumulator; summed once per call.
  float odd[NFR][4];
#pragma unroll
  for (int j = 0; j < NFR; ++j) odd[j][0] = odd[j][1] = odd[j][2] = odd[j][3] = 0.f;
  for (int t0 = 0; t0 < TS; t0 += 16) {
    Frag4 a0, a1;
    Frag2 b0[NFR], b1[NFR];
    load_a_t(a0, A_s + t0 * lda + m0, lda, g, c);
    load_a_t(a1, A_s + (t0 + 8) * lda + m0, lda, g, c);
#pragma unroll
    for (int j = 0; j < NFR; ++j)
      if (j < nvalid) {
        load_b_t(b0[j], B_s + t0 * ldb + n0 + 8 * j, ldb, g, c);
        load_b_t(b1[j], B_s + (t0 + 8) * ldb + n0 + 8 * j, ldb, g, c);
      }
#pragma unroll
    for (int term = 0; term < 3; ++term) {  // term-major order: see mma_layer
#pragma unroll
      for (int j = 0; j < NFR; ++j)
        if (j < nvalid) mma_tf32(acc[j], term == 0 ? a0.lo : a0.hi, term == 1 ? b0[j].lo : b0[j].hi);
#pragma unroll
      for (int j = 0; j < NFR; ++j)
        if (j < nvalid) mma_tf32(odd[j], term == 0 ? a1.lo : a1.hi, term == 1 ? b1[j].lo : b1[j].hi);
    }
  }
#pragma unroll
  for (int j = 0; j < NFR; ++j)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[j][e] += odd[j][e];
}

// ---- arguments shared by every policy kernel (policy.cu, fvp_rows.cuh, pol_rows.cuh, fvp_jac.cu) ----
enum Mode { M_FWD = 0, M_EVAL = 1, M_GRAD = 2, M_FVP = 3 };

struct Args {
  const float* params;
  const float* vec;
  const float* obs;
  const float* act;
  const float* w;
  const float* old_ll;
  const float* old_mean;
  const float* old_log_std;
  float* mean_out;
  float* ll_out;
  const float* in_shift;
  const float* in_scale;
  const float* out_shift;
  const float* out_scale;
  float* partials;    // [grid][dstride]
  double* spartials;  // [grid][2]
  const int32_t* stop_flag;
  long long* phase_cycles;  // debug: [grid][8] cycles per phase (P0,P1,P2,P3a,P3b,P4,P5,P6+loop)
  long long N;
  int n, m, h1, h2, dstride;
  int prefetch;   // 1: double-buffered obs/act staging (Layout::XR/AR present)
  float* cache;   // FVP rows kernel: forward activations in fragment order (fill or use), may be null
  int cache_mode; // 0 none, 1 fill, 2 use, 3 use h1|h2 and the Jacobian rows G1 (fvp_jac.cu), 5 fill in the tile order of fvp_tc.cu
};

// Jacobian-cached FVP (fvp_jac.cu), act_dim <= 4: shape test, extra cache floats after the h1|h2 region,
// the fill pass and the product (-1 when the shape does not fit)
bool jac_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2, int* HP_out, int* NPK_out, size_t* smem_out);
long long jac_cache_floats(int m, int h1, int h2, long long N);
int launch_fvp_jac(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s);
int launch_jac_fill(mjrl_ctx* ctx, const float* params, int n, int m, int h1, int h2, float* cache, long long N,
                    cudaStream_t s);

// tcgen05 / tensor-memory FVP (fvp_tc.cu), hidden 64 x 64, obs_dim <= 15, act_dim <= 8: shape test, floats of its
// tile-ordered copy of the h1|h2 cache, the relayout pass from the fragment-ordered cache and the product
bool tc_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2);
long long tc_cache_floats(long long N);
int launch_tc_fill(mjrl_ctx* ctx, const float* frag_cache, float* tc_cache, long long N, cudaStream_t s);
int launch_fvp_tc(mjrl_ctx* ctx, Args& a, const float* tc_cache, int* grid_out, cudaStream_t s);
// three-pass tcgen05 product for wide observation / action vectors (fvp_tc_wide.cu): shape test, cache floats
// ([h1 | h2 planes][x~ image][DZ1][GZ1]), the x~ image pass (once per gradient pass) and the product
bool tcw_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2);
long long tcw_cache_floats(int n, long long N);
int launch_tcw_ximg(mjrl_ctx* ctx, const Args& a, float* cache, cudaStream_t s);
int launch_fvp_tcw(mjrl_ctx* ctx, Args& a, float* cache, int* grid_out, cudaStream_t s);
struct TcwRegions { float *h12, *ximg, *z1, *gz1; int KB; };   // regions of the wide cache buffer
TcwRegions tcw_regions(float* cache, int n, long long N);
int launch_tcw_l0_fwd(mjrl_ctx* ctx, const TcwRegions& r, const float* weights, long long N, int n, int m, int grid,
                      const int32_t* stop_flag, cudaStream_t s);
int launch_tcw_l0_grad(mjrl_ctx* ctx, const TcwRegions& r, float* partials, int dstride, long long N, int n, int m,
                       int grid, const int32_t* stop_flag, cudaStream_t s);
// gradient / surrogate+KL passes on tcgen05 (pol_tc.cu); `planes` = tile-ordered h1 | h2 region the gradient pass
// fills (null for an evaluation), `wide` = the regions of the three-pass form or null for the fused narrow form
// surrogate / KL evaluation as a forward pass with two 128-timestep tiles in flight (pol_fwd.cu)
int launch_pol_fwd(mjrl_ctx* ctx, Args& a, const TcwRegions* wide, int grid, cudaStream_t s);
template <int MODE>
int launch_pol_tc(mjrl_ctx* ctx, Args& a, float* planes, const TcwRegions* wide, int* grid_out, cudaStream_t s);

}  // namespace pol
