// Fisher-vector product, "rows in registers" variant (included by policy.cu, namespace pol).
//
// Same math as mlp_kernel<.., M_FVP, ..> (npg_cg.py:55-74 replaced by the analytic Gauss-Newton
// product), different dataflow.  The CTA is two independent HALVES of 4 warps; a half walks its own
// tiles of 64 timesteps and synchronises only with itself (named barriers), so the two warps that
// share an SM sub-partition drift apart and one feeds the tensor pipe while the other is in an
// epilogue or at a barrier.  Warp w of a half owns rows 16w..16w+15 of the half's tile and keeps the
// WHOLE forward / tangent / backward chain of those rows in registers:
//
//   The C fragment of mma.m16n8k8 holds (row g, cols 2c,2c+1 | row g+8, cols 2c,2c+1).  Read as an
//   A fragment with the k-slot map  slot c <-> col 2c,  slot c+4 <-> col 2c+1  it IS the next
//   layer's A operand (any permutation of k is fine when B uses the same one), and under that map
//   a B fragment (k-slots c, c+4 of output n = g) is the adjacent pair W[n][8j+2c], W[n][8j+2c+1] of
//   the weight matrix in its natural [out][in] layout: ONE 8-byte shared load.
//
// So activations never round-trip through shared memory and the layers need no CTA barrier; shared
// memory holds only the weights (natural layout, row stride = 8 mod 32 words: conflict-free 8-byte
// fragment loads) and, for the weight gradients  gW[o][i] = sum_t gz[t][o] a[t][i]  that contract
// over timesteps of ALL warps, transposed planes [feature][t] written from the C fragments; their
// fragments are again adjacent pairs (t, t+1).  Bias gradients ride along as a ones column.
// 3xTF32 error compensation as in policy_tile.cuh; every split happens in registers.
#pragma once

namespace pol {
namespace rows {

constexpr int TSR = 64;         // timesteps per half-CTA tile (4 warps x 16 rows)
constexpr int LDT = TSR + 8;    // row stride of the transposed planes (= 8 mod 32)
constexpr int HTHREADS = 128;   // threads of a half
__host__ __device__ constexpr int ldx_of(int NP) { return ((NP + 8) % 16 == 8) ? NP + 8 : NP + 16; }

struct RLayout {  // offsets in floats
  int NP, LDX, LDW;
  int X, XT, PA, PB, UT, XR, W0, V0, W1, V1, W2, V2, W2T, b0, vb0, b1, vb1, vb2, fisher, ishift, iden;
  int half_stride, total;
  __host__ __device__ RLayout(int HP, int NPK, int n) {
    NP = 8 * NPK; LDX = ldx_of(NP); LDW = HP + 8;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 3) & ~3; return r; };
    // per-half regions (half 1 at + half_stride); PA|PB of half 0 double as the [d] staging vector
    PA = take(HP * LDT); PB = take(HP * LDT);
    X = take(TSR * LDX); XT = take(NP * LDT); UT = take(8 * LDT);
    XR = take(2 * TSR * n);
    half_stride = o;
    o *= 2;
    // weights: tf32 hi plane immediately followed by the lo plane (split once at staging)
    W0 = take(2 * HP * LDX); V0 = take(2 * HP * LDX);
    W1 = take(2 * HP * LDW); V1 = take(2 * HP * LDW);
    W2 = take(2 * 8 * LDW); V2 = take(2 * 8 * LDW); W2T = take(2 * HP * 8);
    b0 = take(HP); vb0 = take(HP); b1 = take(HP); vb1 = take(HP); vb2 = take(8); fisher = take(8);
    ishift = take(NP); iden = take(NP);
    total = o;
  }
};

// float2 (adjacent pair) -> B fragment (k-slot c = first, c+4 = second)
__device__ __forceinline__ void split_b(Frag2& f, const float2 v) {
  split_tf32(v.x, f.hi[0], f.lo[0]);
  split_tf32(v.y, f.hi[1], f.lo[1]);
}
// pre-split weight planes: hi pair at p, lo pair at p + plane
__device__ __forceinline__ void load_b(Frag2& f, const float* __restrict__ p, int plane) {
  const float2 h = *reinterpret_cast<const float2*>(p);
  const float2 l = *reinterpret_cast<const float2*>(p + plane);
  f.hi[0] = __float_as_uint(h.x); f.hi[1] = __float_as_uint(h.y);
  f.lo[0] = __float_as_uint(l.x); f.lo[1] = __float_as_uint(l.y);
}
// same from two rows (k-slots c / c+4 in different rows of the plane): four 4-byte loads
__device__ __forceinline__ void load_b_rows(Frag2& f, const float* __restrict__ p, int row_stride, int plane) {
  f.hi[0] = __float_as_uint(p[0]); f.hi[1] = __float_as_uint(p[row_stride]);
  f.lo[0] = __float_as_uint(p[plane]); f.lo[1] = __float_as_uint(p[plane + row_stride]);
}
// C fragment (or two adjacent pairs of rows g / g+8) -> A fragment under the slot map
__device__ __forceinline__ void split_a(Frag4& f, float r0c0, float r0c1, float r1c0, float r1c1) {
  split_tf32(r0c0, f.hi[0], f.lo[0]);   // a0 (g,   slot c)
  split_tf32(r1c0, f.hi[1], f.lo[1]);   // a1 (g+8, slot c)
  split_tf32(r0c1, f.hi[2], f.lo[2]);   // a2 (g,   slot c+4)
  split_tf32(r1c1, f.hi[3], f.lo[3]);   // a3 (g+8, slot c+4)
}
__device__ __forceinline__ void split_a(Frag4& f, const float (&cfrag)[4]) {
  split_a(f, cfrag[0], cfrag[1], cfrag[2], cfrag[3]);
}
// one 3xTF32 term (0: lo*hi, 1: hi*lo, 2: hi*hi) of acc += a * b
__device__ __forceinline__ void mma_term(float (&acc)[4], const Frag4& a, const Frag2& b, int term) {
  mma_tf32(acc, term == 0 ? a.lo : a.hi, term == 1 ? b.lo : b.hi);
}

__device__ __forceinline__ void half_sync(int half) {
  asm volatile("bar.sync %0, %1;" :: "r"(half + 1), "n"(HTHREADS) : "memory");
}

// float2 slots of the activation cache per (tile, warp): [layer 0/1][block nf][row g / g+8][lane]
__device__ __forceinline__ float2 ld_cache(const float2* p) {
  float2 v;
  asm("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
  return v;
}

// USE_CACHE = false: compute h1, h2 (and store them in fragment order when a.cache is set);
// USE_CACHE = true : read them back -- during CG the policy is fixed, only the vector changes, so the
// forward products and their tanh are done once per update instead of once per iteration.
template <int HP, int NPK, bool USE_CACHE>
__global__ void __launch_bounds__(NTHREADS, 1) fvp_rows_kernel(const Args a) {
  constexpr int NF = HP / 8;          // 8-wide feature blocks of a hidden layer
  constexpr int MT = HP / 16;         // 16-row bands of a weight gradient
  constexpr int NP = 8 * NPK, LDX = ldx_of(NP), LDW = HP + 8;
  constexpr int NW1 = MT * NF / 4;    // gW1 fragments per warp of a half, all in one band
  constexpr int NW0 = (MT * NPK + 3) / 4;
  constexpr int NW2 = (NF + 3) / 4;
  constexpr int G = 2;                // n-blocks whose MMAs are interleaved term by term
  static_assert(NF % NW1 == 0 && NF % G == 0, "fragment ownership");
  extern __shared__ __align__(16) float sm[];
  __shared__ uint64_t bar[2][2];
  if (a.stop_flag && *a.stop_flag) return;

  const RLayout L(HP, NPK, a.n);
  const Flat F(a.n, a.m, a.h1, a.h2);
  const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int half = tid >> 7, htid = tid & (HTHREADS - 1), hw = htid >> 5;   // half, thread / warp in half
  const int n = a.n, m = a.m;
  float* hs = sm + (size_t)half * L.half_stride;
  float* sX = hs + L.X; float* sXT = hs + L.XT; float* sPA = hs + L.PA; float* sPB = hs + L.PB;
  float* sUT = hs + L.UT; float* sXR = hs + L.XR;
  const float* sW0 = sm + L.W0; const float* sV0 = sm + L.V0; const float* sW1 = sm + L.W1;
  const float* sV1 = sm + L.V1; const float* sW2 = sm + L.W2;
  const float* sV2 = sm + L.V2; const float* sW2T = sm + L.W2T;
  constexpr int PL0 = HP * LDX, PL1 = HP * LDW, PL2 = 8 * LDW, PL2T = HP * 8;   // hi -> lo plane distance

  // ---- one-time staging: zero everything, then weights in their natural [out][in] layout ------
  for (int i = tid; i < L.total; i += NTHREADS) sm[i] = 0.f;
  if (tid == 0) {
    tma::mbar_init(&bar[0][0], 1); tma::mbar_init(&bar[0][1], 1);
    tma::mbar_init(&bar[1][0], 1); tma::mbar_init(&bar[1][1], 1);
    tma::mbar_init_fence();
  }
  __syncthreads();
  {
    const float* P = a.params; const float* V = a.vec;
    auto put = [&](int at, int plane, float w) {
      uint32_t hi, lo;
      split_tf32(w, hi, lo);
      sm[at] = __uint_as_float(hi); sm[at + plane] = __uint_as_float(lo);
    };
    for (int i = tid; i < F.h1 * n; i += NTHREADS) {
      const int o = i / n, k = i - o * n;
      put(L.W0 + o * LDX + k, PL0, __ldg(P + F.oW0 + i));
      put(L.V0 + o * LDX + k, PL0, __ldg(V + F.oW0 + i));
    }
    for (int i = tid; i < F.h2 * F.h1; i += NTHREADS) {
      const int o = i / F.h1, k = i - o * F.h1;
      put(L.W1 + o * LDW + k, PL1, __ldg(P + F.oW1 + i));
      put(L.V1 + o * LDW + k, PL1, __ldg(V + F.oW1 + i));
    }
    for (int i = tid; i < m * F.h2; i += NTHREADS) {
      const int j = i / F.h2, k = i - j * F.h2;
      const float w = __ldg(P + F.oW2 + i);
      put(L.W2 + j * LDW + k, PL2, w);
      put(L.W2T + k * 8 + j, PL2T, w);
      put(L.V2 + j * LDW + k, PL2, __ldg(V + F.oW2 + i));
    }
    for (int i = tid; i < F.h1; i += NTHREADS) { sm[L.b0 + i] = __ldg(P + F.ob0 + i); sm[L.vb0 + i] = __ldg(V + F.ob0 + i); }
    for (int i = tid; i < F.h2; i += NTHREADS) { sm[L.b1 + i] = __ldg(P + F.ob1 + i); sm[L.vb1 + i] = __ldg(V + F.ob1 + i); }
    for (int j = tid; j < m; j += NTHREADS) {
      sm[L.vb2 + j] = __ldg(V + F.ob2 + j);
      const float sg = expf(__ldg(P + F.ols + j));
      const float os = a.out_scale ? __ldg(a.out_scale + j) : 1.f;
      // d^2 KL / d mu^2 = 2/(2 sigma^2 + 1e-8)  (gaussian_mlp.py:137-138), chained through out_scale twice
      sm[L.fisher + j] = os * os * (2.f / (2.f * sg * sg + 1e-8f));
    }
    for (int i = tid; i < NP; i += NTHREADS) {
      sm[L.ishift + i] = (a.in_shift && i < n) ? __ldg(a.in_shift + i) : 0.f;
      sm[L.iden + i] = (a.in_scale && i < n ? __ldg(a.in_scale + i) : 1.f) + 1e-8f;
    }
  }

  // persistent weight-gradient accumulators (C fragments, unique owner per element within a half)
  float accW1[NW1][4], accB1[4], accW0[NW0][4], accW2[NW2][4], gb2acc[2] = {0.f, 0.f};
#pragma unroll
  for (int j = 0; j < NW1; ++j) accW1[j][0] = accW1[j][1] = accW1[j][2] = accW1[j][3] = 0.f;
#pragma unroll
  for (int j = 0; j < NW0; ++j) accW0[j][0] = accW0[j][1] = accW0[j][2] = accW0[j][3] = 0.f;
#pragma unroll
  for (int j = 0; j < NW2; ++j) accW2[j][0] = accW2[j][1] = accW2[j][2] = accW2[j][3] = 0.f;
  accB1[0] = accB1[1] = accB1[2] = accB1[3] = 0.f;
  const int mt1 = (hw * NW1) / NF, nf1 = (hw * NW1) % NF;   // gW1 band / first block of this warp
  const bool own_b1 = nf1 == 0;                               // this warp also sums the band's gb1

  const int ntiles = (int)((a.N + TSR - 1) / TSR);
  const int tile_elems = TSR * n;
  const bool base_ok = (reinterpret_cast<uintptr_t>(a.obs) & 15) == 0;
  auto tile_rows = [&](int tile) { return tile + 1 < ntiles ? TSR : (int)(a.N - (long long)tile * TSR); };
  auto bulk_ok = [&](int tile) { return base_ok && ((tile_rows(tile) * n) & 3) == 0; };
  auto start_copy = [&](int tile, int buf) {  // one thread of the half
    if (bulk_ok(tile)) {
      const uint32_t bytes = (uint32_t)(tile_rows(tile) * n) * 4u;
      tma::mbar_expect(&bar[half][buf], bytes);
      tma::bulk_g2s(sXR + (size_t)buf * tile_elems, a.obs + (long long)tile * tile_elems, bytes, &bar[half][buf]);
    }
  };
  __syncthreads();
  const int tile_first = 2 * blockIdx.x + half, tile_step = 2 * gridDim.x;
  if (htid == 0 && tile_first < ntiles) start_copy(tile_first, 0);

  const int rA = 16 * hw + g, rB = rA + 8;   // this thread's two rows of the half's tile
  int it = 0;
  for (int tile = tile_first; tile < ntiles; tile += tile_step, ++it) {
    const int buf = it & 1;
    const long long t0 = (long long)tile * TSR;
    const int rows = tile_rows(tile);
    if (htid == 0 && tile + tile_step < ntiles) start_copy(tile + tile_step, buf ^ 1);
    // ---- P0: raw rows -> transformed X [t][k] and X^T [k][t], ones column at k = n ----------
    {
      float* raw = sXR + (size_t)buf * tile_elems;
      if (bulk_ok(tile)) {
        tma::mbar_wait(&bar[half][buf], (it >> 1) & 1);
      } else {
        stream_rows<HTHREADS>(a.obs + t0 * n, rows * n, n, htid, [&](int r, int k, float v) { raw[r * n + k] = v; });
        half_sync(half);
      }
      const int cnt = rows * n;
      const int dq = HTHREADS / n, dr = HTHREADS - dq * n;
      int r = htid / n, k = htid - r * n;
      for (int i = htid; i < cnt; i += HTHREADS) {
        const float v = (raw[i] - sm[L.ishift + k]) / sm[L.iden + k];
        sX[r * LDX + k] = v;
        sXT[k * LDT + r] = v;
        r += dq; k += dr;
        if (k >= n) { k -= n; ++r; }
      }
      if (rows < TSR)
        for (int i = rows * n + htid; i < tile_elems; i += HTHREADS) {
          const int rr = i / n, kk = i - rr * n;
          sX[rr * LDX + kk] = 0.f; sXT[kk * LDT + rr] = 0.f;
        }
      if (htid < TSR) {
        const float one = htid < rows ? 1.f : 0.f;
        sX[htid * LDX + n] = one; sXT[n * LDT + htid] = one;
      }
    }
    half_sync(half);

    // ---- layer 0: z1 = b0 + X W0^T, dz1 = vb0 + X V0^T ------------------------------------------
    float h1[NF][4], d1[NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 b = *reinterpret_cast<const float2*>(sm + L.b0 + 8 * nf + 2 * c);
      const float2 vb = *reinterpret_cast<const float2*>(sm + L.vb0 + 8 * nf + 2 * c);
      h1[nf][0] = b.x; h1[nf][1] = b.y; h1[nf][2] = b.x; h1[nf][3] = b.y;
      d1[nf][0] = vb.x; d1[nf][1] = vb.y; d1[nf][2] = vb.x; d1[nf][3] = vb.y;
    }
    // this warp's slice of the activation cache
    float2* cslot = reinterpret_cast<float2*>(a.cache) + ((size_t)tile * 4 + hw) * (4 * NF * 32) + lane;
    if (USE_CACHE) {
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        const float2 lo = ld_cache(cslot + (2 * nf) * 32), hi = ld_cache(cslot + (2 * nf + 1) * 32);
        h1[nf][0] = lo.x; h1[nf][1] = lo.y; h1[nf][2] = hi.x; h1[nf][3] = hi.y;
      }
    }
#pragma unroll
    for (int ks = 0; ks < NPK; ++ks) {
      const float2 xa = *reinterpret_cast<const float2*>(sX + rA * LDX + 8 * ks + 2 * c);
      const float2 xb = *reinterpret_cast<const float2*>(sX + rB * LDX + 8 * ks + 2 * c);
      Frag4 ax;
      split_a(ax, xa.x, xa.y, xb.x, xb.y);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bw[G], bv[G];
#pragma unroll
        for (int i = 0; i < G; ++i) {
          if (!USE_CACHE) load_b(bw[i], sW0 + (8 * (n0 + i) + g) * LDX + 8 * ks + 2 * c, PL0);
          load_b(bv[i], sV0 + (8 * (n0 + i) + g) * LDX + 8 * ks + 2 * c, PL0);
        }
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          if (!USE_CACHE) {
#pragma unroll
            for (int i = 0; i < G; ++i) mma_term(h1[n0 + i], ax, bw[i], term);
          }
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(d1[n0 + i], ax, bv[i], term);
        }
      }
    }
#pragma unroll
    for (int nf = 0; nf < NF; ++nf)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float h = USE_CACHE ? h1[nf][e] : tanh_fast(h1[nf][e]);
        h1[nf][e] = h;
        d1[nf][e] = (1.f - h * h) * d1[nf][e];
      }
    if (!USE_CACHE && a.cache) {
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        cslot[(2 * nf) * 32] = make_float2(h1[nf][0], h1[nf][1]);
        cslot[(2 * nf + 1) * 32] = make_float2(h1[nf][2], h1[nf][3]);
      }
    }
    // H1^T plane for gW1
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      float* p0 = sPA + (8 * nf + 2 * c) * LDT;
      p0[rA] = h1[nf][0]; p0[LDT + rA] = h1[nf][1]; p0[rB] = h1[nf][2]; p0[LDT + rB] = h1[nf][3];
    }

    // ---- layer 1: z2 = b1 + H1 W1^T, dz2 = vb1 + H1 V1^T + D1 W1^T ------------------------------
    float h2[NF][4], d2[NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 b = *reinterpret_cast<const float2*>(sm + L.b1 + 8 * nf + 2 * c);
      const float2 vb = *reinterpret_cast<const float2*>(sm + L.vb1 + 8 * nf + 2 * c);
      h2[nf][0] = b.x; h2[nf][1] = b.y; h2[nf][2] = b.x; h2[nf][3] = b.y;
      d2[nf][0] = vb.x; d2[nf][1] = vb.y; d2[nf][2] = vb.x; d2[nf][3] = vb.y;
    }
    if (USE_CACHE) {   // issued here, consumed after the layer's MMAs
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        const float2 lo = ld_cache(cslot + (2 * (NF + nf)) * 32), hi = ld_cache(cslot + (2 * (NF + nf) + 1) * 32);
        h2[nf][0] = lo.x; h2[nf][1] = lo.y; h2[nf][2] = hi.x; h2[nf][3] = hi.y;
      }
    }
#pragma unroll
    for (int ks = 0; ks < NF; ++ks) {
      Frag4 ah, ad;
      split_a(ah, h1[ks]);
      split_a(ad, d1[ks]);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bw[G], bv[G];
#pragma unroll
        for (int i = 0; i < G; ++i) {
          load_b(bw[i], sW1 + (8 * (n0 + i) + g) * LDW + 8 * ks + 2 * c, PL1);
          load_b(bv[i], sV1 + (8 * (n0 + i) + g) * LDW + 8 * ks + 2 * c, PL1);
        }
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          if (!USE_CACHE) {
#pragma unroll
            for (int i = 0; i < G; ++i) mma_term(h2[n0 + i], ah, bw[i], term);
          }
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(d2[n0 + i], ah, bv[i], term);
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(d2[n0 + i], ad, bw[i], term);
        }
      }
    }
#pragma unroll
    for (int nf = 0; nf < NF; ++nf)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float h = USE_CACHE ? h2[nf][e] : tanh_fast(h2[nf][e]);
        h2[nf][e] = h;
        d2[nf][e] = (1.f - h * h) * d2[nf][e];
      }
    if (!USE_CACHE && a.cache) {
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        cslot[(2 * (NF + nf)) * 32] = make_float2(h2[nf][0], h2[nf][1]);
        cslot[(2 * (NF + nf) + 1) * 32] = make_float2(h2[nf][2], h2[nf][3]);
      }
    }

    // ---- head tangent: dy = H2 V2^T + D2 W2^T;  u = fisher * (dy + vb2), zero past N -----------
    float u[4];
    {
      // four accumulators (product x k parity): a single one would chain 6*NF dependent MMAs
      float uu[2][2][4];
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int y = 0; y < 2; ++y) uu[x][y][0] = uu[x][y][1] = uu[x][y][2] = uu[x][y][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < NF; ++ks) {
        Frag4 ah, ad;
        Frag2 bw, bv;
        split_a(ah, h2[ks]);
        split_a(ad, d2[ks]);
        load_b(bw, sW2 + g * LDW + 8 * ks + 2 * c, PL2);
        load_b(bv, sV2 + g * LDW + 8 * ks + 2 * c, PL2);
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          mma_term(uu[ks & 1][0], ah, bv, term);
          mma_term(uu[ks & 1][1], ad, bw, term);
        }
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) u[e] = (uu[0][0][e] + uu[0][1][e]) + (uu[1][0][e] + uu[1][1][e]);
    }
    {
      const float2 fj = *reinterpret_cast<const float2*>(sm + L.fisher + 2 * c);
      const float2 vj = *reinterpret_cast<const float2*>(sm + L.vb2 + 2 * c);
      const bool va = rA < rows, vb = rB < rows;
      u[0] = va ? fj.x * (u[0] + vj.x) : 0.f; u[1] = va ? fj.y * (u[1] + vj.y) : 0.f;
      u[2] = vb ? fj.x * (u[2] + vj.x) : 0.f; u[3] = vb ? fj.y * (u[3] + vj.y) : 0.f;
      float* p0 = sUT + (2 * c) * LDT;
      p0[rA] = u[0]; p0[LDT + rA] = u[1]; p0[rB] = u[2]; p0[LDT + rB] = u[3];
    }

    // ---- back through the head: gz2 = (U W2) * (1 - h2^2) ---------------------------------------
    float gz[NF][4];   // gz2, later gz1
    {
      Frag4 au;
      split_a(au, u);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bw[G];
#pragma unroll
        for (int i = 0; i < G; ++i) {
          load_b(bw[i], sW2T + (8 * (n0 + i) + g) * 8 + 2 * c, PL2T);
          gz[n0 + i][0] = gz[n0 + i][1] = gz[n0 + i][2] = gz[n0 + i][3] = 0.f;
        }
#pragma unroll
        for (int term = 0; term < 3; ++term)
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(gz[n0 + i], au, bw[i], term);
      }
    }
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
#pragma unroll
      for (int e = 0; e < 4; ++e) gz[nf][e] *= (1.f - h2[nf][e] * h2[nf][e]);
      float* p0 = sPB + (8 * nf + 2 * c) * LDT;   // GZ2^T plane for gW1
      p0[rA] = gz[nf][0]; p0[LDT + rA] = gz[nf][1]; p0[rB] = gz[nf][2]; p0[LDT + rB] = gz[nf][3];
    }

    // ---- gz1 = (GZ2 W1) * (1 - h1^2) --------------------------------------------------------------
    {
      float gh[NF][4];
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) gh[nf][0] = gh[nf][1] = gh[nf][2] = gh[nf][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < NF; ++ks) {
        Frag4 ag;
        split_a(ag, gz[ks]);
#pragma unroll
        for (int n0 = 0; n0 < NF; n0 += G) {
          Frag2 bw[G];
#pragma unroll
          for (int i = 0; i < G; ++i)
            load_b_rows(bw[i], sW1 + (8 * ks + 2 * c) * LDW + 8 * (n0 + i) + g, LDW, PL1);   // B[k = o][n = i] = W1[o][i]
#pragma unroll
          for (int term = 0; term < 3; ++term)
#pragma unroll
            for (int i = 0; i < G; ++i) mma_term(gh[n0 + i], ag, bw[i], term);
        }
      }
#pragma unroll
      for (int nf = 0; nf < NF; ++nf)
#pragma unroll
        for (int e = 0; e < 4; ++e) gz[nf][e] = gh[nf][e] * (1.f - h1[nf][e] * h1[nf][e]);
    }
    half_sync(half);   // (A) H1^T, GZ2^T, U^T of the half's warps visible

    // ---- gW1[o][i] += sum_t GZ2[t][o] H1[t][i]   (+ gb1 through a ones column) -----------------
#pragma unroll 2
    for (int k8 = 0; k8 < TSR; k8 += 8) {
      const float2 a0 = *reinterpret_cast<const float2*>(sPB + (16 * mt1 + g) * LDT + k8 + 2 * c);
      const float2 a1 = *reinterpret_cast<const float2*>(sPB + (16 * mt1 + g + 8) * LDT + k8 + 2 * c);
      Frag4 ag;
      split_a(ag, a0.x, a0.y, a1.x, a1.y);
#pragma unroll
      for (int j0 = 0; j0 < NW1; j0 += 4) {
        Frag2 bh[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (j0 + j < NW1)
            split_b(bh[j], *reinterpret_cast<const float2*>(sPA + (8 * (nf1 + j0 + j) + g) * LDT + k8 + 2 * c));
#pragma unroll
        for (int term = 0; term < 3; ++term)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j0 + j < NW1) mma_term(accW1[j0 + j], ag, bh[j], term);
      }
      if (own_b1) {   // warp-uniform; B = ones in column 0: hi = 1, lo = 0
        const uint32_t one = g == 0 ? 0x3f800000u : 0u;
        const uint32_t bo[2] = {one, one};
        mma_tf32(accB1, ag.lo, bo);
        mma_tf32(accB1, ag.hi, bo);
      }
    }
    half_sync(half);   // (B) planes free
    // GZ1^T -> PB, H2^T -> PA
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      float* p0 = sPB + (8 * nf + 2 * c) * LDT;
      p0[rA] = gz[nf][0]; p0[LDT + rA] = gz[nf][1]; p0[rB] = gz[nf][2]; p0[LDT + rB] = gz[nf][3];
      float* q0 = sPA + (8 * nf + 2 * c) * LDT;
      q0[rA] = h2[nf][0]; q0[LDT + rA] = h2[nf][1]; q0[rB] = h2[nf][2]; q0[LDT + rB] = h2[nf][3];
    }
    half_sync(half);   // (C)
    // ---- gW0[o][i] += sum_t GZ1[t][o] X[t][i] (column n of X is ones: gb0);  gW2[j][f] += sum_t U[t][j] H2[t][f]
#pragma unroll
    for (int s = 0; s < NW0; ++s) {
      const int f0 = hw + 4 * s;
      if (f0 < MT * NPK) {   // warp-uniform
        const int mt = f0 / NPK, nf = f0 - mt * NPK;
#pragma unroll 4
        for (int k8 = 0; k8 < TSR; k8 += 8) {
          const float2 a0 = *reinterpret_cast<const float2*>(sPB + (16 * mt + g) * LDT + k8 + 2 * c);
          const float2 a1 = *reinterpret_cast<const float2*>(sPB + (16 * mt + g + 8) * LDT + k8 + 2 * c);
          Frag4 ag;
          Frag2 bx;
          split_a(ag, a0.x, a0.y, a1.x, a1.y);
          split_b(bx, *reinterpret_cast<const float2*>(sXT + (8 * nf + g) * LDT + k8 + 2 * c));
#pragma unroll
          for (int term = 0; term < 3; ++term) mma_term(accW0[s], ag, bx, term);
        }
      }
    }
#pragma unroll
    for (int s = 0; s < NW2; ++s) {
      const int nf = hw + 4 * s;
      if (nf < NF) {
#pragma unroll 4
        for (int k8 = 0; k8 < TSR; k8 += 8) {
          const float2 a0 = *reinterpret_cast<const float2*>(sUT + g * LDT + k8 + 2 * c);
          Frag4 au;   // rows 8..15 of U^T are zero: no load
          Frag2 bh;
          split_a(au, a0.x, a0.y, 0.f, 0.f);
          split_b(bh, *reinterpret_cast<const float2*>(sPA + (8 * nf + g) * LDT + k8 + 2 * c));
#pragma unroll
          for (int term = 0; term < 3; ++term) mma_term(accW2[s], au, bh, term);
        }
      }
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) {   // gb2[j] += sum_t U[t][j], j = hw + 4 s (m <= 8)
      const int j = hw + 4 * s;
      if (j < m) {
        const float2 v = *reinterpret_cast<const float2*>(sUT + j * LDT + 2 * lane);
        float sum = v.x + v.y;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        gb2acc[s] += sum;
      }
    }
    half_sync(half);   // (D) X, planes, U^T free for the half's next tile
  }

  // ---- per-CTA partial gradient: within a half every element has exactly one owner; half 0 writes
  // the flat [d] staging vector, half 1 adds to it (fixed order) ------------------------------------
  __syncthreads();
  float* sG = sm + L.PA;
  for (int i = tid; i < F.d; i += NTHREADS) sG[i] = 0.f;
  __syncthreads();
  const int h1n = F.h1, h2n = F.h2;
  for (int hh = 0; hh < 2; ++hh) {
    if (half == hh) {
#pragma unroll
      for (int j = 0; j < NW1; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int o = 16 * mt1 + g + (e >> 1) * 8, i = 8 * (nf1 + j) + 2 * c + (e & 1);
          if (o < h2n && i < h1n) sG[F.oW1 + o * h1n + i] += accW1[j][e];
        }
      if (own_b1 && c == 0) {
        const int o = 16 * mt1 + g;
        if (o < h2n) sG[F.ob1 + o] += accB1[0];
        if (o + 8 < h2n) sG[F.ob1 + o + 8] += accB1[2];
      }
#pragma unroll
      for (int s = 0; s < NW0; ++s) {
        const int f0 = hw + 4 * s;
        if (f0 < MT * NPK) {
          const int mt = f0 / NPK, nf = f0 - mt * NPK;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int o = 16 * mt + g + (e >> 1) * 8, i = 8 * nf + 2 * c + (e & 1);
            if (o < h1n) {
              if (i < n) sG[F.oW0 + o * n + i] += accW0[s][e];
              else if (i == n) sG[F.ob0 + o] += accW0[s][e];
            }
          }
        }
      }
#pragma unroll
      for (int s = 0; s < NW2; ++s) {
        const int nf = hw + 4 * s;
        if (nf < NF && g < m) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int f = 8 * nf + 2 * c + e;
            if (f < h2n) sG[F.oW2 + g * h2n + f] += accW2[s][e];
          }
        }
      }
#pragma unroll
      for (int s = 0; s < 2; ++s)
        if (hw + 4 * s < m && lane == 0) sG[F.ob2 + hw + 4 * s] += gb2acc[s];   // log_std block: finalize kernel
    }
    __syncthreads();
  }
  float* out = a.partials + (size_t)blockIdx.x * a.dstride;
  for (int i = tid; i < F.d; i += NTHREADS) out[i] = sG[i];
}

}  // namespace rows
}  // namespace pol
