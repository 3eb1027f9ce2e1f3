// Fisher-vector product with the per-timestep output Jacobian cached: the third dataflow of the product
// (after the tile kernel and fvp_rows.cuh), for small action dimensions (m <= 4).
//
// Inside one CG solve the policy is fixed and only the vector v changes (cg_solve.py:10-20,
// npg_cg.py:103-106).  fvp_rows.cuh already keeps the activations h1, h2 of every timestep in HBM.
// The Gauss-Newton product is  F v = sum_t J_t^T M J_t v  with J_t = d mu_t / d theta (m rows), and
// J_t factors through the m backpropagated rows
//     G2_j = W2[j,:] * (1 - h2^2)             (d mu_j / d z2, one multiply per element)
//     G1_j = (G2_j W1) * (1 - h1^2)           (d mu_j / d z1, a 64x64 product per j)
// which do NOT depend on v.  With G1 cached next to h1, h2 (m*H more floats per timestep):
//     q1 = dW0 x + db0,  q2 = dW1 h1 + db1,  q3 = dW2 h2 + db2      (the only forward-side products left)
//     (J v)_j = q3_j + <G2_j, q2> + <G1_j, q1>                       (element-wise, quad shuffles)
//     u = M (J v);   gz2 = sum_j u_j G2_j,  gz1 = sum_j u_j G1_j     (element-wise)
//     gW2 += u^T h2,  gW1 += gz2^T h1,  gW0 += gz1^T x               (as before)
// so the tangent propagation (D1 W1^T, D2 W2^T) and the backward data products (U W2, GZ2 W1) -- 43 % of
// the MMAs of fvp_rows_kernel<use cache> -- disappear; the price is m*H*4 more bytes per timestep from
// HBM (768 B at m = 3, H = 64) and one fill pass per solve (jac_fill_kernel: m backward products).
// Same fragment conventions, half-CTAs, planes and partial-sum layout as fvp_rows.cuh.
#include <stdlib.h>

#include "policy_tile.cuh"
#include "fvp_rows.cuh"

namespace pol {
namespace jac {

using rows::load_b; using rows::load_b_rows; using rows::split_a; using rows::split_b; using rows::mma_term;
using rows::ld_cache; using rows::ldx_of; using rows::half_sync;
using rows::TSR; using rows::LDT; using rows::HTHREADS;

constexpr int MJ = 4;   // largest action dimension this variant takes

// G1 is read twice per tile (for J v, then for gz1): let the first read allocate in L1
__device__ __forceinline__ float2 ld_g(const float2* p) {
  float2 v;
  asm("ld.global.nc.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
  return v;
}
// one warp pulls `lines` 128-byte lines starting at p towards L2 (no registers, no wait)
__device__ __forceinline__ void prefetch_l2(const void* p, int lines, int lane) {
  const char* q = reinterpret_cast<const char*>(p);
  for (int i = lane; i < lines; i += 32) asm volatile("prefetch.global.L2 [%0];" :: "l"(q + (size_t)i * 128));
}

struct JLayout {  // offsets in floats
  int NP, LDX, LDW;
  int X, XT, PA, PB, UT, XR, V0, V1, V2, W2, vb0, vb1, vb2, fisher, ishift, iden;
  int half_stride, total;
  __host__ __device__ JLayout(int HP, int NPK, int n, int NH = 2) {
    NP = 8 * NPK; LDX = ldx_of(NP); LDW = HP + 8;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 3) & ~3; return r; };
    PA = take(HP * LDT); PB = take(HP * LDT);     // PA|PB of half 0 double as the [d] staging vector
    X = take(TSR * LDX); XT = take(NP * LDT); UT = take(8 * LDT);
    XR = take(2 * TSR * n);
    half_stride = o;
    o *= NH;
    V0 = take(2 * HP * LDX); V1 = take(2 * HP * LDW); V2 = take(2 * 8 * LDW);   // tf32 hi plane | lo plane
    W2 = take(MJ * HP);
    vb0 = take(HP); vb1 = take(HP); vb2 = take(8); fisher = take(8);
    ishift = take(NP); iden = take(NP);
    total = o;
  }
};

// G1 slots per (tile, warp): [j][block nf][row g / g+8][lane] float2, after the h1|h2 region of the cache
__host__ __device__ inline long long h_region_floats(long long ntiles, int HP) { return ntiles * TSR * 2 * HP; }

// NH independent groups of 4 warps per CTA ("halves" in fvp_rows.cuh; here 2 or 3): with 3 the CTA has 12
// warps of <= 168 registers, i.e. three instruction streams per SM sub-partition instead of two -- this
// variant is bound by instruction issue latency, not by the tensor pipe (ncu: pipe 35 % busy with NH = 2).
// MK = static bound of the action dimension (3 or 4): sizes the register-resident rows.
template <int HP, int NPK, int NH, int MK>
__global__ void __launch_bounds__(NH * HTHREADS, 1) fvp_jac_kernel(const Args a) {
  constexpr int NTHREADS = NH * HTHREADS;   // shadows pol::NTHREADS
  constexpr int NF = HP / 8, MT = HP / 16;
  constexpr int NP = 8 * NPK, LDX = ldx_of(NP), LDW = HP + 8;
  constexpr int NW1 = MT * NF / 4, NW0 = (MT * NPK + 3) / 4, NW2 = (NF + 3) / 4;
  constexpr int G = 2;
  static_assert(NF % NW1 == 0 && NF % G == 0, "fragment ownership");
  extern __shared__ __align__(16) float sm[];
  __shared__ uint64_t bar[NH][2];
  if (a.stop_flag && *a.stop_flag) return;

  const JLayout L(HP, NPK, a.n, NH);
  const Flat F(a.n, a.m, a.h1, a.h2);
  const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int half = tid >> 7, htid = tid & (HTHREADS - 1), hw = htid >> 5;
  const int n = a.n, m = a.m;
  float* hs = sm + (size_t)half * L.half_stride;
  float* sX = hs + L.X; float* sXT = hs + L.XT; float* sPA = hs + L.PA; float* sPB = hs + L.PB;
  float* sUT = hs + L.UT; float* sXR = hs + L.XR;
  const float* sV0 = sm + L.V0; const float* sV1 = sm + L.V1; const float* sV2 = sm + L.V2;
  const float* sW2 = sm + L.W2;
  constexpr int PL0 = HP * LDX, PL1 = HP * LDW, PL2 = 8 * LDW;

  // cache slices of a (tile, warp): h1|h2 (fvp_rows layout) and, after that region, the m rows G1
  const int ntiles_all = (int)((a.N + TSR - 1) / TSR);
  const float2* hbase = reinterpret_cast<const float2*>(a.cache);
  const float2* gbase = reinterpret_cast<const float2*>(a.cache + h_region_floats(ntiles_all, HP));
  constexpr int HSLOT = 4 * NF * 32;                 // float2 per (tile, warp) of the h1|h2 region
  const size_t gslot_stride = (size_t)m * 2 * NF * 32;
  auto prefetch_tile = [&](int tile) {               // the loads of that tile then hit L2
    if (tile < ntiles_all) {
      prefetch_l2(hbase + ((size_t)tile * 4 + hw) * HSLOT, HSLOT * 8 / 128, lane);
      prefetch_l2(gbase + ((size_t)tile * 4 + hw) * gslot_stride, (int)(gslot_stride * 8 / 128), lane);
    }
  };
  prefetch_tile(NH * blockIdx.x + half);

  for (int i = tid; i < L.total; i += NTHREADS) sm[i] = 0.f;
  if (tid == 0) {
    for (int hh = 0; hh < NH; ++hh) { tma::mbar_init(&bar[hh][0], 1); tma::mbar_init(&bar[hh][1], 1); }
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
      put(L.V0 + o * LDX + k, PL0, __ldg(V + F.oW0 + i));
    }
    for (int i = tid; i < F.h2 * F.h1; i += NTHREADS) {
      const int o = i / F.h1, k = i - o * F.h1;
      put(L.V1 + o * LDW + k, PL1, __ldg(V + F.oW1 + i));
    }
    for (int i = tid; i < m * F.h2; i += NTHREADS) {
      const int j = i / F.h2, k = i - j * F.h2;
      put(L.V2 + j * LDW + k, PL2, __ldg(V + F.oW2 + i));
      sm[L.W2 + j * HP + k] = __ldg(P + F.oW2 + i);
    }
    for (int i = tid; i < F.h1; i += NTHREADS) sm[L.vb0 + i] = __ldg(V + F.ob0 + i);
    for (int i = tid; i < F.h2; i += NTHREADS) sm[L.vb1 + i] = __ldg(V + F.ob1 + i);
    for (int j = tid; j < m; j += NTHREADS) {
      sm[L.vb2 + j] = __ldg(V + F.ob2 + j);
      const float sg = expf(__ldg(P + F.ols + j));
      const float os = a.out_scale ? __ldg(a.out_scale + j) : 1.f;
      sm[L.fisher + j] = os * os * (2.f / (2.f * sg * sg + 1e-8f));   // as fvp_rows.cuh
    }
    for (int i = tid; i < NP; i += NTHREADS) {
      sm[L.ishift + i] = (a.in_shift && i < n) ? __ldg(a.in_shift + i) : 0.f;
      // reciprocal of (in_scale + 1e-8): the input transform is a multiply here (x~ within 1 ulp of the
      // division the cache-filling passes use; it only feeds q1 and gW0, far inside the 1e-4 gate)
      sm[L.iden + i] = 1.f / ((a.in_scale && i < n ? __ldg(a.in_scale + i) : 1.f) + 1e-8f);
    }
  }

  float accW1[NW1][4], accB1[4], accW0[NW0][4], accW2[NW2][4], gb2acc[2] = {0.f, 0.f};
#pragma unroll
  for (int j = 0; j < NW1; ++j) accW1[j][0] = accW1[j][1] = accW1[j][2] = accW1[j][3] = 0.f;
#pragma unroll
  for (int j = 0; j < NW0; ++j) accW0[j][0] = accW0[j][1] = accW0[j][2] = accW0[j][3] = 0.f;
#pragma unroll
  for (int j = 0; j < NW2; ++j) accW2[j][0] = accW2[j][1] = accW2[j][2] = accW2[j][3] = 0.f;
  accB1[0] = accB1[1] = accB1[2] = accB1[3] = 0.f;
  const int mt1 = (hw * NW1) / NF, nf1 = (hw * NW1) % NF;
  const bool own_b1 = nf1 == 0;

  const int ntiles = ntiles_all;
  const int tile_elems = TSR * n;
  const bool base_ok = (reinterpret_cast<uintptr_t>(a.obs) & 15) == 0;
  auto tile_rows = [&](int tile) { return tile + 1 < ntiles ? TSR : (int)(a.N - (long long)tile * TSR); };
  auto bulk_ok = [&](int tile) { return base_ok && ((tile_rows(tile) * n) & 3) == 0; };
  auto start_copy = [&](int tile, int buf) {
    if (bulk_ok(tile)) {
      const uint32_t bytes = (uint32_t)(tile_rows(tile) * n) * 4u;
      tma::mbar_expect(&bar[half][buf], bytes);
      tma::bulk_g2s(sXR + (size_t)buf * tile_elems, a.obs + (long long)tile * tile_elems, bytes, &bar[half][buf]);
    }
  };
  __syncthreads();
  const int tile_first = NH * blockIdx.x + half, tile_step = NH * gridDim.x;
  if (htid == 0 && tile_first < ntiles) start_copy(tile_first, 0);

  const int rA = 16 * hw + g, rB = rA + 8;
  int it = 0;
  for (int tile = tile_first; tile < ntiles; tile += tile_step, ++it) {
    const int buf = it & 1;
    const long long t0 = (long long)tile * TSR;
    const int rows = tile_rows(tile);
    if (htid == 0 && tile + tile_step < ntiles) start_copy(tile + tile_step, buf ^ 1);
    prefetch_tile(tile + tile_step);
    const float2* cslot = hbase + ((size_t)tile * 4 + hw) * HSLOT + lane;
    const float2* gslot = gbase + ((size_t)tile * 4 + hw) * gslot_stride + lane;
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
        const float v = (raw[i] - sm[L.ishift + k]) * sm[L.iden + k];
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

    // ---- q1 = vb0 + X V0^T ; q2 = vb1 + H1 V1^T -------------------------------------------------
    float q1[NF][4], q2[NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 v0 = *reinterpret_cast<const float2*>(sm + L.vb0 + 8 * nf + 2 * c);
      const float2 v1 = *reinterpret_cast<const float2*>(sm + L.vb1 + 8 * nf + 2 * c);
      q1[nf][0] = v0.x; q1[nf][1] = v0.y; q1[nf][2] = v0.x; q1[nf][3] = v0.y;
      q2[nf][0] = v1.x; q2[nf][1] = v1.y; q2[nf][2] = v1.x; q2[nf][3] = v1.y;
    }
#pragma unroll
    for (int ks = 0; ks < NPK; ++ks) {
      const float2 xa = *reinterpret_cast<const float2*>(sX + rA * LDX + 8 * ks + 2 * c);
      const float2 xb = *reinterpret_cast<const float2*>(sX + rB * LDX + 8 * ks + 2 * c);
      Frag4 ax;
      split_a(ax, xa.x, xa.y, xb.x, xb.y);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bv[G];
#pragma unroll
        for (int i = 0; i < G; ++i) load_b(bv[i], sV0 + (8 * (n0 + i) + g) * LDX + 8 * ks + 2 * c, PL0);
#pragma unroll
        for (int term = 0; term < 3; ++term)
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(q1[n0 + i], ax, bv[i], term);
      }
    }
    // h1 one 8-column block at a time (never all 32 values in registers): H1^T plane + A fragment
#pragma unroll
    for (int ks = 0; ks < NF; ++ks) {
      const float2 lo = ld_cache(cslot + (2 * ks) * 32), hi = ld_cache(cslot + (2 * ks + 1) * 32);
      {
        float* p0 = sPA + (8 * ks + 2 * c) * LDT;
        p0[rA] = lo.x; p0[LDT + rA] = lo.y; p0[rB] = hi.x; p0[LDT + rB] = hi.y;
      }
      Frag4 ah;
      split_a(ah, lo.x, lo.y, hi.x, hi.y);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bv[G];
#pragma unroll
        for (int i = 0; i < G; ++i) load_b(bv[i], sV1 + (8 * (n0 + i) + g) * LDW + 8 * ks + 2 * c, PL1);
#pragma unroll
        for (int term = 0; term < 3; ++term)
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(q2[n0 + i], ah, bv[i], term);
      }
    }
    // Jacobian rows of output 0 are requested before the head product: their latency hides behind its MMAs
    float gk[MK][NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 lo = ld_cache(gslot + (size_t)(2 * nf) * 32), hi = ld_cache(gslot + (size_t)(2 * nf + 1) * 32);
      gk[0][nf][0] = lo.x; gk[0][nf][1] = lo.y; gk[0][nf][2] = hi.x; gk[0][nf][3] = hi.y;
    }
    // ---- h2 block by block: q3 = vb2 + H2 V2^T (rows g / g+8, columns j = 2c, 2c+1), and
    //      q2 <- (1 - h2^2) q2 so that <G2_j, q2> = sum_col W2[j][col] q2[col] --------------------------
    float q3[4];
    {
      float uu[2][4];
      uu[0][0] = uu[0][1] = uu[0][2] = uu[0][3] = 0.f;
      uu[1][0] = uu[1][1] = uu[1][2] = uu[1][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < NF; ++ks) {
        const float2 lo = ld_g(cslot + (2 * (NF + ks)) * 32), hi = ld_g(cslot + (2 * (NF + ks) + 1) * 32);   // re-read twice below
        Frag4 ah;
        Frag2 bv;
        split_a(ah, lo.x, lo.y, hi.x, hi.y);
        load_b(bv, sV2 + g * LDW + 8 * ks + 2 * c, PL2);
#pragma unroll
        for (int term = 0; term < 3; ++term) mma_term(uu[ks & 1], ah, bv, term);
        q2[ks][0] *= 1.f - lo.x * lo.x; q2[ks][1] *= 1.f - lo.y * lo.y;
        q2[ks][2] *= 1.f - hi.x * hi.x; q2[ks][3] *= 1.f - hi.y * hi.y;
      }
      const float2 vj = *reinterpret_cast<const float2*>(sm + L.vb2 + 2 * c);
      q3[0] = uu[0][0] + uu[1][0] + vj.x; q3[1] = uu[0][1] + uu[1][1] + vj.y;
      q3[2] = uu[0][2] + uu[1][2] + vj.x; q3[3] = uu[0][3] + uu[1][3] + vj.y;
    }
    // ---- (J v)_j.  Layer-2 part first (it frees q2), then the layer-1 part, whose rows G1_j stay in
    //      registers for gz1 below: every cached row is read from HBM exactly once per product -------
    float pa[MK], pb[MK];
#pragma unroll
    for (int j = 0; j < MK; ++j) {
      pa[j] = 0.f; pb[j] = 0.f;
      if (j < m) {
#pragma unroll
        for (int nf = 0; nf < NF; ++nf) {
          const float2 w = *reinterpret_cast<const float2*>(sW2 + j * HP + 8 * nf + 2 * c);
          pa[j] = fmaf(w.x, q2[nf][0], pa[j]); pa[j] = fmaf(w.y, q2[nf][1], pa[j]);
          pb[j] = fmaf(w.x, q2[nf][2], pb[j]); pb[j] = fmaf(w.y, q2[nf][3], pb[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < MK; ++j) {
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        if (j > 0) {
          float2 lo = make_float2(0.f, 0.f), hi = make_float2(0.f, 0.f);
          if (j < m) {
            lo = ld_cache(gslot + (size_t)(j * 2 * NF + 2 * nf) * 32);
            hi = ld_cache(gslot + (size_t)(j * 2 * NF + 2 * nf + 1) * 32);
          }
          gk[j][nf][0] = lo.x; gk[j][nf][1] = lo.y; gk[j][nf][2] = hi.x; gk[j][nf][3] = hi.y;
        }
        pa[j] = fmaf(gk[j][nf][0], q1[nf][0], pa[j]); pa[j] = fmaf(gk[j][nf][1], q1[nf][1], pa[j]);
        pb[j] = fmaf(gk[j][nf][2], q1[nf][2], pb[j]); pb[j] = fmaf(gk[j][nf][3], q1[nf][3], pb[j]);
      }
    }
    // u_j = fisher_j (J v)_j for both rows of this thread (zero past N)
    float uA[MK], uB[MK];
    const bool va = rA < rows, vb = rB < rows;
#pragma unroll
    for (int j = 0; j < MK; ++j) {
      uA[j] = 0.f; uB[j] = 0.f;
      if (j < m) {
        float sa = pa[j], sb = pb[j];
        sa += __shfl_xor_sync(0xffffffffu, sa, 1); sb += __shfl_xor_sync(0xffffffffu, sb, 1);
        sa += __shfl_xor_sync(0xffffffffu, sa, 2); sb += __shfl_xor_sync(0xffffffffu, sb, 2);
        const int src = (lane & ~3) | (j >> 1);
        const float ya = __shfl_sync(0xffffffffu, q3[j & 1], src);
        const float yb = __shfl_sync(0xffffffffu, q3[2 + (j & 1)], src);
        const float fj = sm[L.fisher + j];
        uA[j] = va ? fj * (sa + ya) : 0.f;
        uB[j] = vb ? fj * (sb + yb) : 0.f;
        if (c == 0) { sUT[j * LDT + rA] = uA[j]; sUT[j * LDT + rB] = uB[j]; }
      }
    }
    // ---- gz2 = (sum_j u_j W2[j,:]) (1 - h2^2) -> GZ2^T plane;  gz1 = sum_j u_j G1_j (kept) ----------
    float gz[NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 h2lo = ld_g(cslot + (2 * (NF + nf)) * 32), h2hi = ld_g(cslot + (2 * (NF + nf) + 1) * 32);   // L1 hit
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
      gz[nf][0] = gz[nf][1] = gz[nf][2] = gz[nf][3] = 0.f;
#pragma unroll
      for (int j = 0; j < MK; ++j)
        if (j < m) {
          const float2 w = *reinterpret_cast<const float2*>(sW2 + j * HP + 8 * nf + 2 * c);
          s0 = fmaf(uA[j], w.x, s0); s1 = fmaf(uA[j], w.y, s1);
          s2 = fmaf(uB[j], w.x, s2); s3 = fmaf(uB[j], w.y, s3);
          gz[nf][0] = fmaf(uA[j], gk[j][nf][0], gz[nf][0]); gz[nf][1] = fmaf(uA[j], gk[j][nf][1], gz[nf][1]);
          gz[nf][2] = fmaf(uB[j], gk[j][nf][2], gz[nf][2]); gz[nf][3] = fmaf(uB[j], gk[j][nf][3], gz[nf][3]);
        }
      float* p0 = sPB + (8 * nf + 2 * c) * LDT;
      p0[rA] = s0 * (1.f - h2lo.x * h2lo.x); p0[LDT + rA] = s1 * (1.f - h2lo.y * h2lo.y);
      p0[rB] = s2 * (1.f - h2hi.x * h2hi.x); p0[LDT + rB] = s3 * (1.f - h2hi.y * h2hi.y);
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
      if (own_b1) {
        const uint32_t one = g == 0 ? 0x3f800000u : 0u;
        const uint32_t bo[2] = {one, one};
        mma_tf32(accB1, ag.lo, bo);
        mma_tf32(accB1, ag.hi, bo);
      }
    }
    half_sync(half);   // (B) planes free
    // GZ1^T -> PB, H2^T -> PA (h2 read back from the cache: an L2 hit, frees 32 registers across gW1)
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      float* p0 = sPB + (8 * nf + 2 * c) * LDT;
      p0[rA] = gz[nf][0]; p0[LDT + rA] = gz[nf][1]; p0[rB] = gz[nf][2]; p0[LDT + rB] = gz[nf][3];
      const float2 lo = ld_g(cslot + (2 * (NF + nf)) * 32), hi = ld_g(cslot + (2 * (NF + nf) + 1) * 32);
      float* q0 = sPA + (8 * nf + 2 * c) * LDT;
      q0[rA] = lo.x; q0[LDT + rA] = lo.y; q0[rB] = hi.x; q0[LDT + rB] = hi.y;
    }
    half_sync(half);   // (C)
    // ---- gW0 += GZ1^T X (column n of X is ones: gb0);  gW2 += U^T H2 ------------------------------
#pragma unroll
    for (int s = 0; s < NW0; ++s) {
      const int f0 = hw + 4 * s;
      if (f0 < MT * NPK) {
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
          Frag4 au;
          Frag2 bh;
          split_a(au, a0.x, a0.y, 0.f, 0.f);
          split_b(bh, *reinterpret_cast<const float2*>(sPA + (8 * nf + g) * LDT + k8 + 2 * c));
#pragma unroll
          for (int term = 0; term < 3; ++term) mma_term(accW2[s], au, bh, term);
        }
      }
    }
    {   // gb2[j] += sum_t U[t][j], j = hw (m <= 4)
      if (hw < m) {
        const float2 v = *reinterpret_cast<const float2*>(sUT + hw * LDT + 2 * lane);
        float sum = v.x + v.y;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        gb2acc[0] += sum;
      }
    }
    half_sync(half);   // (D)
  }

  // ---- per-CTA partial gradient (as fvp_rows.cuh) ------------------------------------------------
  __syncthreads();
  float* sG = sm + L.PA;
  for (int i = tid; i < F.d; i += NTHREADS) sG[i] = 0.f;
  __syncthreads();
  const int h1n = F.h1, h2n = F.h2;
  for (int hh = 0; hh < NH; ++hh) {
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
      if (hw < m && lane == 0) sG[F.ob2 + hw] += gb2acc[0];
    }
    __syncthreads();
  }
  float* out = a.partials + (size_t)blockIdx.x * a.dstride;
  for (int i = tid; i < F.d; i += NTHREADS) out[i] = sG[i];
}

// G1_j = ((W2[j,:] * (1 - h2^2)) W1) * (1 - h1^2) for every timestep, from the cached h1, h2: the
// backward data product of fvp_rows.cuh run once per output j.  Warps are independent (16 timesteps each).
template <int HP>
__global__ void __launch_bounds__(256, 2) jac_fill_kernel(const float* __restrict__ params, int n, int m, int h1n,
                                                        int h2n, float* __restrict__ cache, long long ntiles) {
  constexpr int NF = HP / 8, LDW = HP + 8, PL1 = HP * LDW, G = 2;
  extern __shared__ __align__(16) float sm[];
  float* sW1 = sm;                    // W1 TRANSPOSED ([in][out]), tf32 hi plane | lo plane
  float* sW2 = sm + 2 * HP * LDW;     // [MJ][HP]
  const Flat F(n, m, h1n, h2n);
  const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, c = lane & 3, warp = tid >> 5;
  for (int i = tid; i < 2 * HP * LDW + MJ * HP; i += 256) sm[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < h2n * h1n; i += 256) {
    const int o = i / h1n, k = i - o * h1n;
    uint32_t hi, lo;
    split_tf32(__ldg(params + F.oW1 + i), hi, lo);
    sW1[k * LDW + o] = __uint_as_float(hi); sW1[PL1 + k * LDW + o] = __uint_as_float(lo);   // W1^T [in][out]
  }
  for (int i = tid; i < m * h2n; i += 256) {
    const int j = i / h2n, k = i - j * h2n;
    sW2[j * HP + k] = __ldg(params + F.oW2 + i);
  }
  __syncthreads();
  float2* gbase = reinterpret_cast<float2*>(cache + h_region_floats(ntiles, HP));
  const long long nwt = ntiles * 4;
  for (long long wt = (long long)blockIdx.x * 8 + warp; wt < nwt; wt += (long long)gridDim.x * 8) {
    const float2* cslot = reinterpret_cast<const float2*>(cache) + (size_t)wt * (4 * NF * 32) + lane;
    float2* gslot = gbase + (size_t)wt * ((size_t)m * 2 * NF * 32) + lane;
    float f2[NF][4];   // 1 - h2^2; h1 is read again at store time (an L1/L2 hit) to stay under 128 registers
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 b0 = ld_cache(cslot + (2 * (NF + nf)) * 32), b1 = ld_cache(cslot + (2 * (NF + nf) + 1) * 32);
      f2[nf][0] = 1.f - b0.x * b0.x; f2[nf][1] = 1.f - b0.y * b0.y; f2[nf][2] = 1.f - b1.x * b1.x; f2[nf][3] = 1.f - b1.y * b1.y;
    }
    for (int j = 0; j < m; ++j) {
      // n-blocks outermost: 8 accumulator registers live instead of 4*NF (the A fragments are re-split per
      // pair of blocks: a few ALU ops against 1 KB of spills)
#pragma unroll 1
      for (int n0 = 0; n0 < NF; n0 += G) {
        float gh[G][4];
#pragma unroll
        for (int i = 0; i < G; ++i) gh[i][0] = gh[i][1] = gh[i][2] = gh[i][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < NF; ++ks) {
          const float2 w = *reinterpret_cast<const float2*>(sW2 + j * HP + 8 * ks + 2 * c);
          Frag4 ag;
          split_a(ag, w.x * f2[ks][0], w.y * f2[ks][1], w.x * f2[ks][2], w.y * f2[ks][3]);
          Frag2 bw[G];
#pragma unroll
          for (int i = 0; i < G; ++i)
            load_b(bw[i], sW1 + (8 * (n0 + i) + g) * LDW + 8 * ks + 2 * c, PL1);   // B[k = o][n = i] = W1[o][i] = W1^T[i][o]: one 8-byte load
#pragma unroll
          for (int term = 0; term < 3; ++term)
#pragma unroll
            for (int i = 0; i < G; ++i) mma_term(gh[i], ag, bw[i], term);
        }
#pragma unroll
        for (int i = 0; i < G; ++i) {
          const int nf = n0 + i;
          const float2 a0 = ld_g(cslot + (2 * nf) * 32), a1 = ld_g(cslot + (2 * nf + 1) * 32);
          gslot[(size_t)(j * 2 * NF + 2 * nf) * 32] =
              make_float2(gh[i][0] * (1.f - a0.x * a0.x), gh[i][1] * (1.f - a0.y * a0.y));
          gslot[(size_t)(j * 2 * NF + 2 * nf + 1) * 32] =
              make_float2(gh[i][2] * (1.f - a1.x * a1.x), gh[i][3] * (1.f - a1.y * a1.y));
        }
      }
    }
  }
}

}  // namespace jac

// ---- host side (declared in policy_tile.cuh) -----------------------------------------------------
bool jac_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2, int* HP_out, int* NPK_out, size_t* smem_out) {
  const int hmax = h1 > h2 ? h1 : h2;
  if (h1 <= 0 || h2 <= 0 || hmax > 64 || m > jac::MJ || m < 1 || n + 1 > 24) return false;
  static const bool off = getenv("MJRL_B200_FVP_JAC") && getenv("MJRL_B200_FVP_JAC")[0] == '0';
  if (off) return false;
  const int HP = hmax <= 32 ? 32 : 64, NPK = (n + 1 + 7) / 8;
  const Flat F(n, m, h1, h2);
  const jac::JLayout L(HP, NPK, n);
  const size_t smem = sizeof(float) * (size_t)L.total;
  if (smem + 64 > (size_t)ctx->max_smem_optin || F.d > 2 * HP * rows::LDT) return false;
  if (HP_out) *HP_out = HP;
  if (NPK_out) *NPK_out = NPK;
  if (smem_out) *smem_out = smem;
  return true;
}

long long jac_cache_floats(int m, int h1, int h2, long long N) {
  const int HP = (h1 > h2 ? h1 : h2) <= 32 ? 32 : 64;
  const long long ntiles = (N + rows::TSR - 1) / rows::TSR;
  return ntiles * rows::TSR * (long long)m * HP;
}

template <int HP, int NPK, int NH, int MK>
static int launch_jac_one(const Args& a, int grid, size_t smem, cudaStream_t s) {
  auto kern = jac::fvp_jac_kernel<HP, NPK, NH, MK>;
  static size_t attr = 0;
  if (smem > attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr = smem;
  }
  kern<<<grid, NH * rows::HTHREADS, smem, s>>>(a);
  return mjrl_check_launch("fvp_jac_kernel");
}

// a.cache = [h1|h2 region][G1 region], both filled; returns -1 when the shape does not fit
int launch_fvp_jac(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s) {
  int HP, NPK;
  size_t smem;
  if (!a.cache || !jac_fits(ctx, a.n, a.m, a.h1, a.h2, &HP, &NPK, &smem)) return -1;
  const Flat F(a.n, a.m, a.h1, a.h2);
  // two groups of 4 warps per CTA.  (Three groups -- 12 warps of 168 registers, 222 KB of shared memory --
  // measured slower: 0.85 ms against 0.63 ms per product on the Hopper shape)
  constexpr int NH = 2;
  const long long ntiles = (a.N + NH * rows::TSR - 1) / (NH * rows::TSR);
  int grid = (int)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  a.dstride = (F.d + 31) & ~31;
  a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
  if (!a.partials) return MJRL_E_STATE;
#define JAC_CASE(H, K)                                                           \
  if (HP == H && NPK == K)                                                       \
    return a.m <= 3 ? launch_jac_one<H, K, NH, 3>(a, grid, smem, s) : launch_jac_one<H, K, NH, 4>(a, grid, smem, s)
  JAC_CASE(64, 1); JAC_CASE(64, 2); JAC_CASE(64, 3);
  JAC_CASE(32, 1); JAC_CASE(32, 2); JAC_CASE(32, 3);
#undef JAC_CASE
  return -1;
}

int launch_jac_fill(mjrl_ctx* ctx, const float* params, int n, int m, int h1, int h2, float* cache, long long N,
                    cudaStream_t s) {
  int HP;
  if (!jac_fits(ctx, n, m, h1, h2, &HP, nullptr, nullptr)) return -1;
  const long long ntiles = (N + rows::TSR - 1) / rows::TSR;
  const long long nblk = (ntiles * 4 + 7) / 8;
  const int cap = 4 * ctx->sm_count;
  const int grid = (int)(nblk < cap ? nblk : cap);
  const size_t smem = sizeof(float) * (size_t)(2 * HP * (HP + 8) + jac::MJ * HP);
  if (HP == 64) {
    static bool set = false;
    if (!set) {
      cudaFuncSetAttribute(jac::jac_fill_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      set = true;
    }
    jac::jac_fill_kernel<64><<<grid, 256, smem, s>>>(params, n, m, h1, h2, cache, ntiles);
  } else {
    jac::jac_fill_kernel<32><<<grid, 256, smem, s>>>(params, n, m, h1, h2, cache, ntiles);
  }
  return mjrl_check_launch("jac_fill_kernel");
}

}  // namespace pol
