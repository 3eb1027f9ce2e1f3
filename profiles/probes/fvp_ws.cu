// NEGATIVE RESULT, kept for the record (not built into libmjrl_b200.so since round 2): warp-specialised variant of the
// activation-cached rows product, measured 0.794 ms against 0.775 ms for the half-CTA kernel.
// Cached Fisher-vector product, warp-specialised variant of fvp_rows.cuh (same math, same fragment
// conventions, same activation-cache layout; hidden 64, obs_dim <= 14).
//
// fvp_rows_kernel runs 8 warps per SM (2 per sub-partition): every warp walks the register-resident
// forward-tangent / backward chain AND its share of the weight-gradient MMAs, 254 registers each, so no
// third warp fits and the tensor pipe idles whenever both warps of a sub-partition are in an epilogue,
// a split or at a barrier (ncu: pipe 60 % busy, issue 33 %).  Here the CTA has 12 warps:
//
//   warps 4..11  "chain": rows 16(w-4).. of a 128-timestep tile through tangent layers, head, backward
//                data products -- no weight-gradient accumulators, so the chain fits the 168 registers
//                a 384-thread CTA gets;
//   warps 0..3   "wgrad": one per sub-partition, own ALL weight-gradient fragments (gW1 band w, gW0,
//                gW2, bias columns) and contract over the 128 timesteps of the transposed planes the
//                chain warps publish.
//
// Chain and wgrad hand the planes back and forth through named barriers (bar.arrive by the side that
// does not wait, bar.sync by the side that does): FULL1 (H1^T, GZ2^T, U^T stored) -> gW1 -> FREE1 ->
// (GZ1^T, H2^T stored) FULL2 -> gW0, gW2, gb2 -> FREE2.  Three independent MMA streams per
// sub-partition instead of two.
#include "policy_tile.cuh"
#include "fvp_rows.cuh"

namespace pol {
namespace ws {


using rows::load_b; using rows::load_b_rows; using rows::split_a; using rows::split_b; using rows::mma_term;
using rows::ld_cache; using rows::ldx_of;

constexpr int WTHREADS = 384, CHAIN_THREADS = 256, WG_THREADS = 128;
constexpr int TS = 128;          // timesteps per tile (8 chain warps x 16 rows)
constexpr int LDT = TS + 8;      // transposed-plane row stride (= 8 mod 32)
constexpr int BAR_CHAIN = 1, BAR_FULL1 = 2, BAR_FREE1 = 3, BAR_FULL2 = 4, BAR_FREE2 = 5;

struct WLayout {  // offsets in floats
  int NP, LDX, LDW;
  int PA, PB, X, XT, UT, XR, W0, V0, W1, V1, W2, V2, W2T, b0, vb0, b1, vb1, vb2, fisher, ishift, iden, total;
  __host__ __device__ WLayout(int HP, int NPK, int n) {
    NP = 8 * NPK; LDX = ldx_of(NP); LDW = HP + 8;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 3) & ~3; return r; };
    PA = take(HP * LDT); PB = take(HP * LDT);   // contiguous: reused as the [d] staging vector
    X = take(TS * LDX); XT = take(2 * NP * LDT); UT = take(8 * LDT);
    XR = take(TS * n);
    W0 = take(2 * HP * LDX); V0 = take(2 * HP * LDX);
    W1 = take(2 * HP * LDW); V1 = take(2 * HP * LDW);
    W2 = take(2 * 8 * LDW); V2 = take(2 * 8 * LDW); W2T = take(2 * HP * 8);
    b0 = take(HP); vb0 = take(HP); b1 = take(HP); vb1 = take(HP); vb2 = take(8); fisher = take(8);
    ishift = take(NP); iden = take(NP);
    total = o;
  }
};

__device__ __forceinline__ void bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(count) : "memory");
}

template <int HP, int NPK>
__global__ void __launch_bounds__(WTHREADS, 1) fvp_ws_kernel(const Args a) {
  constexpr int NF = HP / 8, MT = HP / 16;
  constexpr int NP = 8 * NPK, LDX = ldx_of(NP), LDW = HP + 8;
  constexpr int NW1 = MT * NF / 4;              // gW1 fragments per wgrad warp: one band
  constexpr int NW0 = (MT * NPK + 3) / 4, NW2 = (NF + 3) / 4;
  constexpr int G = 2;
  static_assert(NW1 == NF, "one gW1 band per wgrad warp (HP = 64)");
  extern __shared__ __align__(16) float sm[];
  __shared__ uint64_t bar;
  if (a.stop_flag && *a.stop_flag) return;

  const WLayout L(HP, NPK, a.n);
  const Flat F(a.n, a.m, a.h1, a.h2);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, c = lane & 3;
  const int n = a.n, m = a.m;
  float* sX = sm + L.X; float* sXT = sm + L.XT; float* sPA = sm + L.PA; float* sPB = sm + L.PB;
  float* sUT = sm + L.UT; float* sXR = sm + L.XR;
  const float* sV0 = sm + L.V0; const float* sW1 = sm + L.W1; const float* sV1 = sm + L.V1;
  const float* sW2 = sm + L.W2; const float* sV2 = sm + L.V2; const float* sW2T = sm + L.W2T;
  constexpr int PL0 = HP * LDX, PL1 = HP * LDW, PL2 = 8 * LDW, PL2T = HP * 8;

  // ---- one-time staging (all 12 warps) ----------------------------------------------------------
  for (int i = tid; i < L.total; i += WTHREADS) sm[i] = 0.f;
  if (tid == 0) { tma::mbar_init(&bar, 1); tma::mbar_init_fence(); }
  __syncthreads();
  {
    const float* P = a.params; const float* V = a.vec;
    auto put = [&](int at, int plane, float w) {
      uint32_t hi, lo;
      split_tf32(w, hi, lo);
      sm[at] = __uint_as_float(hi); sm[at + plane] = __uint_as_float(lo);
    };
    for (int i = tid; i < F.h1 * n; i += WTHREADS) {
      const int o = i / n, k = i - o * n;
      put(L.V0 + o * LDX + k, PL0, __ldg(V + F.oW0 + i));
    }
    for (int i = tid; i < F.h2 * F.h1; i += WTHREADS) {
      const int o = i / F.h1, k = i - o * F.h1;
      put(L.W1 + o * LDW + k, PL1, __ldg(P + F.oW1 + i));
      put(L.V1 + o * LDW + k, PL1, __ldg(V + F.oW1 + i));
    }
    for (int i = tid; i < m * F.h2; i += WTHREADS) {
      const int j = i / F.h2, k = i - j * F.h2;
      const float w = __ldg(P + F.oW2 + i);
      put(L.W2 + j * LDW + k, PL2, w);
      put(L.W2T + k * 8 + j, PL2T, w);
      put(L.V2 + j * LDW + k, PL2, __ldg(V + F.oW2 + i));
    }
    for (int i = tid; i < F.h1; i += WTHREADS) sm[L.vb0 + i] = __ldg(V + F.ob0 + i);
    for (int i = tid; i < F.h2; i += WTHREADS) sm[L.vb1 + i] = __ldg(V + F.ob1 + i);
    for (int j = tid; j < m; j += WTHREADS) {
      sm[L.vb2 + j] = __ldg(V + F.ob2 + j);
      const float sg = expf(__ldg(P + F.ols + j));
      const float os = a.out_scale ? __ldg(a.out_scale + j) : 1.f;
      sm[L.fisher + j] = os * os * (2.f / (2.f * sg * sg + 1e-8f));
    }
    for (int i = tid; i < NP; i += WTHREADS) {
      sm[L.ishift + i] = (a.in_shift && i < n) ? __ldg(a.in_shift + i) : 0.f;
      sm[L.iden + i] = (a.in_scale && i < n ? __ldg(a.in_scale + i) : 1.f) + 1e-8f;
    }
  }
  __syncthreads();

  const int ntiles = (int)((a.N + TS - 1) / TS);
  const long long nslabs = ((a.N + rows::TSR - 1) / rows::TSR) * (rows::TSR / 16);   // 16-row slabs the cache holds
  const int h1n = F.h1, h2n = F.h2;
  float* sG = sPA;   // [d] staging at the very end

  if (warp < 4) {
    // =========================== wgrad warps ===================================================
    const int ww = warp;
    float accW1[NW1][4], accB1[4], accW0[NW0][4], accW2[NW2][4], gb2acc[2] = {0.f, 0.f};
#pragma unroll
    for (int j = 0; j < NW1; ++j) accW1[j][0] = accW1[j][1] = accW1[j][2] = accW1[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < NW0; ++j) accW0[j][0] = accW0[j][1] = accW0[j][2] = accW0[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < NW2; ++j) accW2[j][0] = accW2[j][1] = accW2[j][2] = accW2[j][3] = 0.f;
    accB1[0] = accB1[1] = accB1[2] = accB1[3] = 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const float* xt = sXT + (it & 1) * NP * LDT;
      bar_sync(BAR_FULL1, WTHREADS);
      // ---- gW1[o][i] += sum_t GZ2[t][o] H1[t][i], band o = 16 ww .. ; gb1 through a ones column ----
#pragma unroll 2
      for (int k8 = 0; k8 < TS; k8 += 8) {
        const float2 a0 = *reinterpret_cast<const float2*>(sPB + (16 * ww + g) * LDT + k8 + 2 * c);
        const float2 a1 = *reinterpret_cast<const float2*>(sPB + (16 * ww + g + 8) * LDT + k8 + 2 * c);
        Frag4 ag;
        split_a(ag, a0.x, a0.y, a1.x, a1.y);
#pragma unroll
        for (int j0 = 0; j0 < NW1; j0 += 4) {
          Frag2 bh[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            split_b(bh[j], *reinterpret_cast<const float2*>(sPA + (8 * (j0 + j) + g) * LDT + k8 + 2 * c));
#pragma unroll
          for (int term = 0; term < 3; ++term)
#pragma unroll
            for (int j = 0; j < 4; ++j) mma_term(accW1[j0 + j], ag, bh[j], term);
        }
        const uint32_t one = g == 0 ? 0x3f800000u : 0u;
        const uint32_t bo[2] = {one, one};
        mma_tf32(accB1, ag.lo, bo);
        mma_tf32(accB1, ag.hi, bo);
      }
      bar_arrive(BAR_FREE1, WTHREADS);
      bar_sync(BAR_FULL2, WTHREADS);
      // ---- gW0[o][i] += sum_t GZ1[t][o] X[t][i] (column n of X is ones: gb0) ----------------------
#pragma unroll
      for (int s = 0; s < NW0; ++s) {
        const int f0 = ww + 4 * s;
        if (f0 < MT * NPK) {
          const int mt = f0 / NPK, nf = f0 - mt * NPK;
#pragma unroll 4
          for (int k8 = 0; k8 < TS; k8 += 8) {
            const float2 a0 = *reinterpret_cast<const float2*>(sPB + (16 * mt + g) * LDT + k8 + 2 * c);
            const float2 a1 = *reinterpret_cast<const float2*>(sPB + (16 * mt + g + 8) * LDT + k8 + 2 * c);
            Frag4 ag;
            Frag2 bx;
            split_a(ag, a0.x, a0.y, a1.x, a1.y);
            split_b(bx, *reinterpret_cast<const float2*>(xt + (8 * nf + g) * LDT + k8 + 2 * c));
#pragma unroll
            for (int term = 0; term < 3; ++term) mma_term(accW0[s], ag, bx, term);
          }
        }
      }
      // ---- gW2[j][f] += sum_t U[t][j] H2[t][f] ----------------------------------------------------
#pragma unroll
      for (int s = 0; s < NW2; ++s) {
        const int nf = ww + 4 * s;
        if (nf < NF) {
#pragma unroll 4
          for (int k8 = 0; k8 < TS; k8 += 8) {
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
#pragma unroll
      for (int s = 0; s < 2; ++s) {   // gb2[j] += sum_t U[t][j], j = ww + 4 s
        const int j = ww + 4 * s;
        if (j < m) {
          const float4 v = *reinterpret_cast<const float4*>(sUT + j * LDT + 4 * lane);
          float sum = (v.x + v.y) + (v.z + v.w);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
          gb2acc[s] += sum;
        }
      }
      bar_arrive(BAR_FREE2, WTHREADS);
    }
    // ---- all tiles done: publish the CTA's partial gradient (unique owner per element) ------------
    __syncthreads();   // (E1) chain warps are done with the planes
    for (int i = tid; i < F.d; i += WG_THREADS) sG[i] = 0.f;
    bar_sync(6, WG_THREADS);
#pragma unroll
    for (int j = 0; j < NW1; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int o = 16 * ww + g + (e >> 1) * 8, i = 8 * j + 2 * c + (e & 1);
        if (o < h2n && i < h1n) sG[F.oW1 + o * h1n + i] = accW1[j][e];
      }
    if (c == 0) {
      const int o = 16 * ww + g;
      if (o < h2n) sG[F.ob1 + o] = accB1[0];
      if (o + 8 < h2n) sG[F.ob1 + o + 8] = accB1[2];
    }
#pragma unroll
    for (int s = 0; s < NW0; ++s) {
      const int f0 = ww + 4 * s;
      if (f0 < MT * NPK) {
        const int mt = f0 / NPK, nf = f0 - mt * NPK;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int o = 16 * mt + g + (e >> 1) * 8, i = 8 * nf + 2 * c + (e & 1);
          if (o < h1n) {
            if (i < n) sG[F.oW0 + o * n + i] = accW0[s][e];
            else if (i == n) sG[F.ob0 + o] = accW0[s][e];
          }
        }
      }
    }
#pragma unroll
    for (int s = 0; s < NW2; ++s) {
      const int nf = ww + 4 * s;
      if (nf < NF && g < m) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int f = 8 * nf + 2 * c + e;
          if (f < h2n) sG[F.oW2 + g * h2n + f] = accW2[s][e];
        }
      }
    }
#pragma unroll
    for (int s = 0; s < 2; ++s)
      if (ww + 4 * s < m && lane == 0) sG[F.ob2 + ww + 4 * s] = gb2acc[s];   // log_std block: finalize kernel
    __syncthreads();   // (E2)
  } else {
    // =========================== chain warps ===================================================
    const int cw = warp - 4, ctid = tid - WG_THREADS;   // 0..7, 0..255
    const int rA = 16 * cw + g, rB = rA + 8;
    const int tile_elems = TS * n;
    const bool base_ok = (reinterpret_cast<uintptr_t>(a.obs) & 15) == 0;
    auto tile_rows = [&](int tile) { return tile + 1 < ntiles ? TS : (int)(a.N - (long long)tile * TS); };
    auto bulk_ok = [&](int tile) { return base_ok && ((tile_rows(tile) * n) & 3) == 0; };
    auto start_copy = [&](int tile) {   // one thread
      if (bulk_ok(tile)) {
        const uint32_t bytes = (uint32_t)(tile_rows(tile) * n) * 4u;
        tma::mbar_expect(&bar, bytes);
        tma::bulk_g2s(sXR, a.obs + (long long)tile * tile_elems, bytes, &bar);
      }
    };
    if (ctid == 0 && (int)blockIdx.x < ntiles) start_copy(blockIdx.x);
    int it = 0, nbulk = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const long long t0 = (long long)tile * TS;
      const int rows = tile_rows(tile);
      float* xt = sXT + (it & 1) * NP * LDT;
      // ---- this warp's cached activations: issued first, consumed by the layers ------------------
      const long long slab = (long long)tile * (TS / 16) + cw;
      const bool slab_ok = slab < nslabs;
      const float2* cslot = reinterpret_cast<const float2*>(a.cache) + (size_t)(slab_ok ? slab : 0) * (4 * NF * 32) + lane;
      float h1[NF][4], d1[NF][4];
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        float2 lo = make_float2(0.f, 0.f), hi = lo;
        if (slab_ok) { lo = ld_cache(cslot + (2 * nf) * 32); hi = ld_cache(cslot + (2 * nf + 1) * 32); }
        h1[nf][0] = lo.x; h1[nf][1] = lo.y; h1[nf][2] = hi.x; h1[nf][3] = hi.y;
        const float2 vb = *reinterpret_cast<const float2*>(sm + L.vb0 + 8 * nf + 2 * c);
        d1[nf][0] = vb.x; d1[nf][1] = vb.y; d1[nf][2] = vb.x; d1[nf][3] = vb.y;
      }
      // ---- P0: raw rows -> X [t][k], X^T [k][t] (parity buffer), ones column at k = n ---------------
      {
        if (bulk_ok(tile)) {
          tma::mbar_wait(&bar, nbulk & 1);
          ++nbulk;
        } else {
          stream_rows<CHAIN_THREADS>(a.obs + t0 * n, rows * n, n, ctid, [&](int r, int k, float v) { sXR[r * n + k] = v; });
          bar_sync(BAR_CHAIN, CHAIN_THREADS);
        }
        const int cnt = rows * n;
        const int dq = CHAIN_THREADS / n, dr = CHAIN_THREADS - dq * n;
        int r = ctid / n, k = ctid - r * n;
        for (int i = ctid; i < cnt; i += CHAIN_THREADS) {
          const float v = (sXR[i] - sm[L.ishift + k]) / sm[L.iden + k];
          sX[r * LDX + k] = v;
          xt[k * LDT + r] = v;
          r += dq; k += dr;
          if (k >= n) { k -= n; ++r; }
        }
        if (rows < TS)
          for (int i = rows * n + ctid; i < tile_elems; i += CHAIN_THREADS) {
            const int rr = i / n, kk = i - rr * n;
            sX[rr * LDX + kk] = 0.f; xt[kk * LDT + rr] = 0.f;
          }
        if (ctid < TS) {
          const float one = ctid < rows ? 1.f : 0.f;
          sX[ctid * LDX + n] = one; xt[n * LDT + ctid] = one;
        }
      }
      bar_sync(BAR_CHAIN, CHAIN_THREADS);   // X visible; XR consumed
      if (ctid == 0 && tile + (int)gridDim.x < ntiles) start_copy(tile + gridDim.x);

      // ---- layer 0 tangent: dz1 = vb0 + X V0^T;  d1 = (1 - h1^2) dz1 --------------------------------
#pragma unroll
      for (int ks = 0; ks < NPK; ++ks) {
        const float2 xa = *reinterpret_cast<const float2*>(sX + rA * LDX + 8 * ks + 2 * c);
        const float2 xb = *reinterpret_cast<const float2*>(sX + rB * LDX + 8 * ks + 2 * c);
        Frag4 ax;
        split_a(ax, xa.x, xa.y, xb.x, xb.y);
#pragma unroll
        for (int n0 = 0; n0 < NF; n0 += 4) {
          Frag2 bv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) load_b(bv[i], sV0 + (8 * (n0 + i) + g) * LDX + 8 * ks + 2 * c, PL0);
#pragma unroll
          for (int term = 0; term < 3; ++term)
#pragma unroll
            for (int i = 0; i < 4; ++i) mma_term(d1[n0 + i], ax, bv[i], term);
        }
      }
#pragma unroll
      for (int nf = 0; nf < NF; ++nf)
#pragma unroll
        for (int e = 0; e < 4; ++e) d1[nf][e] = (1.f - h1[nf][e] * h1[nf][e]) * d1[nf][e];

      // ---- layer 1 tangent: dz2 = vb1 + H1 V1^T + D1 W1^T;  d2 = (1 - h2^2) dz2 ---------------------
      float h2[NF][4], d2[NF][4];
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        float2 lo = make_float2(0.f, 0.f), hi = lo;
        if (slab_ok) { lo = ld_cache(cslot + (2 * (NF + nf)) * 32); hi = ld_cache(cslot + (2 * (NF + nf) + 1) * 32); }
        h2[nf][0] = lo.x; h2[nf][1] = lo.y; h2[nf][2] = hi.x; h2[nf][3] = hi.y;
        const float2 vb = *reinterpret_cast<const float2*>(sm + L.vb1 + 8 * nf + 2 * c);
        d2[nf][0] = vb.x; d2[nf][1] = vb.y; d2[nf][2] = vb.x; d2[nf][3] = vb.y;
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
        for (int e = 0; e < 4; ++e) d2[nf][e] = (1.f - h2[nf][e] * h2[nf][e]) * d2[nf][e];

      // ---- head tangent: dy = H2 V2^T + D2 W2^T;  u = fisher * (dy + vb2), zero past N -------------
      float u[4];
      {
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
        const float2 fj = *reinterpret_cast<const float2*>(sm + L.fisher + 2 * c);
        const float2 vj = *reinterpret_cast<const float2*>(sm + L.vb2 + 2 * c);
        const bool va = rA < rows, vb = rB < rows;
        u[0] = va ? fj.x * (u[0] + vj.x) : 0.f; u[1] = va ? fj.y * (u[1] + vj.y) : 0.f;
        u[2] = vb ? fj.x * (u[2] + vj.x) : 0.f; u[3] = vb ? fj.y * (u[3] + vj.y) : 0.f;
      }
      // ---- gz2 = (U W2) * (1 - h2^2) ---------------------------------------------------------------
      float gz[NF][4];
      {
        Frag4 au;
        split_a(au, u);
#pragma unroll
        for (int n0 = 0; n0 < NF; n0 += 4) {
          Frag2 bw[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            load_b(bw[i], sW2T + (8 * (n0 + i) + g) * 8 + 2 * c, PL2T);
            gz[n0 + i][0] = gz[n0 + i][1] = gz[n0 + i][2] = gz[n0 + i][3] = 0.f;
          }
#pragma unroll
          for (int term = 0; term < 3; ++term)
#pragma unroll
            for (int i = 0; i < 4; ++i) mma_term(gz[n0 + i], au, bw[i], term);
        }
#pragma unroll
        for (int nf = 0; nf < NF; ++nf)
#pragma unroll
          for (int e = 0; e < 4; ++e) gz[nf][e] *= (1.f - h2[nf][e] * h2[nf][e]);
      }
      // ---- publish H1^T, GZ2^T, U^T once the wgrad warps are done with the previous tile ----------
      if (it > 0) bar_sync(BAR_FREE2, WTHREADS);
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        float* p0 = sPA + (8 * nf + 2 * c) * LDT;
        p0[rA] = h1[nf][0]; p0[LDT + rA] = h1[nf][1]; p0[rB] = h1[nf][2]; p0[LDT + rB] = h1[nf][3];
        float* q0 = sPB + (8 * nf + 2 * c) * LDT;
        q0[rA] = gz[nf][0]; q0[LDT + rA] = gz[nf][1]; q0[rB] = gz[nf][2]; q0[LDT + rB] = gz[nf][3];
      }
      {
        float* p0 = sUT + (2 * c) * LDT;
        p0[rA] = u[0]; p0[LDT + rA] = u[1]; p0[rB] = u[2]; p0[LDT + rB] = u[3];
      }
      bar_arrive(BAR_FULL1, WTHREADS);
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
          for (int n0 = 0; n0 < NF; n0 += 4) {
            Frag2 bw[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
              load_b_rows(bw[i], sW1 + (8 * ks + 2 * c) * LDW + 8 * (n0 + i) + g, LDW, PL1);
#pragma unroll
            for (int term = 0; term < 3; ++term)
#pragma unroll
              for (int i = 0; i < 4; ++i) mma_term(gh[n0 + i], ag, bw[i], term);
          }
        }
#pragma unroll
        for (int nf = 0; nf < NF; ++nf)
#pragma unroll
          for (int e = 0; e < 4; ++e) gz[nf][e] = gh[nf][e] * (1.f - h1[nf][e] * h1[nf][e]);
      }
      // ---- publish GZ1^T, H2^T once gW1 of this tile has consumed the planes ----------------------
      bar_sync(BAR_FREE1, WTHREADS);
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        float* p0 = sPB + (8 * nf + 2 * c) * LDT;
        p0[rA] = gz[nf][0]; p0[LDT + rA] = gz[nf][1]; p0[rB] = gz[nf][2]; p0[LDT + rB] = gz[nf][3];
        float* q0 = sPA + (8 * nf + 2 * c) * LDT;
        q0[rA] = h2[nf][0]; q0[LDT + rA] = h2[nf][1]; q0[rB] = h2[nf][2]; q0[LDT + rB] = h2[nf][3];
      }
      bar_arrive(BAR_FULL2, WTHREADS);
    }
    __syncthreads();   // (E1)
    __syncthreads();   // (E2) the wgrad warps have written the [d] staging vector
  }
  float* out = a.partials + (size_t)blockIdx.x * a.dstride;
  for (int i = tid; i < F.d; i += WTHREADS) out[i] = sG[i];
}

}  // namespace ws

int launch_fvp_ws(mjrl_ctx* ctx, const Args& a, int grid, cudaStream_t s) {
  const int hmax = a.h1 > a.h2 ? a.h1 : a.h2;
  if (hmax <= 32 || hmax > 64 || a.m > 8 || a.n + 1 > 16 || a.n + 1 <= 8) return -1;   // HP = 64, NPK = 2
  const ws::WLayout L(64, 2, a.n);
  const size_t smem = sizeof(float) * (size_t)L.total;
  const Flat F(a.n, a.m, a.h1, a.h2);
  if (smem + 64 > (size_t)ctx->max_smem_optin || F.d > 2 * 64 * ws::LDT) return -1;
  auto kern = ws::fvp_ws_kernel<64, 2>;
  static size_t attr = 0;
  if (smem > attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr = smem;
  }
  kern<<<grid, ws::WTHREADS, smem, s>>>(a);
  return mjrl_check_launch("fvp_ws_kernel");
}

}  // namespace pol
