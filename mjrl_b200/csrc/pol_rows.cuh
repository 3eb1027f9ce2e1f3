// Policy gradient (flat_vpg) and surrogate/KL evaluation in the "rows in registers" form of
// fvp_rows.cuh (included after it by policy.cu; same fragment conventions, same half-CTA tiling,
// same activation-cache layout, so a gradient pass can hand its h1/h2 to the FVPs that follow).
//
//   M_GRAD  batch_reinforce.py:37-55: forward, Gaussian log-likelihood, LR = exp(LL - LL_old),
//           weight c_t = LR_t w_t, backward of sum_t c_t through the network and log_std.
//   M_EVAL  batch_reinforce.py:37-49 + gaussian_mlp.py:124-140: surrogate sum and KL(old || new) sum.
#pragma once

namespace pol {
namespace rows {

struct GLayout {  // offsets in floats
  int NP, LDX, LDW;
  int X, XT, PA, PB, UT, ET, S0, S1, XR, W0, W1, W2, W2T, b0, b1, b2, oscale, oshift, sig, isig, ls, ols_, ishift, iden;
  int half_stride, total;
  __host__ __device__ GLayout(int HP, int NPK, int n, bool grad) {
    NP = 8 * NPK; LDX = ldx_of(NP); LDW = HP + 8;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 3) & ~3; return r; };
    PA = take(grad ? HP * LDT : 0); PB = take(grad ? HP * LDT : 0);
    X = take(TSR * LDX); XT = take(grad ? NP * LDT : 0); UT = take(grad ? 8 * LDT : 0); ET = take(grad ? 8 * LDT : 0);
    S0 = take(TSR); S1 = take(TSR);
    XR = take(2 * TSR * n);
    half_stride = o;
    o *= 2;
    W0 = take(2 * HP * LDX); W1 = take(2 * HP * LDW); W2 = take(2 * 8 * LDW); W2T = take(grad ? 2 * HP * 8 : 0);
    b0 = take(HP); b1 = take(HP); b2 = take(8); oscale = take(8); oshift = take(8); sig = take(8); isig = take(8);
    ls = take(8); ols_ = take(8); ishift = take(NP); iden = take(NP);
    total = o;
  }
};

template <int HP, int NPK, int MODE>
__global__ void __launch_bounds__(NTHREADS, 1) pol_rows_kernel(const Args a) {
  constexpr bool GRAD = MODE == M_GRAD;
  constexpr int NF = HP / 8, MT = HP / 16;
  constexpr int NP = 8 * NPK, LDX = ldx_of(NP), LDW = HP + 8;
  constexpr int NW1 = MT * NF / 4, NW0 = (MT * NPK + 3) / 4, NW2 = (NF + 3) / 4;
  constexpr int G = NF < 4 ? NF : 4;
  extern __shared__ __align__(16) float sm[];
  __shared__ uint64_t bar[2][2];

  const GLayout L(HP, NPK, a.n, GRAD);
  const Flat F(a.n, a.m, a.h1, a.h2);
  const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int half = tid >> 7, htid = tid & (HTHREADS - 1), hw = htid >> 5;
  const int n = a.n, m = a.m;
  float* hs = sm + (size_t)half * L.half_stride;
  float* sX = hs + L.X; float* sXT = hs + L.XT; float* sPA = hs + L.PA; float* sPB = hs + L.PB;
  float* sUT = hs + L.UT; float* sET = hs + L.ET; float* sS0 = hs + L.S0; float* sS1 = hs + L.S1;
  float* sXR = hs + L.XR;
  const float* sW0 = sm + L.W0; const float* sW1 = sm + L.W1; const float* sW2 = sm + L.W2;
  const float* sW2T = sm + L.W2T;
  constexpr int PL0 = HP * LDX, PL1 = HP * LDW, PL2 = 8 * LDW, PL2T = HP * 8;

  for (int i = tid; i < L.total; i += NTHREADS) sm[i] = 0.f;
  if (tid == 0) {
    tma::mbar_init(&bar[0][0], 1); tma::mbar_init(&bar[0][1], 1);
    tma::mbar_init(&bar[1][0], 1); tma::mbar_init(&bar[1][1], 1);
    tma::mbar_init_fence();
  }
  __syncthreads();
  {
    const float* P = a.params;
    auto put = [&](int at, int plane, float w) {
      uint32_t hi, lo;
      split_tf32(w, hi, lo);
      sm[at] = __uint_as_float(hi); sm[at + plane] = __uint_as_float(lo);
    };
    for (int i = tid; i < F.h1 * n; i += NTHREADS) {
      const int o = i / n, k = i - o * n;
      put(L.W0 + o * LDX + k, PL0, __ldg(P + F.oW0 + i));
    }
    for (int i = tid; i < F.h2 * F.h1; i += NTHREADS) {
      const int o = i / F.h1, k = i - o * F.h1;
      put(L.W1 + o * LDW + k, PL1, __ldg(P + F.oW1 + i));
    }
    for (int i = tid; i < m * F.h2; i += NTHREADS) {
      const int j = i / F.h2, k = i - j * F.h2;
      const float w = __ldg(P + F.oW2 + i);
      put(L.W2 + j * LDW + k, PL2, w);
      if (GRAD) put(L.W2T + k * 8 + j, PL2T, w);
    }
    for (int i = tid; i < F.h1; i += NTHREADS) sm[L.b0 + i] = __ldg(P + F.ob0 + i);
    for (int i = tid; i < F.h2; i += NTHREADS) sm[L.b1 + i] = __ldg(P + F.ob1 + i);
    for (int j = tid; j < 8; j += NTHREADS) {
      const bool in = j < m;
      const float lsv = in ? __ldg(P + F.ols + j) : 0.f;
      const float sg = expf(lsv);
      sm[L.b2 + j] = in ? __ldg(P + F.ob2 + j) : 0.f;
      sm[L.oscale + j] = (a.out_scale && in) ? __ldg(a.out_scale + j) : 1.f;
      sm[L.oshift + j] = (a.out_shift && in) ? __ldg(a.out_shift + j) : 0.f;
      sm[L.ls + j] = lsv; sm[L.sig + j] = sg;
      sm[L.ols_ + j] = (MODE == M_EVAL && in) ? __ldg(a.old_log_std + j) : 0.f;
    }
    for (int i = tid; i < NP; i += NTHREADS) {
      sm[L.ishift + i] = (a.in_shift && i < n) ? __ldg(a.in_shift + i) : 0.f;
      sm[L.iden + i] = (a.in_scale && i < n ? __ldg(a.in_scale + i) : 1.f) + 1e-8f;
    }
  }

  float accW1[GRAD ? NW1 : 1][4], accB1[4], accW0[GRAD ? NW0 : 1][4], accW2[GRAD ? NW2 : 1][4];
  float gb2acc[2] = {0.f, 0.f}, glsacc[2] = {0.f, 0.f};
  double surr_acc = 0.0, kl_acc = 0.0;   // warp 0 of each half, lane-replicated
  if (GRAD) {
#pragma unroll
    for (int j = 0; j < NW1; ++j) accW1[j][0] = accW1[j][1] = accW1[j][2] = accW1[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < NW0; ++j) accW0[j][0] = accW0[j][1] = accW0[j][2] = accW0[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < NW2; ++j) accW2[j][0] = accW2[j][1] = accW2[j][2] = accW2[j][3] = 0.f;
    accB1[0] = accB1[1] = accB1[2] = accB1[3] = 0.f;
  }
  const int mt1 = (hw * NW1) / NF, nf1 = (hw * NW1) % NF;
  const bool own_b1 = nf1 == 0;

  const int ntiles = (int)((a.N + TSR - 1) / TSR);
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
  const int tile_first = 2 * blockIdx.x + half, tile_step = 2 * gridDim.x;
  if (htid == 0 && tile_first < ntiles) start_copy(tile_first, 0);

  const int rA = 16 * hw + g, rB = rA + 8;
  // per-thread constants of its two action columns j = 2c, 2c+1
  const float2 b2j = *reinterpret_cast<const float2*>(sm + L.b2 + 2 * c);
  int it = 0;
  for (int tile = tile_first; tile < ntiles; tile += tile_step, ++it) {
    const int buf = it & 1;
    const long long t0 = (long long)tile * TSR;
    const int rows = tile_rows(tile);
    if (htid == 0 && tile + tile_step < ntiles) start_copy(tile + tile_step, buf ^ 1);
    // per-row inputs of the scalar tail: issued now, consumed after the layers
    const bool vA = rA < rows, vB = rB < rows;
    const long long gA = t0 + rA, gB = t0 + rB;
    float actA[2] = {0.f, 0.f}, actB[2] = {0.f, 0.f}, ollA = 0.f, ollB = 0.f, wA = 0.f, wB = 0.f;
    float omA[2] = {0.f, 0.f}, omB[2] = {0.f, 0.f};
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int j = 2 * c + e;
      if (j < m) {
        if (vA) actA[e] = ldg_stream(a.act + gA * m + j);
        if (vB) actB[e] = ldg_stream(a.act + gB * m + j);
        if (MODE == M_EVAL) {
          if (vA) omA[e] = ldg_stream(a.old_mean + gA * m + j);
          if (vB) omB[e] = ldg_stream(a.old_mean + gB * m + j);
        }
      }
    }
    if (a.old_ll) { if (vA) ollA = ldg_stream(a.old_ll + gA); if (vB) ollB = ldg_stream(a.old_ll + gB); }
    if (a.w) { if (vA) wA = ldg_stream(a.w + gA); if (vB) wB = ldg_stream(a.w + gB); }
    // ---- P0 -----------------------------------------------------------------------------------
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
        if (GRAD) sXT[k * LDT + r] = v;
        r += dq; k += dr;
        if (k >= n) { k -= n; ++r; }
      }
      if (rows < TSR)
        for (int i = rows * n + htid; i < tile_elems; i += HTHREADS) {
          const int rr = i / n, kk = i - rr * n;
          sX[rr * LDX + kk] = 0.f;
          if (GRAD) sXT[kk * LDT + rr] = 0.f;
        }
      if (GRAD && htid < TSR) {
        const float one = htid < rows ? 1.f : 0.f;
        sX[htid * LDX + n] = one; sXT[n * LDT + htid] = one;   // W0 has a zero column there
      }
    }
    half_sync(half);

    // ---- layer 0 --------------------------------------------------------------------------------
    float h1[NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 b = *reinterpret_cast<const float2*>(sm + L.b0 + 8 * nf + 2 * c);
      h1[nf][0] = b.x; h1[nf][1] = b.y; h1[nf][2] = b.x; h1[nf][3] = b.y;
    }
#pragma unroll
    for (int ks = 0; ks < NPK; ++ks) {
      const float2 xa = *reinterpret_cast<const float2*>(sX + rA * LDX + 8 * ks + 2 * c);
      const float2 xb = *reinterpret_cast<const float2*>(sX + rB * LDX + 8 * ks + 2 * c);
      Frag4 ax;
      split_a(ax, xa.x, xa.y, xb.x, xb.y);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bw[G];
#pragma unroll
        for (int i = 0; i < G; ++i) load_b(bw[i], sW0 + (8 * (n0 + i) + g) * LDX + 8 * ks + 2 * c, PL0);
#pragma unroll
        for (int term = 0; term < 3; ++term)
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(h1[n0 + i], ax, bw[i], term);
      }
    }
#pragma unroll
    for (int nf = 0; nf < NF; ++nf)
#pragma unroll
      for (int e = 0; e < 4; ++e) h1[nf][e] = tanh_fast(h1[nf][e]);
    float2* cslot = reinterpret_cast<float2*>(a.cache) + ((size_t)tile * 4 + hw) * (4 * NF * 32) + lane;
    // cache_mode 5: h1, h2 go straight into the tile order of the tensor-memory product (fvp_tc.cu: tiles of 96
    // timesteps, per tile an h1 and an h2 plane of 2 x [96 rows][128 B], 32-byte atoms swizzled by row mod 4);
    // the 4 lanes of a quad write the 32 contiguous bytes of one atom
    float* tcA = nullptr; float* tcB = nullptr;
    if (GRAD && a.cache && a.cache_mode == 5) {
      const long long tA = gA / 96, tB = gB / 96;
      const int rA96 = (int)(gA - tA * 96), rB96 = (int)(gB - tB * 96);
      tcA = a.cache + (size_t)tA * (2 * 96 * 64) + rA96 * 32 + 2 * c;
      tcB = a.cache + (size_t)tB * (2 * 96 * 64) + rB96 * 32 + 2 * c;
    }
    auto tc_off = [](int nf, int r96) { return (nf >> 2) * (96 * 32) + (((nf & 3) ^ (r96 & 3)) << 3); };
    const int sA = (int)(gA % 96) & 3, sB = (int)(gB % 96) & 3;
    if (GRAD) {
      if (tcA) {
#pragma unroll
        for (int nf = 0; nf < NF; ++nf) {
          *reinterpret_cast<float2*>(tcA + tc_off(nf, sA)) = make_float2(h1[nf][0], h1[nf][1]);
          *reinterpret_cast<float2*>(tcB + tc_off(nf, sB)) = make_float2(h1[nf][2], h1[nf][3]);
        }
      } else if (a.cache) {
#pragma unroll
        for (int nf = 0; nf < NF; ++nf) {
          cslot[(2 * nf) * 32] = make_float2(h1[nf][0], h1[nf][1]);
          cslot[(2 * nf + 1) * 32] = make_float2(h1[nf][2], h1[nf][3]);
        }
      }
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        float* p0 = sPA + (8 * nf + 2 * c) * LDT;
        p0[rA] = h1[nf][0]; p0[LDT + rA] = h1[nf][1]; p0[rB] = h1[nf][2]; p0[LDT + rB] = h1[nf][3];
      }
    }

    // ---- layer 1 --------------------------------------------------------------------------------
    float h2[NF][4];
#pragma unroll
    for (int nf = 0; nf < NF; ++nf) {
      const float2 b = *reinterpret_cast<const float2*>(sm + L.b1 + 8 * nf + 2 * c);
      h2[nf][0] = b.x; h2[nf][1] = b.y; h2[nf][2] = b.x; h2[nf][3] = b.y;
    }
#pragma unroll
    for (int ks = 0; ks < NF; ++ks) {
      Frag4 ah;
      split_a(ah, h1[ks]);
#pragma unroll
      for (int n0 = 0; n0 < NF; n0 += G) {
        Frag2 bw[G];
#pragma unroll
        for (int i = 0; i < G; ++i) load_b(bw[i], sW1 + (8 * (n0 + i) + g) * LDW + 8 * ks + 2 * c, PL1);
#pragma unroll
        for (int term = 0; term < 3; ++term)
#pragma unroll
          for (int i = 0; i < G; ++i) mma_term(h2[n0 + i], ah, bw[i], term);
      }
    }
#pragma unroll
    for (int nf = 0; nf < NF; ++nf)
#pragma unroll
      for (int e = 0; e < 4; ++e) h2[nf][e] = tanh_fast(h2[nf][e]);
    if (GRAD && tcA) {
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        *reinterpret_cast<float2*>(tcA + 96 * 64 + tc_off(nf, sA)) = make_float2(h2[nf][0], h2[nf][1]);
        *reinterpret_cast<float2*>(tcB + 96 * 64 + tc_off(nf, sB)) = make_float2(h2[nf][2], h2[nf][3]);
      }
    } else if (GRAD && a.cache) {
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        cslot[(2 * (NF + nf)) * 32] = make_float2(h2[nf][0], h2[nf][1]);
        cslot[(2 * (NF + nf) + 1) * 32] = make_float2(h2[nf][2], h2[nf][3]);
      }
    }

    // ---- head: mu = (H2 W2^T + b2) * out_scale + out_shift ---------------------------------------
    float mu[4];
    {
      float yy[2][4];
#pragma unroll
      for (int x = 0; x < 2; ++x) yy[x][0] = yy[x][1] = yy[x][2] = yy[x][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < NF; ++ks) {
        Frag4 ah;
        Frag2 bw;
        split_a(ah, h2[ks]);
        load_b(bw, sW2 + g * LDW + 8 * ks + 2 * c, PL2);
#pragma unroll
        for (int term = 0; term < 3; ++term) mma_term(yy[ks & 1], ah, bw, term);
      }
      const float2 osc = *reinterpret_cast<const float2*>(sm + L.oscale + 2 * c);
      const float2 osh = *reinterpret_cast<const float2*>(sm + L.oshift + 2 * c);
      mu[0] = ((yy[0][0] + yy[1][0]) + b2j.x) * osc.x + osh.x;
      mu[1] = ((yy[0][1] + yy[1][1]) + b2j.y) * osc.y + osh.y;
      mu[2] = ((yy[0][2] + yy[1][2]) + b2j.x) * osc.x + osh.x;
      mu[3] = ((yy[0][3] + yy[1][3]) + b2j.y) * osc.y + osh.y;
    }
    // ---- Gaussian tail in fragment layout: this thread holds columns j = 2c, 2c+1 of rows rA, rB;
    // sums over j are quad reductions (lanes of equal g) -----------------------------------------
    float u[4] = {0.f, 0.f, 0.f, 0.f}, ev[4] = {0.f, 0.f, 0.f, 0.f};
    {
      float zz[4], ssA = 0.f, ssB = 0.f, sls = 0.f, klA = 0.f, klB = 0.f;
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = 2 * c + e;
        const bool in = j < m;
        const float sg = sm[L.sig + j], lsj = sm[L.ls + j];
        zz[e] = in ? (actA[e] - mu[e]) / sg : 0.f;
        zz[2 + e] = in ? (actB[e] - mu[2 + e]) / sg : 0.f;
        ssA = fmaf(zz[e], zz[e], ssA);
        ssB = fmaf(zz[2 + e], zz[2 + e], ssB);
        if (in) sls += lsj;
        if (in && a.mean_out) {
          if (vA) a.mean_out[gA * m + j] = mu[e];
          if (vB) a.mean_out[gB * m + j] = mu[2 + e];
        }
        if (MODE == M_EVAL && in) {
          // gaussian_mlp.py:137-139, KL(old || new) per dimension
          const float so = expf(sm[L.ols_ + j]);
          const float dr = 2.f * sg * sg + 1e-8f;
          const float dA = omA[e] - mu[e], dB = omB[e] - mu[2 + e];
          klA += (dA * dA + so * so - sg * sg) / dr + lsj - sm[L.ols_ + j];
          klB += (dB * dB + so * so - sg * sg) / dr + lsj - sm[L.ols_ + j];
        }
      }
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1) {
        ssA += __shfl_xor_sync(0xffffffffu, ssA, o); ssB += __shfl_xor_sync(0xffffffffu, ssB, o);
        sls += __shfl_xor_sync(0xffffffffu, sls, o);
        if (MODE == M_EVAL) { klA += __shfl_xor_sync(0xffffffffu, klA, o); klB += __shfl_xor_sync(0xffffffffu, klB, o); }
      }
      // gaussian_mlp.py:107-109: (-0.5*sum z^2 + -sum log_std) + -0.5*m*log(2 pi)
      const float cst = (float)(-0.5 * (double)m * 1.8378770664093453);
      const float llA = (-0.5f * ssA + -sls) + cst, llB = (-0.5f * ssB + -sls) + cst;
      if (a.ll_out && c == 0) { if (vA) a.ll_out[gA] = llA; if (vB) a.ll_out[gB] = llB; }
      const float lrA = (a.old_ll && vA) ? expf(llA - ollA) : 1.f;
      const float lrB = (a.old_ll && vB) ? expf(llB - ollB) : 1.f;
      const float cA = vA ? lrA * wA : 0.f, cB = vB ? lrB * wB : 0.f;
      if (c == 0) {
        sS0[rA] = cA; sS0[rB] = cB;
        if (MODE == M_EVAL) { sS1[rA] = vA ? klA : 0.f; sS1[rB] = vB ? klB : 0.f; }
      }
      if (GRAD) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 2 * c + e;
          if (j < m) {
            const float sg = sm[L.sig + j], osc = sm[L.oscale + j];
            u[e] = cA * (zz[e] / sg) * osc;           // dL/dy_j
            u[2 + e] = cB * (zz[2 + e] / sg) * osc;
            ev[e] = cA * (zz[e] * zz[e] - 1.f);         // dL/dlog_std_j
            ev[2 + e] = cB * (zz[2 + e] * zz[2 + e] - 1.f);
          }
        }
        float* p0 = sUT + (2 * c) * LDT;
        p0[rA] = u[0]; p0[LDT + rA] = u[1]; p0[rB] = u[2]; p0[LDT + rB] = u[3];
        float* q0 = sET + (2 * c) * LDT;
        q0[rA] = ev[0]; q0[LDT + rA] = ev[1]; q0[rB] = ev[2]; q0[LDT + rB] = ev[3];
      }
    }

    float gz[GRAD ? NF : 1][4];
    if (GRAD) {
      // ---- gz2 = (U W2) * (1 - h2^2) ------------------------------------------------------------
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
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
#pragma unroll
        for (int e = 0; e < 4; ++e) gz[nf][e] *= (1.f - h2[nf][e] * h2[nf][e]);
        float* p0 = sPB + (8 * nf + 2 * c) * LDT;
        p0[rA] = gz[nf][0]; p0[LDT + rA] = gz[nf][1]; p0[rB] = gz[nf][2]; p0[LDT + rB] = gz[nf][3];
      }
      // ---- gz1 = (GZ2 W1) * (1 - h1^2) ----------------------------------------------------------
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
            load_b_rows(bw[i], sW1 + (8 * ks + 2 * c) * LDW + 8 * (n0 + i) + g, LDW, PL1);
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
    half_sync(half);   // (A) planes, U^T, E^T, per-row terms visible

    if (hw == 0) {   // fixed-order fp64 sums of the tile's per-timestep terms
      double s0 = (double)sS0[lane] + (double)sS0[lane + 32], s1 = 0.0;
      if (MODE == M_EVAL) s1 = (double)sS1[lane] + (double)sS1[lane + 32];
      s0 = warp_sum(s0);
      if (MODE == M_EVAL) s1 = warp_sum(s1);
      surr_acc += s0; kl_acc += s1;
    }
    if (GRAD) {
      // ---- gW1 (+ gb1 through a ones column) ----------------------------------------------------
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
      half_sync(half);   // (B)
#pragma unroll
      for (int nf = 0; nf < NF; ++nf) {
        float* p0 = sPB + (8 * nf + 2 * c) * LDT;
        p0[rA] = gz[nf][0]; p0[LDT + rA] = gz[nf][1]; p0[rB] = gz[nf][2]; p0[LDT + rB] = gz[nf][3];
        float* q0 = sPA + (8 * nf + 2 * c) * LDT;
        q0[rA] = h2[nf][0]; q0[LDT + rA] = h2[nf][1]; q0[rB] = h2[nf][2]; q0[LDT + rB] = h2[nf][3];
      }
      half_sync(half);   // (C)
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
#pragma unroll
      for (int s = 0; s < 2; ++s) {   // gb2[j] += sum_t U[t][j], gls[j] += sum_t E[t][j], j = hw + 4 s
        const int j = hw + 4 * s;
        if (j < m) {
          const float2 v = *reinterpret_cast<const float2*>(sUT + j * LDT + 2 * lane);
          const float2 w = *reinterpret_cast<const float2*>(sET + j * LDT + 2 * lane);
          float su = v.x + v.y, se = w.x + w.y;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            su += __shfl_xor_sync(0xffffffffu, su, o);
            se += __shfl_xor_sync(0xffffffffu, se, o);
          }
          gb2acc[s] += su; glsacc[s] += se;
        }
      }
    }
    half_sync(half);   // (D)
  }

  // ---- per-CTA results ----------------------------------------------------------------------------
  __shared__ double sred[2][2];
  if (hw == 0 && lane == 0) { sred[half][0] = surr_acc; sred[half][1] = kl_acc; }
  __syncthreads();
  if (tid == 0) {
    a.spartials[2 * blockIdx.x + 0] = sred[0][0] + sred[1][0];
    a.spartials[2 * blockIdx.x + 1] = sred[0][1] + sred[1][1];
  }
  if (GRAD) {
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
          if (hw + 4 * s < m && lane == 0) {
            sG[F.ob2 + hw + 4 * s] += gb2acc[s];
            sG[F.ols + hw + 4 * s] += glsacc[s];
          }
      }
      __syncthreads();
    }
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    for (int i = tid; i < F.d; i += NTHREADS) out[i] = sG[i];
  }
}

}  // namespace rows
}  // namespace pol
