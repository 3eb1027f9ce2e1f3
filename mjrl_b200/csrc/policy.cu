// (c)(d) Fused Gaussian-MLP passes over timestep tiles: forward/log-likelihood, surrogate+KL
// evaluation, vanilla policy gradient, and the analytic Fisher-vector product.
//
// Replaces (reference = bennevans/mjrl):
//   MuNet.forward + mean_LL          mjrl/policies/gaussian_mlp.py:100-110,176-182
//   likelihood_ratio / mean_kl       mjrl/policies/gaussian_mlp.py:124-140
//   CPI_surrogate / kl_old_new / flat_vpg   mjrl/algos/batch_reinforce.py:37-55 (torch autograd)
//   NPG.HVP (double backprop)        mjrl/algos/npg_cg.py:55-74
//
// One persistent CTA per SM walks tiles of TS timesteps (tile = blockIdx.x + k*gridDim.x).
// Per tile the obs rows are read ONCE from HBM (coalesced, contiguous TS*n floats), every
// intermediate stays in shared memory / registers, and nothing but per-CTA partial sums is
// written back: algorithmic bytes 4n (FVP), 4n+8m+12 (grad), 4n+8m+12 (eval) per timestep.
// These passes are fp32-FMA bound, not HBM bound.
#include <stdlib.h>

#include "fvp_tc.cuh"
#include "policy_tile.cuh"

namespace pol {

// Shared-memory carve-up (offsets in floats), identical on host and device.
struct Layout {
  int TS, LDH, LDW, NP, LDX, MP, LDM;
  int X, H1, H2, B1, B2, U, E, ACT, S0, S1, XR, AR;
  int W0t, b0, W1t, b1, W2t, b2, W1, W2, V0t, vb0, V1t, vb1, V2t, vb2;
  int ishift, iden, oscale, oshift, sig, ls, ols_, fisher;
  int total;
  // mma: tensor-core path.  obs width padded to 8 (k-steps of the m16n8k8 MMA), k-major weight
  // stride HP+8 (= 8 mod 32) so B-fragment loads are conflict free; activations keep HP+4 (= 4 mod 8).
  // prefetch: raw staging of the NEXT tile's obs (and actions), filled by cp.async while the
  // current tile computes.
  __host__ __device__ Layout(int HP, int n, int m, int mode, bool mma, bool prefetch = false) {
    const int NF = HP / 4, NT = NTHREADS / NF;
    TS = RT * NT; LDH = HP + 4; LDW = mma ? HP + 8 : HP + 4;
    NP = mma ? ((n + 7) & ~7) : round4(n); LDX = NP + 4; MP = round4(m); LDM = MP + 4;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 3) & ~3; return r; };
    X = take(TS * LDX);
    H1 = take(TS * LDH); H2 = take(TS * LDH);
    const bool bwd = mode == M_GRAD || mode == M_FVP;
    B1 = bwd ? take(TS * LDH) : 0; B2 = bwd ? take(TS * LDH) : 0;
    U = take(TS * LDM);
    E = mode == M_GRAD ? take(TS * LDM) : 0;
    ACT = mode != M_FVP ? take(TS * LDM) : 0;
    S0 = take(TS); S1 = take(TS);
    XR = prefetch ? take(TS * n) : 0;
    AR = (prefetch && mode != M_FVP) ? take(TS * m) : 0;
    W0t = take(NP * LDW); b0 = take(HP); W1t = take(HP * LDW); b1 = take(HP);
    W2t = take(HP * LDM); b2 = take(MP);
    W1 = bwd ? take(HP * LDW) : 0; W2 = bwd ? take(MP * LDH) : 0;
    const bool fvp = mode == M_FVP;
    V0t = fvp ? take(NP * LDW) : 0; vb0 = fvp ? take(HP) : 0; V1t = fvp ? take(HP * LDW) : 0;
    vb1 = fvp ? take(HP) : 0; V2t = fvp ? take(HP * LDM) : 0; vb2 = fvp ? take(MP) : 0;
    ishift = take(NP); iden = take(NP); oscale = take(MP); oshift = take(MP); sig = take(MP);
    ls = take(MP); ols_ = take(MP); fisher = take(MP);
    total = o;
  }
};

// k-major (transposed) copy of a row-major (rows_out x cols_in) weight block of the flat vector:
// dst[k][f] = src[f*cols_in + k], zero padded to [Kpad][ld].
__device__ __forceinline__ void load_wt(float* dst, int ld, int Kpad, int Fpad, const float* src,
                                        int rows_out, int cols_in) {
  for (int i = threadIdx.x; i < Kpad * Fpad; i += NTHREADS) {
    int f = i / Kpad, k = i - f * Kpad;  // consecutive threads walk k: coalesced global reads
    float v = (f < rows_out && k < cols_in) ? __ldg(src + f * cols_in + k) : 0.f;
    dst[k * ld + f] = v;
  }
}
// natural copy: dst[r][c] = src[r*cols + c], zero padded to [Rpad][ld] (first Cpad columns)
__device__ __forceinline__ void load_w(float* dst, int ld, int Rpad, int Cpad, const float* src,
                                       int rows, int cols) {
  for (int i = threadIdx.x; i < Rpad * Cpad; i += NTHREADS) {
    int r = i / Cpad, c = i - r * Cpad;
    dst[r * ld + c] = (r < rows && c < cols) ? __ldg(src + r * cols + c) : 0.f;
  }
}
__device__ __forceinline__ void load_v(float* dst, int pad, const float* src, int cnt, float fill = 0.f) {
  for (int i = threadIdx.x; i < pad; i += NTHREADS) dst[i] = (src && i < cnt) ? __ldg(src + i) : fill;
}

// Tensor-core forward layer on the warp's 32 x 16 tile: z = bias + A Wt (tangent dz = vbias + A Vt
// [+ D Wt]), h = tanh(z), dh = (1 - h^2) dz, written to Hout / Dout ([t][ldo]).
template <bool DUAL, bool HAS_D>
__device__ __forceinline__ void mma_fwd_layer(const float* __restrict__ A, const float* __restrict__ D, int lda,
                                              const float* __restrict__ Wt, const float* __restrict__ Vt, int ldw,
                                              int K, const float* __restrict__ bias, const float* __restrict__ vbias,
                                              float* __restrict__ Hout, float* __restrict__ Dout, int ldo,
                                              int mrow0, int ncol0, int g, int c) {
  float z[2][2][4], dz[2][2][4];
#pragma unroll
  for (int ni = 0; ni < 2; ++ni) {
    const int col = ncol0 + ni * 8 + 2 * c;
    const float b0 = bias[col], b1 = bias[col + 1];
    const float v0 = DUAL ? vbias[col] : 0.f, v1 = DUAL ? vbias[col + 1] : 0.f;
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      z[mi][ni][0] = b0; z[mi][ni][1] = b1; z[mi][ni][2] = b0; z[mi][ni][3] = b1;
      dz[mi][ni][0] = v0; dz[mi][ni][1] = v1; dz[mi][ni][2] = v0; dz[mi][ni][3] = v1;
    }
  }
  mma_layer<DUAL, HAS_D>(z, dz, A + mrow0 * lda, HAS_D ? D + mrow0 * lda : nullptr, lda, Wt + ncol0,
                         DUAL ? Vt + ncol0 : nullptr, ldw, K, g, c);
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int ni = 0; ni < 2; ++ni)
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int row = mrow0 + mi * 16 + g + half * 8, col = ncol0 + ni * 8 + 2 * c;
        const float h0 = tanh_fast(z[mi][ni][2 * half]), h1 = tanh_fast(z[mi][ni][2 * half + 1]);
        *reinterpret_cast<float2*>(Hout + row * ldo + col) = make_float2(h0, h1);
        if (DUAL)
          *reinterpret_cast<float2*>(Dout + row * ldo + col) =
              make_float2((1.f - h0 * h0) * dz[mi][ni][2 * half], (1.f - h1 * h1) * dz[mi][ni][2 * half + 1]);
      }
}

template <int HP, int MODE, int W0SLOTS, bool MMA>
__global__ void __launch_bounds__(NTHREADS, 1) mlp_kernel(const Args a) {
  constexpr int NF = HP / 4, NT = NTHREADS / NF, TS = RT * NT, LDH = HP + 4;
  constexpr bool BWD = MODE == M_GRAD || MODE == M_FVP;
  extern __shared__ __align__(16) float sm[];
  if (MODE == M_FVP && a.stop_flag && *a.stop_flag) return;

  const Layout L(HP, a.n, a.m, MODE, MMA, a.prefetch != 0);
  const Flat F(a.n, a.m, a.h1, a.h2);
  const int tid = threadIdx.x;
  const int NP = L.NP, LDX = L.LDX, MP = L.MP, LDM = L.LDM;
  constexpr int LDW = MMA ? HP + 8 : HP + 4;
  const int n = a.n, m = a.m;

  float* sX = sm + L.X; float* sH1 = sm + L.H1; float* sH2 = sm + L.H2;
  float* sB1 = sm + L.B1; float* sB2 = sm + L.B2; float* sU = sm + L.U; float* sE = sm + L.E;
  float* sAct = sm + L.ACT; float* sS0 = sm + L.S0; float* sS1 = sm + L.S1;

  // ---- one-time: zero activations (padding columns stay zero), stage weights -------------
  for (int i = tid; i < L.W0t; i += NTHREADS) sm[i] = 0.f;
  const float* P = a.params;
  load_wt(sm + L.W0t, LDW, NP, HP, P + F.oW0, F.h1, n);
  load_v(sm + L.b0, HP, P + F.ob0, F.h1);
  load_wt(sm + L.W1t, LDW, HP, HP, P + F.oW1, F.h2, F.h1);
  load_v(sm + L.b1, HP, P + F.ob1, F.h2);
  load_wt(sm + L.W2t, LDM, HP, MP, P + F.oW2, m, F.h2);
  load_v(sm + L.b2, MP, P + F.ob2, m);
  if (BWD) {
    load_w(sm + L.W1, LDW, HP, HP, P + F.oW1, F.h2, F.h1);
    load_w(sm + L.W2, LDH, MP, HP, P + F.oW2, m, F.h2);
  }
  if (MODE == M_FVP) {
    const float* V = a.vec;
    load_wt(sm + L.V0t, LDW, NP, HP, V + F.oW0, F.h1, n);
    load_v(sm + L.vb0, HP, V + F.ob0, F.h1);
    load_wt(sm + L.V1t, LDW, HP, HP, V + F.oW1, F.h2, F.h1);
    load_v(sm + L.vb1, HP, V + F.ob1, F.h2);
    load_wt(sm + L.V2t, LDM, HP, MP, V + F.oW2, m, F.h2);
    load_v(sm + L.vb2, MP, V + F.ob2, m);
  }
  load_v(sm + L.ishift, NP, a.in_shift, n, 0.f);
  for (int i = tid; i < NP; i += NTHREADS)
    sm[L.iden + i] = (a.in_scale && i < n ? __ldg(a.in_scale + i) : 1.f) + 1e-8f;
  load_v(sm + L.oscale, MP, a.out_scale, m, 1.f);
  load_v(sm + L.oshift, MP, a.out_shift, m, 0.f);
  for (int j = tid; j < MP; j += NTHREADS) {
    float ls = j < m ? __ldg(P + F.ols + j) : 0.f;
    float sg = expf(ls);
    sm[L.ls + j] = ls; sm[L.sig + j] = sg;
    float os = (a.out_scale && j < m) ? __ldg(a.out_scale + j) : 1.f;
    // d^2 KL / d mu^2 = 2/(2 sigma^2 + 1e-8)  (gaussian_mlp.py:137-138), chained through out_scale twice
    sm[L.fisher + j] = j < m ? os * os * (2.f / (2.f * sg * sg + 1e-8f)) : 0.f;
    sm[L.ols_ + j] = (MODE == M_EVAL && j < m) ? __ldg(a.old_log_std + j) : 0.f;
  }
  __syncthreads();

  const int tf = tid % NF, tt = tid / NF;  // feature tile / timestep group of the layer products
  const int row0 = tt * RT, col0 = tf * 4;

  // tensor-core path: warp (wm, wn) owns a 32(t) x 16(f) tile of every layer output
  constexpr int WN = HP / 16, WM = 8 / WN;
  static_assert(!MMA || 32 * WM == TS, "MMA warp tiling must cover the timestep tile");
  const int lane = tid & 31, warp = tid >> 5, fg = lane >> 2, fc = lane & 3;
  const int wm = warp % WM, wn = warp / WM;
  const int mrow0 = wm * 32, ncol0 = wn * 16;
  // weight-gradient fragments (16 x 8): gW1 has (HP/16)*(HP/8) of them, NW1 per warp, all in one
  // 16-row band; gW0 has (HP/16)*(NP/8), dealt round-robin to warps (slot = frag / 8)
  constexpr int NW1 = (HP / 16) * (HP / 8) / 8;
  float macc1[MMA ? NW1 : 1][4], macc0[MMA ? W0SLOTS : 1][4], mcb0[4];

  // persistent register accumulators
  float acc0[MMA ? 1 : W0SLOTS][16], acc1[16], acc2[16], cb0[4], cb1[4];
  float gb2acc[4] = {0.f, 0.f, 0.f, 0.f}, glsacc[4] = {0.f, 0.f, 0.f, 0.f};  // warp w: dims w, w+8, ... (lane 0)
  double surr_acc = 0.0, kl_acc = 0.0;
  if (BWD) {
#pragma unroll
    for (int s = 0; s < (MMA ? NW1 : 1); ++s) macc1[s][0] = macc1[s][1] = macc1[s][2] = macc1[s][3] = 0.f;
#pragma unroll
    for (int s = 0; s < (MMA ? W0SLOTS : 1); ++s) macc0[s][0] = macc0[s][1] = macc0[s][2] = macc0[s][3] = 0.f;
    mcb0[0] = mcb0[1] = mcb0[2] = mcb0[3] = 0.f;
#pragma unroll
    for (int s = 0; s < (MMA ? 1 : W0SLOTS); ++s)
#pragma unroll
      for (int e = 0; e < 16; ++e) acc0[s][e] = 0.f;
#pragma unroll
    for (int e = 0; e < 16; ++e) { acc1[e] = 0.f; acc2[e] = 0.f; }
#pragma unroll
    for (int c = 0; c < 4; ++c) { cb0[c] = 0.f; cb1[c] = 0.f; }
  }
  const int nt0 = NF * (NP / 4), nt1 = NF * NF, nt2 = (MP / 4) * NF;
  const TileOwner own1(nt1, tid, 0), own2(nt2, tid, 0);

  // optional per-phase cycle counters (debug: MJRL_B200_PHASE_CYCLES=1), thread 0 of each CTA
  long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long tprev = 0;
#define PHASE_STAMP(q)                                   \
  if (a.phase_cycles && tid == 0) {                      \
    const long long now = clock64();                     \
    ph[q] += now - tprev;                                \
    tprev = now;                                         \
  }
  if (a.phase_cycles && tid == 0) tprev = clock64();
  const long long ntiles = (a.N + TS - 1) / TS;
  // cp.async staging of a tile's raw obs (and action) rows: 16 B chunks when the source is
  // 16 B aligned (TS*n*4 is a multiple of 16, so every tile start is), 4 B copies otherwise.
  auto stage_rows = [&](float* dst, const float* src, int cnt) {
    const bool al = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    const int nvec = al ? (cnt >> 2) : 0;
    for (int i = tid; i < nvec; i += NTHREADS)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                   :: "r"((uint32_t)__cvta_generic_to_shared(dst + 4 * i)), "l"(src + 4 * i) : "memory");
    for (int i = 4 * nvec + tid; i < cnt; i += NTHREADS)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;"
                   :: "r"((uint32_t)__cvta_generic_to_shared(dst + i)), "l"(src + i) : "memory");
  };
  auto stage_tile = [&](long long tile) {
    const long long t0 = tile * TS;
    const int rows = (int)min((long long)TS, a.N - t0);
    stage_rows(sm + L.XR, a.obs + t0 * n, rows * n);
    if (MODE != M_FVP && a.act) stage_rows(sm + L.AR, a.act + t0 * m, rows * m);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (a.prefetch && (long long)blockIdx.x < ntiles) stage_tile(blockIdx.x);
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long t0 = tile * TS;
    const int rows = (int)min((long long)TS, a.N - t0);
    if (a.prefetch) asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // previous tile fully consumed; staged rows of this tile visible
    PHASE_STAMP(7);
    // ---- P0: obs tile (+ actions) -> smem, input transform --------------------------------
    if (a.prefetch) {
      // this tile's raw rows were staged by cp.async during the previous tile
      float* sXR = sm + L.XR; float* sAR = sm + L.AR;
      const int cnt = rows * n;
      for (int i = tid; i < cnt; i += NTHREADS) {
        int r = i / n, c = i - r * n;
        sX[r * LDX + c] = (sXR[i] - sm[L.ishift + c]) / sm[L.iden + c];
      }
      if (rows < TS)
        for (int i = tid; i < (TS - rows) * NP; i += NTHREADS) {
          int r = rows + i / NP, c = i % NP;
          sX[r * LDX + c] = 0.f;
        }
      if (MODE != M_FVP && a.act)
        for (int i = tid; i < rows * m; i += NTHREADS) {
          int r = i / m, c = i - r * m;
          sAct[r * LDM + c] = sAR[i];
        }
    } else {
      const float* src = a.obs + t0 * n;
      const int cnt = rows * n;
      for (int i = tid; i < cnt; i += NTHREADS) {
        int r = i / n, c = i - r * n;
        sX[r * LDX + c] = (ld_stream(src + i) - sm[L.ishift + c]) / sm[L.iden + c];
      }
      if (rows < TS)
        for (int i = tid; i < (TS - rows) * NP; i += NTHREADS) {
          int r = rows + i / NP, c = i % NP;
          sX[r * LDX + c] = 0.f;
        }
      if (MODE != M_FVP && a.act) {
        const float* as = a.act + t0 * m;
        for (int i = tid; i < rows * m; i += NTHREADS) {
          int r = i / m, c = i - r * m;
          sAct[r * LDM + c] = ld_stream(as + i);
        }
      }
    }
    __syncthreads();
    if (a.prefetch && tile + gridDim.x < ntiles) stage_tile(tile + gridDim.x);
    PHASE_STAMP(0);
    // ---- P1: layer 0 ----------------------------------------------------------------------
    if constexpr (MMA) {
      mma_fwd_layer<MODE == M_FVP, false>(sX, nullptr, LDX, sm + L.W0t, sm + L.V0t, LDW, NP, sm + L.b0,
                                          sm + L.vb0, sH1, sB1, LDH, mrow0, ncol0, fg, fc);
    } else {
      float z[RT][4], dz[RT][4];
      const float4 bias = *reinterpret_cast<const float4*>(sm + L.b0 + col0);
#pragma unroll
      for (int r = 0; r < RT; ++r) { z[r][0] = bias.x; z[r][1] = bias.y; z[r][2] = bias.z; z[r][3] = bias.w; }
      if (MODE == M_FVP) {
        const float4 vb = *reinterpret_cast<const float4*>(sm + L.vb0 + col0);
#pragma unroll
        for (int r = 0; r < RT; ++r) { dz[r][0] = vb.x; dz[r][1] = vb.y; dz[r][2] = vb.z; dz[r][3] = vb.w; }
        mm_acc_dual<RT, false>(z, dz, sX + row0 * LDX, nullptr, LDX, sm + L.W0t + col0,
                               sm + L.V0t + col0, LDW, NP);
      } else {
        mm_acc<RT>(z, sX + row0 * LDX, LDX, sm + L.W0t + col0, LDW, NP);
      }
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        float4 h = make_float4(tanh_fast(z[r][0]), tanh_fast(z[r][1]), tanh_fast(z[r][2]), tanh_fast(z[r][3]));
        *reinterpret_cast<float4*>(sH1 + (row0 + r) * LDH + col0) = h;
        if (MODE == M_FVP) {
          float4 d = make_float4((1.f - h.x * h.x) * dz[r][0], (1.f - h.y * h.y) * dz[r][1],
                                 (1.f - h.z * h.z) * dz[r][2], (1.f - h.w * h.w) * dz[r][3]);
          *reinterpret_cast<float4*>(sB1 + (row0 + r) * LDH + col0) = d;
        }
      }
    }
    __syncthreads();
    PHASE_STAMP(1);
    // ---- P2: layer 1 ----------------------------------------------------------------------
    if constexpr (MMA) {
      mma_fwd_layer<MODE == M_FVP, MODE == M_FVP>(sH1, sB1, LDH, sm + L.W1t, sm + L.V1t, LDW, HP,
                                                  sm + L.b1, sm + L.vb1, sH2, sB2, LDH, mrow0, ncol0, fg, fc);
    } else {
      float z[RT][4], dz[RT][4];
      const float4 bias = *reinterpret_cast<const float4*>(sm + L.b1 + col0);
#pragma unroll
      for (int r = 0; r < RT; ++r) { z[r][0] = bias.x; z[r][1] = bias.y; z[r][2] = bias.z; z[r][3] = bias.w; }
      if (MODE == M_FVP) {
        const float4 vb = *reinterpret_cast<const float4*>(sm + L.vb1 + col0);
#pragma unroll
        for (int r = 0; r < RT; ++r) { dz[r][0] = vb.x; dz[r][1] = vb.y; dz[r][2] = vb.z; dz[r][3] = vb.w; }
        mm_acc_dual<RT, true>(z, dz, sH1 + row0 * LDH, sB1 + row0 * LDH, LDH, sm + L.W1t + col0,
                              sm + L.V1t + col0, LDW, HP);
      } else {
        mm_acc<RT>(z, sH1 + row0 * LDH, LDH, sm + L.W1t + col0, LDW, HP);
      }
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        float4 h = make_float4(tanh_fast(z[r][0]), tanh_fast(z[r][1]), tanh_fast(z[r][2]), tanh_fast(z[r][3]));
        *reinterpret_cast<float4*>(sH2 + (row0 + r) * LDH + col0) = h;
        if (MODE == M_FVP) {
          float4 d = make_float4((1.f - h.x * h.x) * dz[r][0], (1.f - h.y * h.y) * dz[r][1],
                                 (1.f - h.z * h.z) * dz[r][2], (1.f - h.w * h.w) * dz[r][3]);
          *reinterpret_cast<float4*>(sB2 + (row0 + r) * LDH + col0) = d;
        }
      }
    }
    __syncthreads();
    if (HP == 64 && MODE == M_GRAD && a.cache_mode == 5 && a.cache) {
      // h1 | h2 of the tile in the plane order of the tcgen05 product (fvp_tc.cuh: 96-timestep tiles, 32-byte atoms)
      for (int i = tid; i < TS * 8; i += NTHREADS) {
        const int tr = i >> 3, qa = i & 7;
        const long long row = t0 + tr;
        if (row >= (a.N + tc::ROWS - 1) / tc::ROWS * tc::ROWS) continue;   // past the last 96-timestep tile
        const long long t96 = row / tc::ROWS;
        const int r96 = (int)(row - t96 * tc::ROWS);
        float* dst = a.cache + (size_t)t96 * (2 * tc::PLANE / 4) + tc::mn_atom(r96, qa) / 4;
        const float4* s1 = reinterpret_cast<const float4*>(sH1 + tr * LDH + 8 * qa);
        const float4* s2 = reinterpret_cast<const float4*>(sH2 + tr * LDH + 8 * qa);
        reinterpret_cast<float4*>(dst)[0] = s1[0]; reinterpret_cast<float4*>(dst)[1] = s1[1];
        reinterpret_cast<float4*>(dst + tc::PLANE / 4)[0] = s2[0]; reinterpret_cast<float4*>(dst + tc::PLANE / 4)[1] = s2[1];
      }
    }
    PHASE_STAMP(2);
    // ---- P3a: head.  4 lanes split K for one timestep (lane q takes k = q, q+4, ...: the
    // activation reads of a warp then cover all 32 banks and the four weight rows of one step
    // sit in four different bank groups), shuffle-reduce ------------------------------------
    {
      const int q = tid & 3;
      constexpr int KQ = HP / 4;
      for (int t = tid >> 2; t < TS; t += NTHREADS / 4) {
        const float* h2 = sH2 + t * LDH + q;
        const float* d2 = sB2 + t * LDH + q;
        for (int jt = 0; jt < MP; jt += 4) {
          float s[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
          for (int k = 0; k < KQ; ++k) {
            const float hv = h2[4 * k];
            const float4 wv = *reinterpret_cast<const float4*>(sm + L.W2t + (q + 4 * k) * LDM + jt);
            if (MODE == M_FVP) {
              const float dv = d2[4 * k];
              const float4 vv = *reinterpret_cast<const float4*>(sm + L.V2t + (q + 4 * k) * LDM + jt);
              s[0] = fmaf(hv, vv.x, s[0]); s[1] = fmaf(hv, vv.y, s[1]);
              s[2] = fmaf(hv, vv.z, s[2]); s[3] = fmaf(hv, vv.w, s[3]);
              s[0] = fmaf(dv, wv.x, s[0]); s[1] = fmaf(dv, wv.y, s[1]);
              s[2] = fmaf(dv, wv.z, s[2]); s[3] = fmaf(dv, wv.w, s[3]);
            } else {
              s[0] = fmaf(hv, wv.x, s[0]); s[1] = fmaf(hv, wv.y, s[1]);
              s[2] = fmaf(hv, wv.z, s[2]); s[3] = fmaf(hv, wv.w, s[3]);
            }
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            s[c] += __shfl_xor_sync(0xffffffffu, s[c], 1);
            s[c] += __shfl_xor_sync(0xffffffffu, s[c], 2);
          }
          if (q == 0) {
            float o[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const int j = jt + c;
              if (MODE == M_FVP) {
                // u = fisher_j * (dy_j + vb2_j); rows past N contribute nothing
                o[c] = (t < rows) ? sm[L.fisher + j] * (s[c] + sm[L.vb2 + j]) : 0.f;
              } else {
                o[c] = (s[c] + sm[L.b2 + j]) * sm[L.oscale + j] + sm[L.oshift + j];  // mu
              }
            }
            *reinterpret_cast<float4*>(sU + t * LDM + jt) = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
      }
    }
    __syncthreads();
    PHASE_STAMP(3);
    // ---- P3b: Gaussian head, one thread per timestep ----------------------------------------
    if (MODE != M_FVP) {
      for (int t = tid; t < TS; t += NTHREADS) {
        const bool valid = t < rows;
        const long long gt = t0 + t;
        float ss = 0.f, kl = 0.f, sum_ls = 0.f;
        const bool need_ll = (MODE != M_FWD) || a.ll_out != nullptr;
        for (int j = 0; j < m; ++j) {
          const float mu = sU[t * LDM + j];
          if (MODE == M_FWD && a.mean_out && valid) a.mean_out[gt * m + j] = mu;
          if (MODE == M_GRAD && a.mean_out && valid) a.mean_out[gt * m + j] = mu;
          if (need_ll) {
            const float z = (sAct[t * LDM + j] - mu) / sm[L.sig + j];
            ss = fmaf(z, z, ss);
            sum_ls += sm[L.ls + j];
            if (MODE == M_GRAD) sE[t * LDM + j] = z;  // stash z, finished below once LR is known
          }
          if (MODE == M_EVAL && valid) {
            // gaussian_mlp.py:137-139, KL(old || new) per dimension
            const float mo = ld_stream(a.old_mean + gt * m + j);
            const float so = expf(sm[L.ols_ + j]), sn = sm[L.sig + j];
            const float nr = (mo - mu) * (mo - mu) + so * so - sn * sn;
            const float dr = 2.f * sn * sn + 1e-8f;
            kl += nr / dr + sm[L.ls + j] - sm[L.ols_ + j];
          }
        }
        // gaussian_mlp.py:107-109: (-0.5*sum z^2 + -sum log_std) + -0.5*m*log(2 pi)
        const float ll = (-0.5f * ss + -sum_ls) + (float)(-0.5 * (double)m * 1.8378770664093453);
        if ((MODE == M_FWD || MODE == M_GRAD) && a.ll_out && valid) a.ll_out[gt] = ll;
        if (MODE == M_EVAL || MODE == M_GRAD) {
          float lr = 1.f;
          if (a.old_ll && valid) lr = expf(ll - ld_stream(a.old_ll + gt));
          const float wv = (a.w && valid) ? ld_stream(a.w + gt) : 0.f;
          const float c = valid ? lr * wv : 0.f;
          sS0[t] = c;
          if (MODE == M_EVAL) sS1[t] = valid ? kl : 0.f;
          if (MODE == M_GRAD) {
            for (int j = 0; j < m; ++j) {
              const float z = sE[t * LDM + j];
              const float sg = sm[L.sig + j];
              sU[t * LDM + j] = c * (z / sg) * sm[L.oscale + j];  // dL/dy_j
              sE[t * LDM + j] = c * (z * z - 1.f);                 // dL/dlog_std_j
            }
            for (int j = m; j < MP; ++j) { sU[t * LDM + j] = 0.f; sE[t * LDM + j] = 0.f; }
          }
        }
      }
      __syncthreads();
    PHASE_STAMP(4);
      if (MODE == M_EVAL || MODE == M_GRAD) {
        if (tid < 32) {  // fixed-order fp64 sums of the tile's per-timestep terms
          double s0 = 0.0, s1 = 0.0;
          for (int t = tid; t < TS; t += 32) { s0 += (double)sS0[t]; if (MODE == M_EVAL) s1 += (double)sS1[t]; }
          s0 = warp_sum(s0);
          if (MODE == M_EVAL) s1 = warp_sum(s1);
          surr_acc += s0; kl_acc += s1;
        }
      }
    }
    if (BWD) {
      // ---- P4: back through the head: gz2 = (U W2) * (1 - h2^2); gW2 += U^T H2 ---------------
      {
        float g[RT][4];
#pragma unroll
        for (int r = 0; r < RT; ++r) { g[r][0] = g[r][1] = g[r][2] = g[r][3] = 0.f; }
        mm_acc<RT>(g, sU + row0 * LDM, LDM, sm + L.W2 + col0, LDH, MP);
#pragma unroll
        for (int r = 0; r < RT; ++r) {
          const float4 h = *reinterpret_cast<const float4*>(sH2 + (row0 + r) * LDH + col0);
          float4 o = make_float4(g[r][0] * (1.f - h.x * h.x), g[r][1] * (1.f - h.y * h.y),
                                 g[r][2] * (1.f - h.z * h.z), g[r][3] * (1.f - h.w * h.w));
          *reinterpret_cast<float4*>(sB2 + (row0 + r) * LDH + col0) = o;
          cb1[0] += o.x; cb1[1] += o.y; cb1[2] += o.z; cb1[3] += o.w;
        }
        if (own2.active)
          wgrad_acc(acc2, sU, LDM, sH2, LDH, 4 * (own2.tile / NF), 4 * (own2.tile % NF), own2.grp,
                    own2.G, TS);
        // fc2.bias / log_std gradients: column sums over the tile, one warp per action dimension
        // (lanes over timesteps, fixed-order shuffle reduction)
#pragma unroll
        for (int ji = 0; ji < 4; ++ji) {
          const int j = ji * 8 + warp;
          if (j < MP) {
            float sacc = 0.f, eacc = 0.f;
            for (int t = lane; t < TS; t += 32) {
              sacc += sU[t * LDM + j];
              if (MODE == M_GRAD) eacc += sE[t * LDM + j];
            }
            sacc = warp_sum(sacc);
            if (MODE == M_GRAD) eacc = warp_sum(eacc);
            gb2acc[ji] += sacc; glsacc[ji] += eacc;
          }
        }
      }
      __syncthreads();
    PHASE_STAMP(5);
      // ---- P5: gz1 = (GZ2 W1) * (1 - h1^2); gW1 += GZ2^T H1 -----------------------------------
      if constexpr (MMA) {
        float gh[2][2][4], unused[2][2][4];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
          for (int ni = 0; ni < 2; ++ni) gh[mi][ni][0] = gh[mi][ni][1] = gh[mi][ni][2] = gh[mi][ni][3] = 0.f;
        mma_layer<false, false>(gh, unused, sB2 + mrow0 * LDH, nullptr, LDH, sm + L.W1 + ncol0, nullptr, LDW,
                                HP, fg, fc);
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
          for (int ni = 0; ni < 2; ++ni)
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const int row = mrow0 + mi * 16 + fg + half * 8, col = ncol0 + ni * 8 + 2 * fc;
              const float2 h = *reinterpret_cast<const float2*>(sH1 + row * LDH + col);
              const float o0 = gh[mi][ni][2 * half] * (1.f - h.x * h.x);
              const float o1 = gh[mi][ni][2 * half + 1] * (1.f - h.y * h.y);
              *reinterpret_cast<float2*>(sB1 + row * LDH + col) = make_float2(o0, o1);
              mcb0[ni * 2] += o0; mcb0[ni * 2 + 1] += o1;
            }
        {
          const int f = warp * NW1;  // NW1 consecutive fragments share one 16-row band
          mma_wgrad<NW1>(macc1, sB2, LDH, sH1, LDH, (f / (HP / 8)) * 16, (f % (HP / 8)) * 8, NW1, TS, fg, fc);
        }
      } else {
        float g[RT][4];
#pragma unroll
        for (int r = 0; r < RT; ++r) { g[r][0] = g[r][1] = g[r][2] = g[r][3] = 0.f; }
        mm_acc<RT>(g, sB2 + row0 * LDH, LDH, sm + L.W1 + col0, LDW, HP);
#pragma unroll
        for (int r = 0; r < RT; ++r) {
          const float4 h = *reinterpret_cast<const float4*>(sH1 + (row0 + r) * LDH + col0);
          float4 o = make_float4(g[r][0] * (1.f - h.x * h.x), g[r][1] * (1.f - h.y * h.y),
                                 g[r][2] * (1.f - h.z * h.z), g[r][3] * (1.f - h.w * h.w));
          *reinterpret_cast<float4*>(sB1 + (row0 + r) * LDH + col0) = o;
          cb0[0] += o.x; cb0[1] += o.y; cb0[2] += o.z; cb0[3] += o.w;
        }
        if (own1.active)
          wgrad_acc(acc1, sB2, LDH, sH1, LDH, 4 * (own1.tile / NF), 4 * (own1.tile % NF), own1.grp,
                    own1.G, TS);
      }
      __syncthreads();
    PHASE_STAMP(6);
      // ---- P6: gW0 += GZ1^T X -----------------------------------------------------------------
      if constexpr (MMA) {
        const int nfn = NP / 8, nf0 = (HP / 16) * nfn;
#pragma unroll
        for (int s = 0; s < W0SLOTS; ++s) {
          const int f = warp + 8 * s;
          if (f < nf0) {
            float (&acc)[1][4] = *reinterpret_cast<float (*)[1][4]>(&macc0[s]);
            mma_wgrad<1>(acc, sB1, LDH, sX, LDX, (f / nfn) * 16, (f % nfn) * 8, 1, TS, fg, fc);
          }
        }
      } else {
#pragma unroll
        for (int s = 0; s < W0SLOTS; ++s) {
          const TileOwner own0(nt0, tid, s);
          if (own0.active)
            wgrad_acc(acc0[s], sB1, LDH, sX, LDX, 4 * (own0.tile / (NP / 4)), 4 * (own0.tile % (NP / 4)),
                      own0.grp, own0.G, TS);
        }
      }
    }
  }

  // ---- per-CTA results ------------------------------------------------------------------------
  if (a.phase_cycles && tid == 0)
    for (int q = 0; q < 8; ++q) a.phase_cycles[blockIdx.x * 8 + q] = ph[q];
  if ((MODE == M_EVAL || MODE == M_GRAD) && tid == 0) {
    a.spartials[2 * blockIdx.x + 0] = surr_acc;
    a.spartials[2 * blockIdx.x + 1] = kl_acc;
  }
  if (BWD) {
    // Reduce thread-group partials in a FIXED order into a flat [d] staging vector (reusing the
    // activation buffers), then one coalesced store of the CTA's partial gradient.
    __syncthreads();
    float* sG = sH1;  // H1,H2,B1,B2 are contiguous: 4*TS*LDH floats >= d (checked on host)
    for (int i = tid; i < F.d; i += NTHREADS) sG[i] = 0.f;
    __syncthreads();
    const int h1 = F.h1, h2 = F.h2;
    int gmax = NT;
    if constexpr (MMA) {
      gmax = max(gmax, max(own2.G, WM * 8));
      // tensor-core fragments: every element has exactly one owner thread
#pragma unroll
      for (int j = 0; j < NW1; ++j) {
        const int f = warp * NW1 + j, o0 = (f / (HP / 8)) * 16 + fg, i0 = (f % (HP / 8)) * 8 + 2 * fc;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int o = o0 + (e >> 1) * 8, i = i0 + (e & 1);
          if (o < h2 && i < h1) sG[F.oW1 + o * h1 + i] = macc1[j][e];
        }
      }
      const int nfn = NP / 8, nf0 = (HP / 16) * nfn;
#pragma unroll
      for (int sl = 0; sl < W0SLOTS; ++sl) {
        const int f = warp + 8 * sl;
        if (f < nf0) {
          const int o0 = (f / nfn) * 16 + fg, i0 = (f % nfn) * 8 + 2 * fc;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int o = o0 + (e >> 1) * 8, i = i0 + (e & 1);
            if (o < h1 && i < n) sG[F.oW0 + o * n + i] = macc0[sl][e];
          }
        }
      }
      __syncthreads();
    } else {
      const TileOwner o0(nt0, tid, 0);
      gmax = max(gmax, max(o0.G, max(own1.G, own2.G)));
    }
    for (int g = 0; g < gmax; ++g) {
      if constexpr (MMA) {
        // fc0 bias gradient: column sums held per (warp row band, fragment row group)
        if (wm * 8 + fg == g) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int f = ncol0 + (q >> 1) * 8 + 2 * fc + (q & 1);
            if (f < h1) sG[F.ob0 + f] += mcb0[q];
          }
        }
      }
#pragma unroll
      for (int s = 0; s < (MMA ? 0 : W0SLOTS); ++s) {
        const TileOwner own0(nt0, tid, s);
        if (own0.active && own0.grp == g) {
          const int o0 = 4 * (own0.tile / (NP / 4)), i0 = 4 * (own0.tile % (NP / 4));
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int o = o0 + (e >> 2), i = i0 + (e & 3);
            if (o < h1 && i < n) sG[F.oW0 + o * n + i] += acc0[s][e];
          }
        }
      }
      if (!MMA && own1.active && own1.grp == g) {
        const int o0 = 4 * (own1.tile / NF), i0 = 4 * (own1.tile % NF);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int o = o0 + (e >> 2), i = i0 + (e & 3);
          if (o < h2 && i < h1) sG[F.oW1 + o * h1 + i] += acc1[e];
        }
      }
      if (own2.active && own2.grp == g) {
        const int j0 = 4 * (own2.tile / NF), f0 = 4 * (own2.tile % NF);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int j = j0 + (e >> 2), f = f0 + (e & 3);
          if (j < m && f < h2) sG[F.oW2 + j * h2 + f] += acc2[e];
        }
      }
      if (tt == g) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int f = col0 + c;
          if (!MMA && f < h1) sG[F.ob0 + f] += cb0[c];
          if (f < h2) sG[F.ob1 + f] += cb1[c];
        }
      }
      __syncthreads();
    }
    if (lane == 0) {
#pragma unroll
      for (int ji = 0; ji < 4; ++ji) {
        const int j = ji * 8 + warp;
        if (j < m) {
          sG[F.ob2 + j] = gb2acc[ji];
          sG[F.ols + j] = glsacc[ji];  // FVP: 0 here, the log_std block is filled by the finalize kernel
        }
      }
    }
    __syncthreads();
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    for (int i = tid; i < F.d; i += NTHREADS) out[i] = sG[i];
  }
}

}  // namespace pol

#include "fvp_rows.cuh"
#include "pol_rows.cuh"
#include "philox.cuh"

// =========================================================================================
// Second-stage reductions (fixed order => bit-reproducible run to run)
// =========================================================================================
namespace pol {

// out[i] = sum_c partials[c][i] in fp64.  FVP: the log_std block [ols, d) is
// count * h(sigma_j) * vec_j with h = 4 s (2 s - eps)/(2 s + eps)^2, s = sigma^2, eps = 1e-8
// (second derivative of the KL expression of gaussian_mlp.py:137-139 at new == old).
__global__ void __launch_bounds__(256)
reduce_partials_kernel(const float* __restrict__ partials, int nparts, int dstride,
                       int d, double* __restrict__ out, int ols, double count,
                       const float* __restrict__ params, const float* __restrict__ vec,
                       const int32_t* __restrict__ stop_flag) {
  // block = 32 outputs x 8 slices of the partials; slices are combined in a fixed order
  __shared__ double sacc[8][33];
  if (stop_flag && *stop_flag) return;
  const int x = threadIdx.x & 31, y = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + x;
  double acc = 0.0;
  const bool net = i < d && !(vec && i >= ols);
  if (net)
    for (int c = y; c < nparts; c += 8) acc += (double)partials[(size_t)c * dstride + i];
  sacc[y][x] = acc;
  __syncthreads();
  if (y != 0 || i >= d) return;
  if (vec && i >= ols) {
    const double ls = (double)params[i];
    const double s = exp(2.0 * ls), eps = 1e-8;
    const double h = 4.0 * s * (2.0 * s - eps) / ((2.0 * s + eps) * (2.0 * s + eps));
    out[i] = count * h * (double)vec[i];
    return;
  }
  double tot = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) tot += sacc[k][x];
  out[i] = tot;
}

__global__ void reduce_scalars_kernel(const double* __restrict__ sp, int nparts, double* __restrict__ out,
                                      int which /*bit0: surr, bit1: kl*/, double* __restrict__ out1) {
  double s0 = 0.0, s1 = 0.0;
  for (int c = threadIdx.x; c < nparts; c += 32) { s0 += sp[2 * c]; s1 += sp[2 * c + 1]; }
  s0 = warp_sum(s0); s1 = warp_sum(s1);
  if (threadIdx.x == 0) {
    if (which & 1) out[0] = s0;
    if (which & 2) (out1 ? out1 : out + 1)[0] = s1;
  }
}

__global__ void clamp_log_std_kernel(float* p, int ols, int m, float lo) {
  int j = threadIdx.x;
  if (j < m) p[ols + j] = fmaxf(p[ols + j], lo);
}

template <int HP, int MODE, int W0SLOTS, bool MMA>
int launch_one(mjrl_ctx* ctx, const Args& a, int grid, size_t smem, cudaStream_t s) {
  auto kern = mlp_kernel<HP, MODE, W0SLOTS, MMA>;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         ctx->max_smem_optin);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr_set = true;
  }
  static const bool dbg = getenv("MJRL_B200_PHASE_CYCLES") != nullptr;
  if (dbg) {
    Args b = a;
    long long* dev; cudaMalloc(&dev, sizeof(long long) * 8 * grid);
    cudaMemset(dev, 0, sizeof(long long) * 8 * grid);
    b.phase_cycles = dev;
    kern<<<grid, NTHREADS, smem, s>>>(b);
    cudaStreamSynchronize(s);
    long long* h = new long long[8 * grid];
    cudaMemcpy(h, dev, sizeof(long long) * 8 * grid, cudaMemcpyDeviceToHost);
    double tot[8] = {0};
    for (int c = 0; c < grid; ++c) for (int q = 0; q < 8; ++q) tot[q] += (double)h[c * 8 + q] / grid;
    fprintf(stderr, "[phase cycles HP=%d mode=%d mma=%d] P0 %.0f P1 %.0f P2 %.0f P3a %.0f P3b %.0f P4 %.0f P5 %.0f P6 %.0f\n",
            HP, MODE, (int)MMA, tot[0], tot[1], tot[2], tot[3], tot[4], tot[5], tot[6], tot[7]);
    delete[] h; cudaFree(dev);
    return mjrl_check_launch("mlp_kernel(dbg)");
  }
  kern<<<grid, NTHREADS, smem, s>>>(a);
  return mjrl_check_launch("mlp_kernel");
}

// Kernel variant: tensor cores (3xTF32 mma.sync) unless MJRL_B200_POLICY_PATH=fma or the shape does
// not fit the tensor-core shared-memory plan (then the fp32-FMA variant, same results to ~1e-6).
inline bool want_mma() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("MJRL_B200_POLICY_PATH");
    v = (e && e[0] == 'f') ? 0 : 1;
  }
  return v == 1;
}

template <int HP, int MODE, bool MMA>
int launch_slots(mjrl_ctx* ctx, const Args& a, int slots, int grid, size_t smem, cudaStream_t s) {
  constexpr bool BWD = MODE == M_GRAD || MODE == M_FVP;
  if constexpr (!BWD) {
    return launch_one<HP, MODE, 1, MMA>(ctx, a, grid, smem, s);
  } else if constexpr (MMA) {
    if (slots <= 1) return launch_one<HP, MODE, 1, true>(ctx, a, grid, smem, s);
    if (slots <= 2) return launch_one<HP, MODE, 2, true>(ctx, a, grid, smem, s);
    if (slots <= 4) return launch_one<HP, MODE, 4, true>(ctx, a, grid, smem, s);
    return launch_one<HP, MODE, 8, true>(ctx, a, grid, smem, s);
  } else {
    return slots <= 1 ? launch_one<HP, MODE, 1, false>(ctx, a, grid, smem, s)
                      : launch_one<HP, MODE, 2, false>(ctx, a, grid, smem, s);
  }
}

// FVP variant selection: "rows in registers" (fvp_rows.cuh) when the shape fits it, unless
// MJRL_B200_FVP=tile asks for the shared-memory tile kernel.
inline bool want_rows() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("MJRL_B200_FVP");
    v = (e && e[0] == 't') ? 0 : 1;
  }
  return v == 1;
}

template <int HP, int NPK, bool USE_CACHE>
int launch_rows_one(mjrl_ctx* ctx, const Args& a, int grid, size_t smem, cudaStream_t s) {
  auto kern = rows::fvp_rows_kernel<HP, NPK, USE_CACHE>;
  static size_t attr = 0;   // the kernel also has a few static bytes: ask for what is needed, not the maximum
  if (smem > attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr = smem;
  }
  kern<<<grid, NTHREADS, smem, s>>>(a);
  return mjrl_check_launch("fvp_rows_kernel");
}

// does the rows variant take this shape?  (HP, NPK, shared memory)
inline bool rows_fits(const mjrl_ctx* ctx, int n, int m, int h1, int h2, int* HP_out, int* NPK_out, size_t* smem_out) {
  const int hmax = h1 > h2 ? h1 : h2;
  if (!want_mma() || !want_rows() || h1 <= 0 || h2 <= 0 || hmax > 64 || m > 8 || n + 1 > 24) return false;
  const int HP = hmax <= 32 ? 32 : 64, NPK = (n + 1 + 7) / 8;
  const Flat F(n, m, h1, h2);
  const rows::RLayout L(HP, NPK, n);
  const size_t smem = sizeof(float) * (size_t)L.total;
  if (smem + 64 > (size_t)ctx->max_smem_optin || F.d > 2 * HP * rows::LDT) return false;
  *HP_out = HP; *NPK_out = NPK; *smem_out = smem;
  return true;
}

// returns -1 when the shape does not fit this variant
inline int launch_fvp_rows(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s) {
  int HP, NPK;
  size_t smem;
  if (!rows_fits(ctx, a.n, a.m, a.h1, a.h2, &HP, &NPK, &smem)) return -1;
  const Flat F(a.n, a.m, a.h1, a.h2);
  const long long ntiles = (a.N + 2 * rows::TSR - 1) / (2 * rows::TSR);   // a CTA walks two tiles at a time
  int grid = (int)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  a.dstride = (F.d + 31) & ~31;
  a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
  if (!a.partials) return MJRL_E_STATE;
  const bool use = a.cache && a.cache_mode == 2;
  if (a.cache_mode != 1 && !use) a.cache = nullptr;
#define ROWS_CASE(H, K)                                                                     \
  if (HP == H && NPK == K)                                                                  \
    return use ? launch_rows_one<H, K, true>(ctx, a, grid, smem, s)                         \
               : launch_rows_one<H, K, false>(ctx, a, grid, smem, s)
  ROWS_CASE(64, 1); ROWS_CASE(64, 2); ROWS_CASE(64, 3);
  ROWS_CASE(32, 1); ROWS_CASE(32, 2); ROWS_CASE(32, 3);
#undef ROWS_CASE
  return -1;
}

template <int HP, int NPK, int MODE>
int launch_polrows_one(mjrl_ctx* ctx, const Args& a, int grid, size_t smem, cudaStream_t s) {
  auto kern = rows::pol_rows_kernel<HP, NPK, MODE>;
  static size_t attr = 0;
  if (smem > attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr = smem;
  }
  kern<<<grid, NTHREADS, smem, s>>>(a);
  return mjrl_check_launch("pol_rows_kernel");
}

// gradient / evaluation in the rows form; returns -1 when the shape does not fit
template <int MODE>
int launch_pol_rows(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s) {
  int HP, NPK;
  size_t smem_fvp;
  if (!rows_fits(ctx, a.n, a.m, a.h1, a.h2, &HP, &NPK, &smem_fvp)) return -1;
  const Flat F(a.n, a.m, a.h1, a.h2);
  const rows::GLayout L(HP, NPK, a.n, MODE == M_GRAD);
  const size_t smem = sizeof(float) * (size_t)L.total;
  if (smem + 128 > (size_t)ctx->max_smem_optin) return -1;
  const long long ntiles = (a.N + 2 * rows::TSR - 1) / (2 * rows::TSR);
  int grid = (int)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  a.dstride = (F.d + 31) & ~31;
  if (MODE == M_GRAD) {
    a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
    if (!a.partials) return MJRL_E_STATE;
  }
  a.spartials = (double*)mjrl_scratch2(ctx, sizeof(double) * 2 * grid);
  if (!a.spartials) return MJRL_E_STATE;
#define POLROWS_CASE(H, K) if (HP == H && NPK == K) return launch_polrows_one<H, K, MODE>(ctx, a, grid, smem, s)
  POLROWS_CASE(64, 1); POLROWS_CASE(64, 2); POLROWS_CASE(64, 3);
  POLROWS_CASE(32, 1); POLROWS_CASE(32, 2); POLROWS_CASE(32, 3);
#undef POLROWS_CASE
  return -1;
}

template <int MODE>
int launch_mode(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s) {
  const int hmax = a.h1 > a.h2 ? a.h1 : a.h2;
  MJRL_REQUIRE(a.h1 > 0 && a.h2 > 0, MJRL_E_SHAPE, "policy: MLP kernels need two hidden layers");
  MJRL_REQUIRE(hmax <= 64, MJRL_E_SHAPE, "policy: hidden sizes (%d,%d) > 64 are not supported yet",
               a.h1, a.h2);
  if constexpr (MODE == M_FVP) {
    const int rc = launch_fvp_rows(ctx, a, grid_out, s);
    if (rc >= 0) return rc;
  }
  if constexpr (MODE == M_GRAD || MODE == M_EVAL) {
    const int rc = launch_pol_rows<MODE>(ctx, a, grid_out, s);
    if (rc >= 0) return rc;
  }
  if (a.cache_mode != 5) a.cache = nullptr;   // the tile kernels below neither fill nor use the fragment-ordered cache
  const int HP = hmax <= 32 ? 32 : 64;
  const Flat F(a.n, a.m, a.h1, a.h2);
  const bool bwd = MODE == M_GRAD || MODE == M_FVP;
  bool mma = want_mma();
  int slots = 1;
  if (mma) {
    const Layout Lm(HP, a.n, a.m, MODE, true);
    const int nf0 = (HP / 16) * (Lm.NP / 8);
    slots = bwd ? (nf0 + 7) / 8 : 1;
    if (sizeof(float) * (size_t)Lm.total > (size_t)ctx->max_smem_optin || slots > 8) mma = false;
  }
  MJRL_REQUIRE(a.m <= 32, MJRL_E_SHAPE, "policy: act_dim %d > 32 is not supported by the MLP kernels", a.m);
  // double-buffer the obs/act rows when the staging area fits next to everything else
  a.prefetch = sizeof(float) * (size_t)Layout(HP, a.n, a.m, MODE, mma, true).total <= (size_t)ctx->max_smem_optin;
  const Layout L(HP, a.n, a.m, MODE, mma, a.prefetch != 0);
  const size_t smem = sizeof(float) * (size_t)L.total;
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE,
               "policy: obs_dim=%d act_dim=%d hidden=%d needs %zu B of shared memory (max %d)", a.n,
               a.m, HP, smem, ctx->max_smem_optin);
  if (bwd)
    MJRL_REQUIRE(F.d <= 4 * L.TS * L.LDH, MJRL_E_SHAPE, "policy: d=%d exceeds the staging area", F.d);
  if (!mma) {
    const int nt0 = (HP / 4) * (L.NP / 4);
    MJRL_REQUIRE(nt0 <= 2 * NTHREADS, MJRL_E_SHAPE, "policy: obs_dim=%d too large (fc0 tiles %d)", a.n, nt0);
    slots = (bwd && nt0 > NTHREADS) ? 2 : 1;
  }
  const long long ntiles = (a.N + L.TS - 1) / L.TS;
  int grid = (int)(ntiles < ctx->sm_count ? ntiles : ctx->sm_count);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  a.dstride = (F.d + 31) & ~31;
  if (bwd) {
    a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
    if (!a.partials) return MJRL_E_STATE;
  }
  if (MODE == M_EVAL || MODE == M_GRAD) {
    a.spartials = (double*)mjrl_scratch2(ctx, sizeof(double) * 2 * grid);
    if (!a.spartials) return MJRL_E_STATE;
  }
  if (HP == 32)
    return mma ? launch_slots<32, MODE, true>(ctx, a, slots, grid, smem, s)
               : launch_slots<32, MODE, false>(ctx, a, slots, grid, smem, s);
  return mma ? launch_slots<64, MODE, true>(ctx, a, slots, grid, smem, s)
             : launch_slots<64, MODE, false>(ctx, a, slots, grid, smem, s);
}

inline void fill_desc(Args& a, const mjrl_policy_desc* d) {
  a.n = d->n; a.m = d->m; a.h1 = d->h1; a.h2 = d->h2;
  a.in_shift = d->in_shift; a.in_scale = d->in_scale;
  a.out_shift = d->out_shift; a.out_scale = d->out_scale;
}

}  // namespace pol

// =========================================================================================
// Affine policy (LinearModel.forward, gaussian_linear.py:171-175): mu = (W0 x~ + b0)*os + osh.
// The zero-hidden-layer case of the same passes; d = m*n + 2m is tiny, so one thread per
// timestep for the head and one thread per weight entry for the outer-product sums.
// =========================================================================================
namespace pol {

constexpr int LIN_TS = 256;

template <int MODE>
__global__ void __launch_bounds__(LIN_TS) linpol_kernel(const Args a) {
  extern __shared__ __align__(16) float sm[];
  if (MODE == M_FVP && a.stop_flag && *a.stop_flag) return;
  const int n = a.n, m = a.m, tid = threadIdx.x;
  const int ns = n | 1, ms = m | 1;
  const Flat F(n, m, 0, 0);
  float* sX = sm;                      // [TS][ns]
  float* sU = sX + LIN_TS * ns;        // [TS][ms]  dL/dy
  float* sE = sU + LIN_TS * ms;        // [TS][ms]  dL/dlog_std terms
  float* sW = sE + LIN_TS * ms;        // [m][n]
  float* sV = sW + m * n;              // [m][n]  (FVP)
  float* sG = sV + m * n;              // [d] accumulators
  float* sp = sG + F.d;                // per-dim constants: b, vb, os, osh, sig, ls, ols, fisher (8*m)
  float* sish = sp + 8 * m;            // [n]
  float* siden = sish + n;             // [n]
  __shared__ double sred[32];
  const float* P = a.params;
  for (int i = tid; i < m * n; i += LIN_TS) { sW[i] = P[i]; sV[i] = MODE == M_FVP ? a.vec[i] : 0.f; }
  for (int i = tid; i < F.d; i += LIN_TS) sG[i] = 0.f;
  for (int j = tid; j < m; j += LIN_TS) {
    const float ls = P[F.ols + j], sg = expf(ls);
    const float os = a.out_scale ? a.out_scale[j] : 1.f;
    sp[j] = P[F.ob0 + j];
    sp[m + j] = MODE == M_FVP ? a.vec[F.ob0 + j] : 0.f;
    sp[2 * m + j] = os;
    sp[3 * m + j] = a.out_shift ? a.out_shift[j] : 0.f;
    sp[4 * m + j] = sg;
    sp[5 * m + j] = ls;
    sp[6 * m + j] = MODE == M_EVAL ? a.old_log_std[j] : 0.f;
    sp[7 * m + j] = os * os * (2.f / (2.f * sg * sg + 1e-8f));
  }
  for (int i = tid; i < n; i += LIN_TS) {
    sish[i] = a.in_shift ? a.in_shift[i] : 0.f;
    siden[i] = (a.in_scale ? a.in_scale[i] : 1.f) + 1e-8f;
  }
  double surr_acc = 0.0, kl_acc = 0.0;
  const long long ntiles = (a.N + LIN_TS - 1) / LIN_TS;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long t0 = tile * LIN_TS;
    const int rows = (int)min((long long)LIN_TS, a.N - t0);
    __syncthreads();
    const float* src = a.obs + t0 * n;
    for (int i = tid; i < rows * n; i += LIN_TS) {
      int r = i / n, c = i - r * n;
      sX[r * ns + c] = (ld_stream(src + i) - sish[c]) / siden[c];
    }
    __syncthreads();
    const int t = tid;
    const bool valid = t < rows;
    const long long gt = t0 + t;
    float surr_t = 0.f, kl_t = 0.f;
    if (valid) {
      float ss = 0.f, sum_ls = 0.f;
      const float* x = sX + t * ns;
      for (int j = 0; j < m; ++j) {
        float y = sp[j], dy = sp[m + j];
        for (int k = 0; k < n; ++k) {
          y = fmaf(sW[j * n + k], x[k], y);
          if (MODE == M_FVP) dy = fmaf(sV[j * n + k], x[k], dy);
        }
        if (MODE == M_FVP) { sU[t * ms + j] = sp[7 * m + j] * dy; continue; }
        const float mu = y * sp[2 * m + j] + sp[3 * m + j];
        if ((MODE == M_FWD || MODE == M_GRAD) && a.mean_out) a.mean_out[gt * m + j] = mu;
        if (a.act) {
          const float z = (ld_stream(a.act + gt * m + j) - mu) / sp[4 * m + j];
          ss = fmaf(z, z, ss);
          sum_ls += sp[5 * m + j];
          if (MODE == M_GRAD) sE[t * ms + j] = z;
        }
        if (MODE == M_EVAL) {
          const float mo = ld_stream(a.old_mean + gt * m + j);
          const float so = expf(sp[6 * m + j]), sn = sp[4 * m + j];
          kl_t += ((mo - mu) * (mo - mu) + so * so - sn * sn) / (2.f * sn * sn + 1e-8f) + sp[5 * m + j] - sp[6 * m + j];
        }
      }
      if (MODE != M_FVP) {
        const float ll = (-0.5f * ss + -sum_ls) + (float)(-0.5 * (double)m * 1.8378770664093453);
        if ((MODE == M_FWD || MODE == M_GRAD) && a.ll_out) a.ll_out[gt] = ll;
        if (MODE == M_EVAL || MODE == M_GRAD) {
          const float lr = a.old_ll ? expf(ll - ld_stream(a.old_ll + gt)) : 1.f;
          const float c = lr * (a.w ? ld_stream(a.w + gt) : 0.f);
          surr_t = c;
          if (MODE == M_GRAD)
            for (int j = 0; j < m; ++j) {
              const float z = sE[t * ms + j];
              sU[t * ms + j] = c * (z / sp[4 * m + j]) * sp[2 * m + j];
              sE[t * ms + j] = c * (z * z - 1.f);
            }
        }
      }
    } else if (MODE == M_GRAD || MODE == M_FVP) {
      for (int j = 0; j < m; ++j) { sU[t * ms + j] = 0.f; sE[t * ms + j] = 0.f; }
    }
    if (MODE == M_EVAL || MODE == M_GRAD) {
      double s0 = block_sum((double)surr_t, sred);
      double s1 = MODE == M_EVAL ? block_sum((double)kl_t, sred) : 0.0;
      if (tid == 0) { surr_acc += s0; kl_acc += s1; }
    }
    if (MODE == M_GRAD || MODE == M_FVP) {
      __syncthreads();
      for (int e = tid; e < m * n + 2 * m; e += LIN_TS) {
        float s = 0.f;
        if (e < m * n) {
          const int j = e / n, k = e - j * n;
          for (int r = 0; r < rows; ++r) s = fmaf(sU[r * ms + j], sX[r * ns + k], s);
          sG[e] += s;
        } else if (e < m * n + m) {
          const int j = e - m * n;
          for (int r = 0; r < rows; ++r) s += sU[r * ms + j];
          sG[F.ob0 + j] += s;
        } else if (MODE == M_GRAD) {
          const int j = e - m * n - m;
          for (int r = 0; r < rows; ++r) s += sE[r * ms + j];
          sG[F.ols + j] += s;
        }
      }
    }
  }
  if ((MODE == M_EVAL || MODE == M_GRAD) && tid == 0) {
    a.spartials[2 * blockIdx.x] = surr_acc;
    a.spartials[2 * blockIdx.x + 1] = kl_acc;
  }
  if (MODE == M_GRAD || MODE == M_FVP) {
    __syncthreads();
    float* out = a.partials + (size_t)blockIdx.x * a.dstride;
    for (int i = tid; i < F.d; i += LIN_TS) out[i] = sG[i];
  }
}

template <int MODE>
int launch_linear(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s) {
  const Flat F(a.n, a.m, 0, 0);
  const int ns = a.n | 1, ms = a.m | 1;
  size_t fl = (size_t)LIN_TS * ns + 2 * (size_t)LIN_TS * ms + 2 * (size_t)a.m * a.n + F.d + 8 * a.m +
              2 * a.n + 8;
  size_t smem = sizeof(float) * fl + sizeof(double) * 40;
  MJRL_REQUIRE(smem + 1024 <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE, "linear policy: n=%d m=%d too large", a.n, a.m);
  const long long ntiles = (a.N + LIN_TS - 1) / LIN_TS;
  int grid = (int)(ntiles < ctx->sm_count * 2 ? ntiles : ctx->sm_count * 2);
  if (grid < 1) grid = 1;
  *grid_out = grid;
  a.dstride = (F.d + 31) & ~31;
  if (MODE == M_GRAD || MODE == M_FVP) {
    a.partials = (float*)mjrl_scratch(ctx, sizeof(float) * (size_t)grid * a.dstride);
    if (!a.partials) return MJRL_E_STATE;
  }
  if (MODE == M_EVAL || MODE == M_GRAD) {
    a.spartials = (double*)mjrl_scratch2(ctx, sizeof(double) * 2 * grid);
    if (!a.spartials) return MJRL_E_STATE;
  }
  auto kern = linpol_kernel<MODE>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
  }
  kern<<<grid, LIN_TS, smem, s>>>(a);
  return mjrl_check_launch("linpol_kernel");
}

template <int MODE>
int launch_any(mjrl_ctx* ctx, Args& a, int* grid_out, cudaStream_t s) {
  if (a.h1 == 0 && a.h2 == 0) return launch_linear<MODE>(ctx, a, grid_out, s);
  return launch_mode<MODE>(ctx, a, grid_out, s);
}

}  // namespace pol

// =========================================================================================
// (d) CG vector algebra, step-size rule;  (e) batched action sampling
// =========================================================================================
namespace pol {

constexpr int CG_THREADS = 1024;

// block-wide fp64 dot helper for the single-CTA CG kernels
__device__ __forceinline__ double cta_sum_bcast(double v, double* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  double r = (l < nw) ? red[l] : 0.0;
  r = warp_sum(r);
  return r;  // every thread holds the total
}

__global__ void __launch_bounds__(CG_THREADS)
cg_init_kernel(mjrl_cg_state st, long long d, const double* __restrict__ b, double b_scale) {
  __shared__ double red[32];
  double acc = 0.0;
  for (long long i = threadIdx.x; i < d; i += CG_THREADS) {
    const double v = b[i] * b_scale;
    st.x[i] = 0.0; st.r[i] = v; st.p[i] = v; st.p32[i] = (float)v;
    acc = fma(v, v, acc);
  }
  const double rr = cta_sum_bcast(acc, red);
  if (threadIdx.x == 0) {
    st.scal[0] = rr; st.scal[1] = 0.0; st.scal[2] = 0.0; st.scal[3] = 0.0; st.scal[4] = rr;
    st.flags[0] = 0; st.flags[1] = 0;
  }
}

// cg_solve.py:10-20 with z = z_sum*inv_count + damping*p (npg_cg.py:74).
// REGS: every thread keeps its <= CG_E elements of p, z, x, r in registers, so the four vectors are
// read once (all loads in flight together) instead of once per pass; same arithmetic and order.
constexpr int CG_E = 8;
template <bool REGS>
__global__ void __launch_bounds__(CG_THREADS)
cg_update_kernel(mjrl_cg_state st, long long d, double inv_count, double damping, double tol) {
  __shared__ double red[32];
  if (st.flags[0]) return;
  const double rdotr = st.scal[0];
  double pz = 0.0;
  double rp[REGS ? CG_E : 1], rz[REGS ? CG_E : 1], rx[REGS ? CG_E : 1], rr[REGS ? CG_E : 1];
  if (REGS) {
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      const long long i = threadIdx.x + (long long)k * CG_THREADS;
      const bool in = i < d;
      rp[k] = in ? st.p[i] : 0.0; rz[k] = in ? st.z[i] : 0.0;
      rx[k] = in ? st.x[i] : 0.0; rr[k] = in ? st.r[i] : 0.0;
    }
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      rz[k] = fma(rz[k], inv_count, damping * rp[k]);
      pz = fma(rp[k], rz[k], pz);
    }
  } else {
    for (long long i = threadIdx.x; i < d; i += CG_THREADS) {
      const double p = st.p[i];
      const double z = fma(st.z[i], inv_count, damping * p);
      st.z[i] = z;
      pz = fma(p, z, pz);
    }
  }
  pz = cta_sum_bcast(pz, red);
  const double v = rdotr / pz;
  double nrr = 0.0;
  if (REGS) {
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      rx[k] = fma(v, rp[k], rx[k]);
      rr[k] = fma(-v, rz[k], rr[k]);
      nrr = fma(rr[k], rr[k], nrr);
    }
  } else {
    for (long long i = threadIdx.x; i < d; i += CG_THREADS) {
      st.x[i] = fma(v, st.p[i], st.x[i]);
      const double r = fma(-v, st.z[i], st.r[i]);
      st.r[i] = r;
      nrr = fma(r, r, nrr);
    }
  }
  nrr = cta_sum_bcast(nrr, red);
  const double mu = nrr / rdotr;
  if (REGS) {
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      const long long i = threadIdx.x + (long long)k * CG_THREADS;
      if (i < d) {
        const double p = fma(mu, rp[k], rr[k]);
        st.z[i] = rz[k]; st.x[i] = rx[k]; st.r[i] = rr[k];
        st.p[i] = p; st.p32[i] = (float)p;
      }
    }
  } else {
    for (long long i = threadIdx.x; i < d; i += CG_THREADS) {
      const double p = fma(mu, st.p[i], st.r[i]);
      st.p[i] = p; st.p32[i] = (float)p;
    }
  }
  if (threadIdx.x == 0) {
    st.scal[0] = nrr; st.scal[1] = pz; st.scal[2] = v; st.scal[3] = mu;
    st.flags[1] += 1;
    if (nrr < tol) st.flags[0] = 1;
  }
}

__global__ void __launch_bounds__(CG_THREADS)
npg_step_kernel(long long d, int m, const double* __restrict__ g, double g_scale,
                const double* __restrict__ x, const float* __restrict__ curr, double delta,
                double const_alpha, double alpha_scale, float min_log_std,
                float* __restrict__ newp, double* __restrict__ scal) {
  __shared__ double red[32];
  double gx = 0.0;
  for (long long i = threadIdx.x; i < d; i += CG_THREADS) gx = fma(g[i] * g_scale, x[i], gx);
  gx = cta_sum_bcast(gx, red);
  double alpha, nstep;
  if (const_alpha >= 0.0) { alpha = const_alpha; nstep = alpha * alpha * gx; }        // npg_cg.py:130-132
  else { nstep = delta; alpha = sqrt(fabs(delta / (gx + 1e-20))); }                     // npg_cg.py:133-135
  const double a = alpha * alpha_scale;
  for (long long i = threadIdx.x; i < d; i += CG_THREADS) {
    float v = (float)((double)curr[i] + a * x[i]);
    if (i >= d - m) v = fmaxf(v, min_log_std);  // set_param_values clamp, gaussian_mlp.py:75-76
    newp[i] = v;
  }
  if (threadIdx.x == 0) { scal[0] = alpha; scal[1] = gx; scal[2] = nstep; scal[3] = a; }
}

// ---- Philox4x32-10 + Box-Muller ------------------------------------------------------------
// act[b][j] = mean[b][j] + exp(log_std_j) * eps, counter = (row + offset, j/4), key = seed
__global__ void act_noise_kernel(const float* __restrict__ mean, const float* __restrict__ params, int ols,
                                 int m, long long B, unsigned long long seed, unsigned long long offset,
                                 float* __restrict__ act) {
  const long long nblk = (m + 3) / 4;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = B * nblk, stride = (long long)gridDim.x * blockDim.x;
  for (; i < total; i += stride) {
    const long long b = i / nblk;
    const int jb = (int)(i - b * nblk);
    const unsigned long long row = (unsigned long long)b + offset;
    uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), (uint32_t)jb, 0u};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    float e[4];
    philox_normal4(c, e);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int j = jb * 4 + q;
      if (j < m) act[b * m + j] = mean[b * m + j] + expf(params[ols + j]) * e[q];
    }
  }
}

}  // namespace pol

// =========================================================================================
// C-ABI entry points (include/mjrl_b200.h)
// =========================================================================================
using namespace pol;

static int check_desc(const mjrl_policy_desc* d, const char* who) {
  MJRL_REQUIRE(d != nullptr, MJRL_E_ARG, "%s: null policy description", who);
  MJRL_REQUIRE(d->n > 0 && d->m > 0, MJRL_E_ARG, "%s: bad obs/act dims %d/%d", who, d->n, d->m);
  MJRL_REQUIRE((d->h1 > 0 && d->h2 > 0) || (d->h1 == 0 && d->h2 == 0), MJRL_E_SHAPE,
               "%s: hidden sizes (%d,%d): exactly two hidden layers, or none", who, d->h1, d->h2);
  return 0;
}

extern "C" {

int64_t mjrl_policy_param_count(const mjrl_policy_desc* d) {
  if (!d || d->n <= 0 || d->m <= 0 || d->h1 < 0 || d->h2 < 0) return -1;
  return Flat(d->n, d->m, d->h1, d->h2).d;
}

int mjrl_policy_clamp_log_std(mjrl_ctx* ctx, const mjrl_policy_desc* desc, float* params,
                              float min_log_std, void* stream) {
  int rc = check_desc(desc, "clamp_log_std");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params, MJRL_E_ARG, "clamp_log_std: null pointer");
  MJRL_REQUIRE(desc->m <= 1024, MJRL_E_SHAPE, "clamp_log_std: act_dim %d > 1024", desc->m);
  const Flat F(desc->n, desc->m, desc->h1, desc->h2);
  clamp_log_std_kernel<<<1, ((desc->m + 31) / 32) * 32, 0, (cudaStream_t)stream>>>(params, F.ols, desc->m, min_log_std);
  return mjrl_check_launch("clamp_log_std");
}

int mjrl_policy_forward(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                        const float* obs, const float* act, int64_t N, float* mean_out,
                        float* ll_out, void* stream) {
  int rc = check_desc(desc, "policy_forward");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && N >= 0, MJRL_E_ARG, "policy_forward: bad argument");
  if (N == 0) return 0;
  MJRL_REQUIRE(obs && (mean_out || ll_out), MJRL_E_ARG, "policy_forward: null pointer");
  MJRL_REQUIRE(!ll_out || act, MJRL_E_ARG, "policy_forward: ll_out needs act");
  Args a = {};
  fill_desc(a, desc);
  a.params = params; a.obs = obs; a.act = ll_out ? act : nullptr; a.N = N;
  a.mean_out = mean_out; a.ll_out = ll_out;
  int grid;
  return launch_any<M_FWD>(ctx, a, &grid, (cudaStream_t)stream);
}

int mjrl_policy_eval(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                     const float* obs, const float* act, const float* w, const float* old_ll,
                     const float* old_mean, const float* old_log_std, int64_t N, double* sums2,
                     void* stream) {
  int rc = check_desc(desc, "policy_eval");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && obs && act && old_ll && old_mean && old_log_std && sums2 && N > 0,
               MJRL_E_ARG, "policy_eval: bad argument");
  Args a = {};
  fill_desc(a, desc);
  a.params = params; a.obs = obs; a.act = act; a.w = w; a.old_ll = old_ll; a.old_mean = old_mean;
  a.old_log_std = old_log_std; a.N = N;
  int grid;
  cudaStream_t s = (cudaStream_t)stream;
  rc = launch_any<M_EVAL>(ctx, a, &grid, s);
  if (rc) return rc;
  reduce_scalars_kernel<<<1, 32, 0, s>>>(a.spartials, grid, sums2, 3, nullptr);
  return mjrl_check_launch("policy_eval reduce");
}

int mjrl_policy_eval_tc(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                        const float* obs, const float* act, const float* w, const float* old_ll,
                        const float* old_mean, const float* old_log_std, int64_t N, double* sums2,
                        float* fvp_cache, void* stream) {
  int rc = check_desc(desc, "policy_eval_tc");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && obs && act && old_ll && old_mean && old_log_std && sums2 && N > 0,
               MJRL_E_ARG, "policy_eval_tc: bad argument");
  const int kind = mjrl_policy_fvp_tc_ok(ctx, desc);
  MJRL_REQUIRE(kind == 1 || (kind == 2 && fvp_cache), MJRL_E_SHAPE,
               "policy_eval_tc: this policy shape has no tcgen05 pass (or the wide form was given no cache)");
  Args a = {};
  fill_desc(a, desc);
  a.params = params; a.obs = obs; a.act = act; a.w = w; a.old_ll = old_ll; a.old_mean = old_mean;
  a.old_log_std = old_log_std; a.N = N;
  int grid;
  cudaStream_t s = (cudaStream_t)stream;
  TcwRegions r = {};
  if (kind == 2) r = tcw_regions(fvp_cache, desc->n, N);
  const int64_t nt = (N + 127) / 128;
  grid = (int)(nt < ctx->sm_count ? nt : ctx->sm_count);
  rc = launch_pol_fwd(ctx, a, kind == 2 ? &r : nullptr, grid, s);
  if (rc) return rc;
  reduce_scalars_kernel<<<1, 32, 0, s>>>(a.spartials, grid, sums2, 3, nullptr);
  return mjrl_check_launch("policy_eval_tc reduce");
}

int mjrl_policy_grad(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                     const float* obs, const float* act, const float* w, const float* old_ll,
                     int64_t N, double* grad_sum, double* surr_sum, float* ll_out, float* mean_out,
                     void* stream) {
  return mjrl_policy_grad_cached(ctx, desc, params, obs, act, w, old_ll, N, grad_sum, surr_sum, ll_out,
                                 mean_out, nullptr, stream);
}

int mjrl_policy_grad_cached(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                            const float* obs, const float* act, const float* w, const float* old_ll,
                            int64_t N, double* grad_sum, double* surr_sum, float* ll_out,
                            float* mean_out, float* fvp_cache, void* stream) {
  int rc = check_desc(desc, "policy_grad");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && obs && act && w && grad_sum && N > 0, MJRL_E_ARG,
               "policy_grad: bad argument");
  Args a = {};
  fill_desc(a, desc);
  a.params = params; a.obs = obs; a.act = act; a.w = w; a.old_ll = old_ll; a.N = N;
  a.ll_out = ll_out; a.mean_out = mean_out;
  a.cache = fvp_cache; a.cache_mode = fvp_cache ? 1 : 0;
  int grid;
  cudaStream_t s = (cudaStream_t)stream;
  rc = launch_any<M_GRAD>(ctx, a, &grid, s);
  if (rc) return rc;
  const Flat F(desc->n, desc->m, desc->h1, desc->h2);
  reduce_partials_kernel<<<(F.d + 31) / 32, 256, 0, s>>>(a.partials, grid, a.dstride, F.d, grad_sum,
                                                            F.ols, 0.0, nullptr, nullptr, nullptr);
  if (surr_sum) reduce_scalars_kernel<<<1, 32, 0, s>>>(a.spartials, grid, surr_sum, 1, nullptr);
  return mjrl_check_launch("policy_grad reduce");
}

static int64_t tc_region_offset(mjrl_ctx* ctx, const mjrl_policy_desc* desc, int64_t N);

// gradient / evaluation passes on tcgen05 (pol_tc.cu) unless MJRL_B200_GRAD_TC=0 (then the mma.sync kernels fill the planes)
static bool grad_on_tc() {
  static const bool off = getenv("MJRL_B200_GRAD_TC") && getenv("MJRL_B200_GRAD_TC")[0] == '0';
  return !off;
}

int mjrl_policy_grad_cached_tc(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                               const float* obs, const float* act, const float* w, const float* old_ll,
                               int64_t N, double* grad_sum, double* surr_sum, float* ll_out,
                               float* mean_out, float* fvp_cache, void* stream) {
  int rc = check_desc(desc, "policy_grad_cached_tc");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && obs && act && grad_sum && fvp_cache && N > 0, MJRL_E_ARG,
               "policy_grad_cached_tc: bad argument");
  MJRL_REQUIRE(mjrl_policy_fvp_tc_ok(ctx, desc), MJRL_E_SHAPE,
               "policy_grad_cached_tc: this policy shape has no tcgen05 product");
  cudaStream_t s = (cudaStream_t)stream;
  if (mjrl_policy_fvp_tc_ok(ctx, desc) == 2) {
    // wide observation / action vectors (fvp_tc_wide.cu): the shared-memory tile kernel writes the planes, one more
    // pass builds the x~ image both first-layer passes of every product read
    Args a = {};
    fill_desc(a, desc);
    a.params = params; a.obs = obs; a.act = act; a.w = w; a.old_ll = old_ll; a.N = N;
    a.ll_out = ll_out; a.mean_out = mean_out;
    a.cache = fvp_cache; a.cache_mode = 5;
    {
      const int64_t ntc = (N + 95) / 96, covered = ((N + 63) / 64) * 64;   // the tile kernel walks 64 timesteps at a time
      if (covered < ntc * 96) {
        cudaError_t e = cudaMemsetAsync(fvp_cache + (ntc - 1) * (2 * 96 * 64), 0, sizeof(float) * 2 * 96 * 64, s);
        if (e != cudaSuccess) { mjrl_set_error("policy_grad_cached_tc: memset: %s", cudaGetErrorString(e)); return (int)e; }
      }
    }
    rc = launch_tcw_ximg(ctx, a, fvp_cache, s);
    if (rc) return rc;
    int grid;
    if (grad_on_tc()) {
      const TcwRegions r = tcw_regions(fvp_cache, desc->n, N);
      a.cache = nullptr; a.cache_mode = 0;
      rc = launch_pol_tc<M_GRAD>(ctx, a, r.h12, &r, &grid, s);
    } else {
      rc = launch_mode<M_GRAD>(ctx, a, &grid, s);
    }
    if (rc) return rc;
    const Flat F(desc->n, desc->m, desc->h1, desc->h2);
    reduce_partials_kernel<<<(F.d + 31) / 32, 256, 0, s>>>(a.partials, grid, a.dstride, F.d, grad_sum,
                                                              F.ols, 0.0, nullptr, nullptr, nullptr);
    if (surr_sum) reduce_scalars_kernel<<<1, 32, 0, s>>>(a.spartials, grid, surr_sum, 1, nullptr);
    return mjrl_check_launch("policy_grad_cached_tc reduce");
  }
  const int64_t off = tc_region_offset(ctx, desc, N);
  MJRL_REQUIRE(off >= 0, MJRL_E_SHAPE, "policy_grad_cached_tc: shape not taken");
  // rows of the last 96-timestep tile that no 64-timestep tile of the gradient pass writes: must be finite
  // (they are operands of the weight-gradient MMAs, multiplied by zero rows)
  {
    const int64_t ntc = (N + 95) / 96, covered = ((N + rows::TSR - 1) / rows::TSR) * rows::TSR;
    if (covered < ntc * 96) {
      cudaError_t e = cudaMemsetAsync(fvp_cache + off + (ntc - 1) * (2 * 96 * 64), 0, sizeof(float) * 2 * 96 * 64, s);
      if (e != cudaSuccess) { mjrl_set_error("policy_grad_cached_tc: memset: %s", cudaGetErrorString(e)); return (int)e; }
    }
  }
  Args a = {};
  fill_desc(a, desc);
  a.params = params; a.obs = obs; a.act = act; a.w = w; a.old_ll = old_ll; a.N = N;
  a.ll_out = ll_out; a.mean_out = mean_out;
  int grid;
  if (grad_on_tc()) {
    rc = launch_pol_tc<M_GRAD>(ctx, a, fvp_cache + off, nullptr, &grid, s);     // pol_tc.cu: fills the planes itself
  } else {
    a.cache = fvp_cache + off; a.cache_mode = 5;
    rc = launch_pol_rows<M_GRAD>(ctx, a, &grid, s);
    MJRL_REQUIRE(rc >= 0, MJRL_E_SHAPE, "policy_grad_cached_tc: rows kernel does not take this shape");
  }
  if (rc) return rc;
  const Flat F(desc->n, desc->m, desc->h1, desc->h2);
  reduce_partials_kernel<<<(F.d + 31) / 32, 256, 0, s>>>(a.partials, grid, a.dstride, F.d, grad_sum,
                                                            F.ols, 0.0, nullptr, nullptr, nullptr);
  if (surr_sum) reduce_scalars_kernel<<<1, 32, 0, s>>>(a.spartials, grid, surr_sum, 1, nullptr);
  return mjrl_check_launch("policy_grad_cached_tc reduce");
}

int64_t mjrl_policy_fvp_cache_count(mjrl_ctx* ctx, const mjrl_policy_desc* desc, int64_t N) {
  if (!ctx || !desc || N <= 0) return 0;
  int HP, NPK;
  size_t smem;
  if (!rows_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, &HP, &NPK, &smem))
    return tcw_fits(ctx, desc->n, desc->m, desc->h1, desc->h2) ? tcw_cache_floats(desc->n, N) : 0;
  const int64_t ntiles = (N + rows::TSR - 1) / rows::TSR;
  int64_t cnt = ntiles * rows::TSR * 2 * HP;   // h1 and h2 of every (padded) timestep
  if (jac_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, nullptr, nullptr, nullptr))
    cnt += jac_cache_floats(desc->m, desc->h1, desc->h2, N);   // + the m Jacobian rows G1 (fvp_jac.cu)
  if (tc_fits(ctx, desc->n, desc->m, desc->h1, desc->h2))
    cnt += tc_cache_floats(N);                                  // + h1|h2 in the tile order of fvp_tc.cu
  return cnt;
}

// float offset of the tcgen05 kernel's tile-ordered region inside the cache buffer (after h1|h2 and G1)
static int64_t tc_region_offset(mjrl_ctx* ctx, const mjrl_policy_desc* desc, int64_t N) {
  int HP, NPK;
  size_t smem;
  if (!rows_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, &HP, &NPK, &smem)) return -1;
  const int64_t ntiles = (N + rows::TSR - 1) / rows::TSR;
  int64_t off = ntiles * rows::TSR * 2 * HP;
  if (jac_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, nullptr, nullptr, nullptr))
    off += jac_cache_floats(desc->m, desc->h1, desc->h2, N);
  return off;
}

int mjrl_policy_fvp_tc_ok(mjrl_ctx* ctx, const mjrl_policy_desc* desc) {
  if (!ctx || !desc) return 0;
  int HP, NPK;
  size_t smem;
  if (rows_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, &HP, &NPK, &smem))
    return HP == 64 && tc_fits(ctx, desc->n, desc->m, desc->h1, desc->h2) ? 1 : 0;
  // shapes the rows-in-registers kernels do not take: the three-pass form, fed by mjrl_policy_grad_cached_tc only
  return tcw_fits(ctx, desc->n, desc->m, desc->h1, desc->h2) ? 2 : 0;
}

int mjrl_policy_fvp_tc_fill(mjrl_ctx* ctx, const mjrl_policy_desc* desc, int64_t N, float* cache, void* stream) {
  int rc = check_desc(desc, "policy_fvp_tc_fill");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && cache && N > 0, MJRL_E_ARG, "policy_fvp_tc_fill: bad argument");
  MJRL_REQUIRE(mjrl_policy_fvp_tc_ok(ctx, desc) == 1, MJRL_E_SHAPE,
               "policy_fvp_tc_fill: this policy shape has no fragment-ordered cache to re-order");
  const int64_t off = tc_region_offset(ctx, desc, N);
  MJRL_REQUIRE(off >= 0, MJRL_E_SHAPE, "policy_fvp_tc_fill: shape not taken");
  return launch_tc_fill(ctx, cache, cache + off, N, (cudaStream_t)stream);
}

int mjrl_policy_fvp_jac_ok(mjrl_ctx* ctx, const mjrl_policy_desc* desc) {
  if (!ctx || !desc) return 0;
  int HP, NPK;
  size_t smem;
  return rows_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, &HP, &NPK, &smem) &&
         jac_fits(ctx, desc->n, desc->m, desc->h1, desc->h2, nullptr, nullptr, nullptr) ? 1 : 0;
}

int mjrl_policy_fvp_jac_fill(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params, int64_t N,
                             float* cache, void* stream) {
  int rc = check_desc(desc, "policy_fvp_jac_fill");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && cache && N > 0, MJRL_E_ARG, "policy_fvp_jac_fill: bad argument");
  MJRL_REQUIRE(mjrl_policy_fvp_jac_ok(ctx, desc), MJRL_E_SHAPE,
               "policy_fvp_jac_fill: this policy shape has no Jacobian-cached product");
  rc = launch_jac_fill(ctx, params, desc->n, desc->m, desc->h1, desc->h2, cache, N, (cudaStream_t)stream);
  MJRL_REQUIRE(rc >= 0, MJRL_E_SHAPE, "policy_fvp_jac_fill: shape not taken");
  return rc;
}

int mjrl_policy_fvp(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                    const float* obs, int64_t N, const float* vec, double* fvp_sum,
                    const int32_t* stop_flag, void* stream) {
  return mjrl_policy_fvp_cached(ctx, desc, params, obs, N, vec, fvp_sum, stop_flag, nullptr, 0, stream);
}

int mjrl_policy_fvp_cached(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                           const float* obs, int64_t N, const float* vec, double* fvp_sum,
                           const int32_t* stop_flag, float* cache, int32_t cache_mode, void* stream) {
  int rc = check_desc(desc, "policy_fvp");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && obs && vec && fvp_sum && N > 0, MJRL_E_ARG, "policy_fvp: bad argument");
  MJRL_REQUIRE(cache_mode >= 0 && cache_mode <= 4 && (cache_mode == 0 || cache), MJRL_E_ARG,
               "policy_fvp: cache_mode %d needs a cache buffer", cache_mode);
  Args a = {};
  fill_desc(a, desc);
  a.params = params; a.obs = obs; a.vec = vec; a.N = N; a.stop_flag = stop_flag;
  a.cache = cache; a.cache_mode = cache_mode;
  int grid;
  cudaStream_t s = (cudaStream_t)stream;
  if (cache_mode == 3) {
    MJRL_REQUIRE(mjrl_policy_fvp_jac_ok(ctx, desc), MJRL_E_SHAPE,
                 "policy_fvp: cache_mode 3 needs a shape mjrl_policy_fvp_jac_ok accepts");
    rc = launch_fvp_jac(ctx, a, &grid, s);
    MJRL_REQUIRE(rc >= 0, MJRL_E_SHAPE, "policy_fvp: Jacobian-cached product does not take this shape");
  } else if (cache_mode == 4) {
    MJRL_REQUIRE(mjrl_policy_fvp_tc_ok(ctx, desc), MJRL_E_SHAPE,
                 "policy_fvp: cache_mode 4 needs a shape mjrl_policy_fvp_tc_ok accepts");
    if (mjrl_policy_fvp_tc_ok(ctx, desc) == 2) {
      rc = launch_fvp_tcw(ctx, a, cache, &grid, s);
    } else {
      const int64_t off = tc_region_offset(ctx, desc, N);
      rc = launch_fvp_tc(ctx, a, off >= 0 ? cache + off : nullptr, &grid, s);
    }
    MJRL_REQUIRE(rc >= 0, MJRL_E_SHAPE, "policy_fvp: tcgen05 product does not take this shape");
  } else {
    rc = launch_any<M_FVP>(ctx, a, &grid, s);
  }
  if (rc) return rc;
  const Flat F(desc->n, desc->m, desc->h1, desc->h2);
  reduce_partials_kernel<<<(F.d + 31) / 32, 256, 0, s>>>(a.partials, grid, a.dstride, F.d, fvp_sum,
                                                            F.ols, (double)N, params, vec, stop_flag);
  return mjrl_check_launch("policy_fvp reduce");
}

int mjrl_cg_init(mjrl_ctx* ctx, const mjrl_cg_state* st, int64_t d, const double* b, double b_scale,
                 void* stream) {
  MJRL_REQUIRE(ctx && st && b && d > 0, MJRL_E_ARG, "cg_init: bad argument");
  MJRL_REQUIRE(st->x && st->r && st->p && st->z && st->scal && st->p32 && st->flags, MJRL_E_ARG,
               "cg_init: null state array");
  cg_init_kernel<<<1, CG_THREADS, 0, (cudaStream_t)stream>>>(*st, d, b, b_scale);
  return mjrl_check_launch("cg_init");
}

int mjrl_cg_update(mjrl_ctx* ctx, const mjrl_cg_state* st, int64_t d, double inv_count, double damping,
                   double residual_tol, void* stream) {
  MJRL_REQUIRE(ctx && st && d > 0, MJRL_E_ARG, "cg_update: bad argument");
  if (d <= (int64_t)CG_E * CG_THREADS)
    cg_update_kernel<true><<<1, CG_THREADS, 0, (cudaStream_t)stream>>>(*st, d, inv_count, damping, residual_tol);
  else
    cg_update_kernel<false><<<1, CG_THREADS, 0, (cudaStream_t)stream>>>(*st, d, inv_count, damping, residual_tol);
  return mjrl_check_launch("cg_update");
}

int mjrl_npg_step(mjrl_ctx* ctx, int64_t d, int32_t m, const double* g, double g_scale,
                  const double* x, const float* curr, double delta, double const_alpha,
                  double alpha_scale, float min_log_std, float* new_params, double* scal4,
                  void* stream) {
  MJRL_REQUIRE(ctx && g && x && curr && new_params && scal4 && d > 0 && m > 0 && m <= d, MJRL_E_ARG,
               "npg_step: bad argument");
  npg_step_kernel<<<1, CG_THREADS, 0, (cudaStream_t)stream>>>(d, m, g, g_scale, x, curr, delta, const_alpha,
                                                               alpha_scale, min_log_std, new_params, scal4);
  return mjrl_check_launch("npg_step");
}

int mjrl_policy_act(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params,
                    const float* obs, int64_t B, uint64_t seed, uint64_t offset, float* act_out,
                    float* mean_out, void* stream) {
  int rc = check_desc(desc, "policy_act");
  if (rc) return rc;
  MJRL_REQUIRE(ctx && params && obs && act_out && mean_out && B > 0, MJRL_E_ARG, "policy_act: bad argument");
  rc = mjrl_policy_forward(ctx, desc, params, obs, nullptr, B, mean_out, nullptr, stream);
  if (rc) return rc;
  const Flat F(desc->n, desc->m, desc->h1, desc->h2);
  const long long total = B * ((desc->m + 3) / 4);
  int grid = (int)((total + 255) / 256);
  if (grid > ctx->sm_count * 8) grid = ctx->sm_count * 8;
  act_noise_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(mean_out, params, F.ols, desc->m, B, seed, offset, act_out);
  return mjrl_check_launch("policy_act");
}

}  // extern "C"
