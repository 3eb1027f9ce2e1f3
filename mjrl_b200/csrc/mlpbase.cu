// (b) MLPBaseline: (n+4) -> 128 -> 128 -> 1 ReLU value network, fp32.
//
// Replaces (reference = bennevans/mjrl, mjrl/baselines/mlp_baseline.py):
//   _features   :37-56   [clip(o,-10,10)/10, al, al^2, al^3, al^4], al = t_in_path/1000
//   predict     :107-115 forward pass over a path's features
//   fit         :59-105  epochs x (floor(N/bs) - 1) Adam steps on minibatches drawn from a
//                        host permutation (np.random.permutation, :84); MSE loss; L2
//                        weight_decay added to the gradient (torch.optim.Adam, :34)
//
// predict is a persistent tile kernel like the policy passes (fp32-FMA bound: 36.9 kflop per
// 4n+8 B).  fit is a SERIAL chain of dependent steps: one persistent CTA keeps
// the fp32 master weights in shared memory for the whole chain, the Adam moments in L2-resident
// global memory, prefetches the next minibatch with cp.async while the current step computes,
// and applies Adam directly from the register tiles that hold the weight gradients (the
// gradient is never materialised).
//
// Flat parameter layout (torch `model.parameters()` order, row-major (out,in)):
//   fc_0.weight [128][F] | fc_0.bias [128] | fc_1.weight [128][128] | fc_1.bias [128] |
//   fc_2.weight [1][128] | fc_2.bias [1]          F = n + 4
#include <algorithm>
#include <stdlib.h>

#include "policy_tile.cuh"

namespace mlpb {

constexpr int H = 128;         // hidden width (mlp_baseline.py:24-28)
constexpr int LDH = H + 4;     // activation / fc_1 row stride: = 4 (mod 32) words
constexpr int MAXB = 64;       // rows of the minibatch tile

struct Offsets {
  int F, oW0, ob0, oW1, ob1, oW2, ob2, D;
  __host__ __device__ explicit Offsets(int n) {
    F = n + 4;
    oW0 = 0; ob0 = H * F; oW1 = ob0 + H; ob1 = oW1 + H * H; oW2 = ob1 + H; ob2 = oW2 + H; D = ob2 + 1;
  }
};

// Padded feature width (k extent of fc_0) and fc_0 row stride.  The stride is = 4 (mod 8) words
// so that the 8 lanes of a quarter warp reading float4s of 8 different weight rows hit 8
// different bank groups.
__host__ __device__ inline int feat_pad(int F) { return (F + 3) & ~3; }
__host__ __device__ inline int w0_stride(int F) {
  int s = feat_pad(F) + 4;
  while ((s & 7) != 4) s += 4;
  return s;
}

// One row of features (mlp_baseline.py:38-55), computed in fp64 like the reference, cast to fp32.
__device__ __forceinline__ void time_feats(int t_in_path, float (&tf)[4]) {
  const double al = (double)t_in_path / 1000.0;
  tf[0] = (float)al;
  tf[1] = (float)(al * al);
  tf[2] = (float)pow(al, 3.0);
  tf[3] = (float)pow(al, 4.0);
}
__device__ __forceinline__ float obs_feat(float o) {
  double v = (double)o;
  v = v < -10.0 ? -10.0 : (v > 10.0 ? 10.0 : v);
  return (float)(v / 10.0);
}

__global__ void features_kernel(const float* __restrict__ obs, const int32_t* __restrict__ offsets, int P,
                                long long N, int n, float* __restrict__ feat) {
  const int F = n + 4;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < N;
       t += (long long)gridDim.x * blockDim.x) {
    const int p = find_path(offsets, P, t);
    float tf[4];
    time_feats((int)(t - offsets[p]), tf);
    const float* o = obs + t * n;
    float* f = feat + t * F;
    for (int k = 0; k < n; ++k) f[k] = obs_feat(ld_stream(o + k));
    f[n] = tf[0]; f[n + 1] = tf[1]; f[n + 2] = tf[2]; f[n + 3] = tf[3];
  }
}

// out[r][c] for the thread's 4 rows x 4 INTERLEAVED output units o = lane + 32 c:
//   acc[r][c] += sum_k A[row0 + r][k] * W[lane + 32 c][k]      (W natural (out,in), stride ldw)
// k-vectorised dot products: A loads are warp broadcasts, W loads are conflict free because
// ldw = 4 (mod 8) words.
template <int R>
__device__ __forceinline__ void dot_acc(float (&acc)[R][4], const float* __restrict__ A, int lda,
                                        const float* __restrict__ W, int ldw, int K) {
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float4 w[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) w[c] = *reinterpret_cast<const float4*>(W + (32 * c) * ldw + k);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(A + r * lda + k);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        acc[r][c] = fmaf(a.x, w[c].x, acc[r][c]);
        acc[r][c] = fmaf(a.y, w[c].y, acc[r][c]);
        acc[r][c] = fmaf(a.z, w[c].z, acc[r][c]);
        acc[r][c] = fmaf(a.w, w[c].w, acc[r][c]);
      }
    }
  }
}

// Stage fc_0 / fc_1 / fc_2 from the flat vector into padded shared-memory images.
__device__ __forceinline__ void stage_weights(const float* __restrict__ P, const Offsets& O, int FP, int LD0,
                                              float* sW0, float* sb0, float* sW1, float* sb1, float* sw2,
                                              float* sb2, int nthreads) {
  for (int i = threadIdx.x; i < H * LD0; i += nthreads) {
    const int o = i / LD0, k = i - o * LD0;
    sW0[i] = k < O.F ? P[O.oW0 + o * O.F + k] : 0.f;
  }
  for (int i = threadIdx.x; i < H * LDH; i += nthreads) {
    const int o = i / LDH, k = i - o * LDH;
    sW1[i] = k < H ? P[O.oW1 + o * H + k] : 0.f;
  }
  for (int i = threadIdx.x; i < H; i += nthreads) {
    sb0[i] = P[O.ob0 + i]; sb1[i] = P[O.ob1 + i]; sw2[i] = P[O.oW2 + i];
  }
  if (threadIdx.x == 0) sb2[0] = P[O.ob2];
  (void)FP;
}

}  // namespace mlpb

// =========================================================================================
// predict: persistent CTAs of 256 threads walk tiles of 64 timesteps.  Warp w owns rows
// 8w..8w+7, lane l owns output units l, l+32, l+64, l+96 of each hidden layer.
// =========================================================================================
namespace mlpb {

constexpr int PT = 256;   // threads
constexpr int TS = 64;    // timesteps per tile
constexpr int R = 8;      // rows per thread

struct Smem {
  int W0, W1, b0, b1, w2, b2, X, H0, H1, total;
  __host__ __device__ Smem(int F, int xbufs) {
    const int FP = feat_pad(F), LD0 = w0_stride(F);
    int o = 0;
    auto take = [&](int c) { int r = o; o += (c + 3) & ~3; return r; };
    W0 = take(H * LD0); W1 = take(H * LDH); b0 = take(H); b1 = take(H); w2 = take(H); b2 = take(4);
    X = take(xbufs * TS * FP); H0 = take(TS * LDH); H1 = take(TS * LDH);
    total = o;
  }
};

// hidden layer on the thread's 8 x 4 tile: out = relu(bias + A W^T)
__device__ __forceinline__ void hidden_layer(const float* __restrict__ A, int lda, const float* __restrict__ W,
                                             int ldw, int K, const float* __restrict__ bias,
                                             float* __restrict__ out, int row0, int lane) {
  float acc[R][4];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float b = bias[lane + 32 * c];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r][c] = b;
  }
  dot_acc<R>(acc, A + row0 * lda, lda, W + lane * ldw, ldw, K);
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) out[(row0 + r) * LDH + lane + 32 * c] = fmaxf(acc[r][c], 0.f);
}

// y[b] = b2 + sum_o w2[o] h1[b][o] for the 64 rows: 4 lanes per row, units q, q+4, q+8, ...
__device__ __forceinline__ float head_row(const float* __restrict__ sH1, const float* __restrict__ sw2, int tid) {
  const int b = tid >> 2, q = tid & 3;
  const float* h = sH1 + b * LDH;
  float s = 0.f;
#pragma unroll 8
  for (int k = 0; k < H / 4; ++k) s = fmaf(h[q + 4 * k], sw2[q + 4 * k], s);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  return s;
}

__global__ void __launch_bounds__(PT, 1)
predict_kernel(const float* __restrict__ obs, const int32_t* __restrict__ offsets, int P, long long N, int n,
               const float* __restrict__ params, double* __restrict__ pred) {
  extern __shared__ __align__(16) float sm[];
  const Offsets O(n);
  const int F = O.F, FP = feat_pad(F), LD0 = w0_stride(F);
  const Smem L(F, 1);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, row0 = warp * R;
  float* sX = sm + L.X; float* sH0 = sm + L.H0; float* sH1 = sm + L.H1;
  stage_weights(params, O, FP, LD0, sm + L.W0, sm + L.b0, sm + L.W1, sm + L.b1, sm + L.w2, sm + L.b2, PT);
  for (int i = tid; i < TS * FP; i += PT) sX[i] = 0.f;
  const long long ntiles = (N + TS - 1) / TS;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long t0 = tile * TS;
    const int rows = (int)min((long long)TS, N - t0);
    __syncthreads();
    const float* src = obs + t0 * n;
    for (int i = tid; i < rows * n; i += PT) {
      const int r = i / n, c = i - r * n;
      sX[r * FP + c] = obs_feat(ld_stream(src + i));
    }
    if (tid < rows) {
      const int p = find_path(offsets, P, t0 + tid);
      float tf[4];
      time_feats((int)(t0 + tid - offsets[p]), tf);
#pragma unroll
      for (int j = 0; j < 4; ++j) sX[tid * FP + n + j] = tf[j];
    }
    __syncthreads();
    hidden_layer(sX, FP, sm + L.W0, LD0, FP, sm + L.b0, sH0, row0, lane);
    __syncthreads();
    hidden_layer(sH0, LDH, sm + L.W1, LDH, H, sm + L.b1, sH1, row0, lane);
    __syncthreads();
    const float y = head_row(sH1, sm + L.w2, tid) + sm[L.b2];
    if ((tid & 3) == 0 && (tid >> 2) < rows) pred[t0 + (tid >> 2)] = (double)y;
  }
}

// sums2 = { sum (y32 - p32)^2 , sum y32^2 } with y32 = (float)returns, p32 = (float)pred
// (mlp_baseline.py:75-81,98-104 work on fp32 arrays); fp64 accumulation, fixed order.
__global__ void __launch_bounds__(256) sqerr_partial_kernel(const double* __restrict__ pred,
                                                            const double* __restrict__ ret, long long N,
                                                            double* __restrict__ partial) {
  __shared__ double red[32];
  double se = 0.0, sy = 0.0;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < N;
       t += (long long)gridDim.x * blockDim.x) {
    const float y = (float)ret[t], p = (float)pred[t];
    const float e = y - p;
    se += (double)(e * e); sy += (double)(y * y);
  }
  se = block_sum(se, red);
  if (threadIdx.x == 0) partial[2 * blockIdx.x] = se;
  sy = block_sum(sy, red);
  if (threadIdx.x == 0) partial[2 * blockIdx.x + 1] = sy;
}
__global__ void sqerr_final_kernel(const double* __restrict__ partial, int nparts, double* __restrict__ sums2) {
  double a = 0.0, b = 0.0;
  for (int c = threadIdx.x; c < nparts; c += 32) { a += partial[2 * c]; b += partial[2 * c + 1]; }
  a = warp_sum(a); b = warp_sum(b);
  if (threadIdx.x == 0) { sums2[0] = a; sums2[1] = b; }
}

}  // namespace mlpb

// =========================================================================================
// fit: ONE persistent CTA runs the whole chain of Adam steps (mlp_baseline.py:83-96).
// Per step (minibatch of <= 64 rows, 256 threads):
//   wait(cp.async batch s) -> prefetch batch s+1 -> L0 -> L1 -> head/dy -> gw2,gb2 ->
//   dz1 (in place of h1) -> gW1 tile in registers, gb1 -> dh0 = dz1 W1 -> dz0 (in place of h0)
//   -> Adam(W1,b1,w2,b2) -> gW0,gb0 -> Adam(W0,b0)
// =========================================================================================
namespace mlpb {

struct AdamC {
  float lr_over_bc1;   // lr / (1 - beta1^t)            (torch: step_size)
  float bc2_sqrt;      // sqrt(1 - beta2^t)
  float one_m_b1, b2, one_m_b2, eps, wd;
};

// torch.optim.Adam single-tensor update (weight_decay added to the gradient; lerp for exp_avg;
// denom = sqrt(v)/sqrt(bc2) + eps; p -= step_size * m/denom).
__device__ __forceinline__ float adam(float w, float g, float& m, float& v, const AdamC& c) {
  g = fmaf(c.wd, w, g);
  m = m + c.one_m_b1 * (g - m);
  v = fmaf(c.one_m_b2 * g, g, v * c.b2);
  const float denom = sqrtf(v) / c.bc2_sqrt + c.eps;
  return w - c.lr_over_bc1 * (m / denom);
}

__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"((uint32_t)__cvta_generic_to_shared(dst_smem)),
               "l"(src) : "memory");
}

struct FitArgs {
  const float* feat;      // [N][F]
  const float* y32;       // [N]
  const int32_t* perm;    // [nsteps * bs]
  float* params;          // [D] in/out
  float* am;              // Adam exp_avg [D] in/out
  float* av;              // Adam exp_avg_sq [D] in/out
  long long nsteps, step0;
  int n, bs;
  double lr, beta1, beta2, eps, wd;
  long long* phase_cycles;   // debug (MJRL_B200_PHASE_CYCLES): 12 per-phase cycle sums of thread 0
};

__global__ void __launch_bounds__(PT, 1) fit_kernel(const FitArgs a) {
  extern __shared__ __align__(16) float sm[];
  __shared__ AdamC s_adam;
  __shared__ float s_dy[TS], s_y[2][TS];
  const Offsets O(a.n);
  const int F = O.F, FP = feat_pad(F), LD0 = w0_stride(F);
  const Smem L(F, 2);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, row0 = warp * R;
  const int bs = a.bs;
  float* sW0 = sm + L.W0; float* sW1 = sm + L.W1; float* sb0 = sm + L.b0; float* sb1 = sm + L.b1;
  float* sw2 = sm + L.w2; float* sb2 = sm + L.b2; float* sH0 = sm + L.H0; float* sH1 = sm + L.H1;
  stage_weights(a.params, O, FP, LD0, sW0, sb0, sW1, sb1, sw2, sb2, PT);
  for (int i = tid; i < 2 * TS * FP; i += PT) sm[L.X + i] = 0.f;
  if (tid < TS) { s_y[0][tid] = 0.f; s_y[1][tid] = 0.f; }
  __syncthreads();

  auto prefetch = [&](long long s) {   // minibatch s -> buffer s & 1
    float* X = sm + L.X + (int)(s & 1) * TS * FP;
    const int32_t* idx = a.perm + s * bs;
    for (int i = tid; i < bs * F; i += PT) {
      const int b = i / F, k = i - b * F;
      cp_async4(X + b * FP + k, a.feat + (long long)__ldg(idx + b) * F + k);
    }
    if (tid < bs) cp_async4(&s_y[s & 1][tid], a.y32 + __ldg(idx + tid));
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (a.nsteps > 0) prefetch(0);

  // wgrad ownership: fc_1 tile rows o0..o0+7, cols [i0, i0+4) and [64+i0, 64+i0+4) (two float4
  // columns 64 apart so the 16 lanes that differ in i0 read 256 contiguous bytes: conflict free)
  const int o0 = 8 * (tid >> 4), i0 = 4 * (tid & 15);

  long long ph[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#define FIT_STAMP(q) if (a.phase_cycles && tid == 0) { const long long now = clock64(); ph[q] += now - tprev; tprev = now; }
  for (long long s = 0; s < a.nsteps; ++s) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (tid == 0) {
      const double t = (double)(a.step0 + s + 1);
      AdamC c;
      c.lr_over_bc1 = (float)(a.lr / (1.0 - pow(a.beta1, t)));
      c.bc2_sqrt = (float)sqrt(1.0 - pow(a.beta2, t));
      c.one_m_b1 = (float)(1.0 - a.beta1); c.b2 = (float)a.beta2; c.one_m_b2 = (float)(1.0 - a.beta2);
      c.eps = (float)a.eps; c.wd = (float)a.wd;
      s_adam = c;
    }
    __syncthreads();   // batch s landed; previous step's weight updates visible
    FIT_STAMP(0);
    if (s + 1 < a.nsteps) prefetch(s + 1);
    const float* X = sm + L.X + (int)(s & 1) * TS * FP;
    const AdamC c = s_adam;
    FIT_STAMP(1);

    // ---- forward ------------------------------------------------------------------------
    hidden_layer(X, FP, sW0, LD0, FP, sb0, sH0, row0, lane);
    __syncthreads();
    FIT_STAMP(2);
    hidden_layer(sH0, LDH, sW1, LDH, H, sb1, sH1, row0, lane);
    __syncthreads();
    FIT_STAMP(3);
    {
      const float y = head_row(sH1, sw2, tid) + sb2[0];
      const int b = tid >> 2;
      // MSELoss (mean over the bs x 1 outputs): dL/dyhat = 2 (yhat - y) / bs
      if ((tid & 3) == 0) s_dy[b] = b < bs ? 2.f * (y - s_y[s & 1][b]) / (float)bs : 0.f;
    }
    __syncthreads();
    // ---- head gradients (threads 0..127: unit o = tid; thread 128: bias) ------------------
    float gw2 = 0.f, gb2 = 0.f;
    if (tid < H) {
#pragma unroll 8
      for (int b = 0; b < TS; ++b) gw2 = fmaf(s_dy[b], sH1[b * LDH + tid], gw2);
    } else if (tid == H) {
      for (int b = 0; b < TS; ++b) gb2 += s_dy[b];
    }
    __syncthreads();
    FIT_STAMP(4);
    // ---- dz1 = dy w2^T * (h1 > 0), in place of h1 --------------------------------------------
    for (int i = tid; i < TS * H; i += PT) {
      const int b = i >> 7, o = i & (H - 1);
      const float h = sH1[b * LDH + o];
      sH1[b * LDH + o] = h > 0.f ? s_dy[b] * sw2[o] : 0.f;
    }
    __syncthreads();
    FIT_STAMP(5);
    // ---- gW1 tile (8 x (4+4)) = sum_b dz1[b][o] h0[b][i]; gb1 ------------------------------------
    float g1[8][8];
#pragma unroll
    for (int x = 0; x < 8; ++x)
#pragma unroll
      for (int y = 0; y < 8; ++y) g1[x][y] = 0.f;
#pragma unroll 2
    for (int b = 0; b < TS; ++b) {
      const float4 d0 = *reinterpret_cast<const float4*>(sH1 + b * LDH + o0);
      const float4 d1 = *reinterpret_cast<const float4*>(sH1 + b * LDH + o0 + 4);
      const float4 h0 = *reinterpret_cast<const float4*>(sH0 + b * LDH + i0);
      const float4 h1 = *reinterpret_cast<const float4*>(sH0 + b * LDH + i0 + 64);
      const float dz[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
      const float hh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
      for (int x = 0; x < 8; ++x)
#pragma unroll
        for (int y = 0; y < 8; ++y) g1[x][y] = fmaf(dz[x], hh[y], g1[x][y]);
    }
    float gb1 = 0.f;
    if (tid < H) {
#pragma unroll 8
      for (int b = 0; b < TS; ++b) gb1 += sH1[b * LDH + tid];
    }
    __syncthreads();   // all reads of h0 done before dz0 overwrites it
    FIT_STAMP(6);
    // ---- dh0 = dz1 W1 (old weights), dz0 = dh0 * (h0 > 0) in place of h0 -------------------
    {
      float acc[R][4];
#pragma unroll
      for (int r = 0; r < R; ++r) acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.f;
      pol::mm_acc<R>(acc, sH1 + row0 * LDH, LDH, sW1 + 4 * lane, LDH, H);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float4* p = reinterpret_cast<float4*>(sH0 + (row0 + r) * LDH + 4 * lane);
        const float4 h = *p;
        *p = make_float4(h.x > 0.f ? acc[r][0] : 0.f, h.y > 0.f ? acc[r][1] : 0.f,
                         h.z > 0.f ? acc[r][2] : 0.f, h.w > 0.f ? acc[r][3] : 0.f);
      }
    }
    __syncthreads();   // old W1 fully consumed; dz0 complete
    FIT_STAMP(7);
    // ---- Adam: fc_1 (register tile), fc_1.bias, fc_2 ------------------------------------------
#pragma unroll
    for (int x = 0; x < 8; ++x) {
      const int gi = O.oW1 + (o0 + x) * H + i0;
      float4* pm = reinterpret_cast<float4*>(a.am + gi);
      float4* pv = reinterpret_cast<float4*>(a.av + gi);
      float* pw = sW1 + (o0 + x) * LDH + i0;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float4 m4 = pm[16 * hh], v4 = pv[16 * hh], w4 = *reinterpret_cast<float4*>(pw + 64 * hh);
        w4.x = adam(w4.x, g1[x][4 * hh + 0], m4.x, v4.x, c);
        w4.y = adam(w4.y, g1[x][4 * hh + 1], m4.y, v4.y, c);
        w4.z = adam(w4.z, g1[x][4 * hh + 2], m4.z, v4.z, c);
        w4.w = adam(w4.w, g1[x][4 * hh + 3], m4.w, v4.w, c);
        pm[16 * hh] = m4; pv[16 * hh] = v4; *reinterpret_cast<float4*>(pw + 64 * hh) = w4;
      }
    }
    if (tid < H) {
      sb1[tid] = adam(sb1[tid], gb1, a.am[O.ob1 + tid], a.av[O.ob1 + tid], c);
      sw2[tid] = adam(sw2[tid], gw2, a.am[O.oW2 + tid], a.av[O.oW2 + tid], c);
    } else if (tid == H) {
      sb2[0] = adam(sb2[0], gb2, a.am[O.ob2], a.av[O.ob2], c);
    }
    FIT_STAMP(8);
    // ---- gW0 = dz0^T X (float4 chunks of k), gb0; Adam ----------------------------------------
    const int kchunks = FP / 4;
    for (int ch = tid; ch < H * kchunks; ch += PT) {
      const int o = ch / kchunks, k0 = 4 * (ch - o * kchunks);
      float g[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int b = 0; b < TS; ++b) {
        const float d = sH0[b * LDH + o];
        const float4 x = *reinterpret_cast<const float4*>(X + b * FP + k0);
        g[0] = fmaf(d, x.x, g[0]); g[1] = fmaf(d, x.y, g[1]); g[2] = fmaf(d, x.z, g[2]); g[3] = fmaf(d, x.w, g[3]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k0 + j;
        if (k < F) {
          const int gi = O.oW0 + o * F + k;
          sW0[o * LD0 + k] = adam(sW0[o * LD0 + k], g[j], a.am[gi], a.av[gi], c);
        }
      }
    }
    if (tid < H) {
      float gb0 = 0.f;
#pragma unroll 8
      for (int b = 0; b < TS; ++b) gb0 += sH0[b * LDH + tid];
      sb0[tid] = adam(sb0[tid], gb0, a.am[O.ob0 + tid], a.av[O.ob0 + tid], c);
    }
    FIT_STAMP(9);
    // the __syncthreads at the top of the next step orders these writes before its reads
  }
  if (a.phase_cycles && tid == 0)
    for (int q = 0; q < 12; ++q) a.phase_cycles[q] = ph[q];
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  // ---- write the master weights back ---------------------------------------------------------
  for (int i = tid; i < H * F; i += PT) a.params[O.oW0 + i] = sW0[(i / F) * LD0 + (i % F)];
  for (int i = tid; i < H * H; i += PT) a.params[O.oW1 + i] = sW1[(i >> 7) * LDH + (i & (H - 1))];
  for (int i = tid; i < H; i += PT) {
    a.params[O.ob0 + i] = sb0[i]; a.params[O.ob1 + i] = sb1[i]; a.params[O.oW2 + i] = sw2[i];
  }
  if (tid == 0) a.params[O.ob2] = sb2[0];
}

__global__ void cast_y_kernel(const double* __restrict__ ret, long long N, float* __restrict__ y32) {
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < N;
       t += (long long)gridDim.x * blockDim.x)
    y32[t] = (float)ret[t];
}

}  // namespace mlpb


// =========================================================================================
// fit, cluster version: ONE thread-block cluster of 8 CTAs runs the serial Adam chain.
//
// The chain is latency bound (15,624 dependent steps at N = 1M, 3.4 Mflop each), so the step is
// split MODEL-parallel over the cluster: CTA c owns hidden units U_c = [16c, 16c+16) of both
// layers -- rows U_c of fc_0 / fc_1, entries U_c of the biases and of fc_2 -- with their Adam
// moments, all resident in its shared memory for the whole chain.  Per step three small
// exchanges go over distributed shared memory as bulk async copies (cp.async.bulk
// shared::cta -> shared::cluster) that signal an mbarrier in the RECEIVING CTA:
//   E1  all-gather  h0[:, U_c]           (4 KB to each peer)        before layer 1
//   E2  all-gather  partial head output  (256 B to each peer)       before dy
//   E3  reduce-scatter of dh0 partials   (4 KB to each peer)        before dz0
// Exchange buffers are double buffered by step parity; because every CTA needs data from every
// peer in every step, no CTA can run more than one step ahead, so no cluster barrier is needed
// inside the loop.  Sums over peers are taken in rank order, so all CTAs compute bit-identical
// dy and the result does not depend on timing.
// =========================================================================================
namespace mlpb {

constexpr int CL = 8;            // CTAs per cluster
constexpr int UC = H / CL;       // hidden units owned per CTA (16)
constexpr int SS = TS * UC + 16; // slice stride (floats) of slice-major [CL][64][16] buffers: = 16 (mod 32)
constexpr int LDU = UC + 4;      // row stride of the per-CTA [64][16] matrices

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_cta(const void* p, uint32_t rank) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p), r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_init1(uint64_t* b) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((uint32_t)__cvta_generic_to_shared(b)) : "memory");
}
__device__ __forceinline__ void mbar_arm(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"((uint32_t)__cvta_generic_to_shared(b)), "r"(bytes) : "memory");
}
// bounded wait: a protocol bug traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_parity(uint64_t* b, uint32_t parity) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(b);
  uint32_t ok = 0;
  for (long long it = 0; !ok; ++it) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (it > (1ll << 24)) __trap();
  }
}
// local shared memory -> the same buffer of CTA `rank`, completing `bytes` on its mbarrier
__device__ __forceinline__ void bulk_to_peer(void* dst_local_equiv, const void* src, uint32_t bytes,
                                             uint64_t* bar_local_equiv, uint32_t rank) {
  const uint32_t d = map_to_cta(dst_local_equiv, rank), m = map_to_cta(bar_local_equiv, rank);
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(src);
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(d), "r"(s), "r"(bytes), "r"(m) : "memory");
}

struct CSmem {   // offsets in floats
  int W0, M0, V0, W1, M1, V1, vec, X, H0, H1, P, RECV, DZ0, YP, total;
  __host__ __device__ explicit CSmem(int F) {
    const int FP = feat_pad(F), LD0 = w0_stride(F);
    int o = 0;
    auto take = [&](int c) { int r = o; o += (c + 3) & ~3; return r; };
    W0 = take(UC * LD0); M0 = take(UC * LD0); V0 = take(UC * LD0);
    W1 = take(UC * LDH); M1 = take(UC * LDH); V1 = take(UC * LDH);
    vec = take(16 * UC);           // b0,b1,w2 (+m,v each) , b2,m,v
    X = take(2 * TS * FP);
    H0 = take(2 * CL * SS);        // all-gathered h0, slice-major, double buffered
    H1 = take(TS * LDU);           // own h1 slice, then dz1
    P = take(CL * SS);             // dh0 partials for every destination slice
    RECV = take(CL * SS);          // dh0 partials received from every source
    DZ0 = take(TS * LDU);
    YP = take(CL * TS);            // head partials from every source
    total = o;
  }
};

__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(PT, 1) fit_cluster_kernel(const FitArgs a) {
  extern __shared__ __align__(16) float sm[];
  __shared__ __align__(8) uint64_t barH[2], barY[2], barR[2];
  __shared__ AdamC s_adam;
  __shared__ float s_dy[TS], s_y[2][TS];
  const Offsets O(a.n);
  const int F = O.F, FP = feat_pad(F), LD0 = w0_stride(F);
  const CSmem L(F);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int c = (int)cluster_rank();
  const int u0 = c * UC;               // first owned unit
  const int bs = a.bs;
  float* sW0 = sm + L.W0; float* sM0 = sm + L.M0; float* sV0 = sm + L.V0;
  float* sW1 = sm + L.W1; float* sM1 = sm + L.M1; float* sV1 = sm + L.V1;
  float* sb0 = sm + L.vec;            float* sb0m = sb0 + UC; float* sb0v = sb0 + 2 * UC;
  float* sb1 = sm + L.vec + 3 * UC;   float* sb1m = sb1 + UC; float* sb1v = sb1 + 2 * UC;
  float* sw2 = sm + L.vec + 6 * UC;   float* sw2m = sw2 + UC; float* sw2v = sw2 + 2 * UC;
  float* sb2 = sm + L.vec + 9 * UC;   // [0] b2, [1] m, [2] v  (replicated in every CTA)
  float* sH1 = sm + L.H1; float* sP = sm + L.P; float* sDZ0 = sm + L.DZ0;

  // ---- stage the owned rows and their Adam moments --------------------------------------------
  for (int i = tid; i < UC * LD0; i += PT) {
    const int u = i / LD0, k = i - u * LD0;
    const int gi = O.oW0 + (u0 + u) * F + k;
    const bool in = k < F;
    sW0[i] = in ? a.params[gi] : 0.f; sM0[i] = in ? a.am[gi] : 0.f; sV0[i] = in ? a.av[gi] : 0.f;
  }
  for (int i = tid; i < UC * LDH; i += PT) {
    const int u = i / LDH, k = i - u * LDH;
    const int gi = O.oW1 + (u0 + u) * H + k;
    const bool in = k < H;
    sW1[i] = in ? a.params[gi] : 0.f; sM1[i] = in ? a.am[gi] : 0.f; sV1[i] = in ? a.av[gi] : 0.f;
  }
  if (tid < UC) {
    sb0[tid] = a.params[O.ob0 + u0 + tid]; sb0m[tid] = a.am[O.ob0 + u0 + tid]; sb0v[tid] = a.av[O.ob0 + u0 + tid];
    sb1[tid] = a.params[O.ob1 + u0 + tid]; sb1m[tid] = a.am[O.ob1 + u0 + tid]; sb1v[tid] = a.av[O.ob1 + u0 + tid];
    sw2[tid] = a.params[O.oW2 + u0 + tid]; sw2m[tid] = a.am[O.oW2 + u0 + tid]; sw2v[tid] = a.av[O.oW2 + u0 + tid];
  }
  if (tid == 0) {
    sb2[0] = a.params[O.ob2]; sb2[1] = a.am[O.ob2]; sb2[2] = a.av[O.ob2];
    for (int p = 0; p < 2; ++p) { mbar_init1(&barH[p]); mbar_init1(&barY[p]); mbar_init1(&barR[p]); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < 2 * TS * FP; i += PT) sm[L.X + i] = 0.f;
  if (tid < TS) { s_y[0][tid] = 0.f; s_y[1][tid] = 0.f; }
  __syncthreads();
  cluster_sync_all();   // every CTA's barriers are initialised before any peer signals them

  // element i = tid + j PT of a minibatch is (row slot i / F, column i % F) in every step: divide once, not per step
  // (ncu: the two runtime divisions per element were 8 % of the chain's stall samples)
  int pfb[4], pfk[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { const int i = tid + j * PT; pfb[j] = i / F; pfk[j] = i - pfb[j] * F; }
  auto prefetch = [&](long long s) {   // minibatch s -> buffer s & 1 (every CTA loads the same rows)
    float* X = sm + L.X + (int)(s & 1) * TS * FP;
    const int32_t* idx = a.perm + s * bs;
    const int cnt = bs * F;
    if (cnt <= 4 * PT) {
      long long row[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) row[j] = tid + j * PT < cnt ? (long long)__ldg(idx + pfb[j]) : 0;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (tid + j * PT < cnt) cp_async4(X + pfb[j] * FP + pfk[j], a.feat + row[j] * F + pfk[j]);
    } else
    for (int i0 = tid; i0 < cnt; i0 += 4 * PT) {
      long long row[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = i0 + j * PT;
        row[j] = i < cnt ? (long long)__ldg(idx + i / F) : 0;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = i0 + j * PT;
        if (i < cnt) { const int b = i / F, k = i - b * F; cp_async4(X + b * FP + k, a.feat + row[j] * F + k); }
      }
    }
    if (tid < bs) cp_async4(&s_y[s & 1][tid], a.y32 + __ldg(idx + tid));
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (a.nsteps > 0) prefetch(0);

  const int u = tid & 15, r0 = 4 * (tid >> 4);        // (unit, 4 rows) tiling of the [64][16] products
  const int gu = tid >> 4, gi0 = 4 * (tid & 15);      // fc_1 gradient: unit gu, inputs gi0..+3 and 64+gi0..+3

  // beta^t by recurrence (thread 0): fp64 pow() every step would cost ~2k cycles of the chain
  double b1t = pow(a.beta1, (double)a.step0), b2t = pow(a.beta2, (double)a.step0);
  long long ph[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#define CL_STAMP(q) if (a.phase_cycles && tid == 0 && c == 0) { const long long now = clock64(); ph[q] += now - tprev; tprev = now; }
  for (long long s = 0; s < a.nsteps; ++s) {
    const int p = (int)(s & 1);
    const uint32_t par = (uint32_t)((s >> 1) & 1);
    float* sH0 = sm + L.H0 + p * CL * SS;
    // Only h0 needs two buffers: a peer that has my E3 data can finish step s and send its next
    // h0 slice while this CTA still reads h0 for the fc_1 gradient.  E2/E3 data of step s+1 can
    // only be sent after this CTA's own step-(s+1) h0 went out, i.e. after it left step s.
    float* sRECV = sm + L.RECV;
    float* sYP = sm + L.YP;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (tid == 0) {
      b1t *= a.beta1; b2t *= a.beta2;
      AdamC ac;
      ac.lr_over_bc1 = (float)(a.lr / (1.0 - b1t));
      ac.bc2_sqrt = (float)sqrt(1.0 - b2t);
      ac.one_m_b1 = (float)(1.0 - a.beta1); ac.b2 = (float)a.beta2; ac.one_m_b2 = (float)(1.0 - a.beta2);
      ac.eps = (float)a.eps; ac.wd = (float)a.wd;
      s_adam = ac;
      mbar_arm(&barH[p], (CL - 1) * TS * UC * 4);
      mbar_arm(&barY[p], (CL - 1) * TS * 4);
      mbar_arm(&barR[p], (CL - 1) * TS * UC * 4);
    }
    __syncthreads();
    if (s + 1 < a.nsteps) prefetch(s + 1);
    const float* X = sm + L.X + p * TS * FP;
    const AdamC ac = s_adam;
    CL_STAMP(0);

    // ---- L0: own slice of h0 -> sH0[c], then all-gather (E1) ---------------------------------
    {
      float acc[4];
      const float b = sb0[u];
      acc[0] = acc[1] = acc[2] = acc[3] = b;
      for (int k = 0; k < FP; k += 4) {
        const float4 w = *reinterpret_cast<const float4*>(sW0 + u * LD0 + k);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float4 x = *reinterpret_cast<const float4*>(X + (r0 + r) * FP + k);
          acc[r] = fmaf(x.x, w.x, acc[r]); acc[r] = fmaf(x.y, w.y, acc[r]);
          acc[r] = fmaf(x.z, w.z, acc[r]); acc[r] = fmaf(x.w, w.w, acc[r]);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) sH0[c * SS + (r0 + r) * UC + u] = fmaxf(acc[r], 0.f);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid < CL && tid != c) bulk_to_peer(sH0 + c * SS, sH0 + c * SS, TS * UC * 4, &barH[p], (uint32_t)tid);
    CL_STAMP(1);
    mbar_wait_parity(&barH[p], par);
    CL_STAMP(2);

    // ---- L1: own slice of h1; head partial -> E2 -----------------------------------------------
    {
      float acc[4];
      const float b = sb1[u];
      acc[0] = acc[1] = acc[2] = acc[3] = b;
#pragma unroll 4
      for (int k = 0; k < H; k += 4) {
        const float4 w = *reinterpret_cast<const float4*>(sW1 + u * LDH + k);
        const float* hs = sH0 + (k >> 4) * SS + (k & 15);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float4 x = *reinterpret_cast<const float4*>(hs + (r0 + r) * UC);
          acc[r] = fmaf(x.x, w.x, acc[r]); acc[r] = fmaf(x.y, w.y, acc[r]);
          acc[r] = fmaf(x.z, w.z, acc[r]); acc[r] = fmaf(x.w, w.w, acc[r]);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) sH1[(r0 + r) * LDU + u] = fmaxf(acc[r], 0.f);
    }
    __syncthreads();
    if (tid < TS) {
      float y = 0.f;
#pragma unroll
      for (int k = 0; k < UC; ++k) y = fmaf(sH1[tid * LDU + k], sw2[k], y);
      sYP[c * TS + tid] = y;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid < CL && tid != c) bulk_to_peer(sYP + c * TS, sYP + c * TS, TS * 4, &barY[p], (uint32_t)tid);
    CL_STAMP(3);
    mbar_wait_parity(&barY[p], par);
    CL_STAMP(4);
    if (tid < TS) {
      float y = sb2[0];
#pragma unroll
      for (int q = 0; q < CL; ++q) y += sYP[q * TS + tid];     // rank order: identical in every CTA
      s_dy[tid] = tid < bs ? 2.f * (y - s_y[p][tid]) / (float)bs : 0.f;
    }
    __syncthreads();
    // ---- head gradients, dz1 in place of h1 ------------------------------------------------------
    float gw2 = 0.f, gb2 = 0.f;
    if (tid < UC) {
      for (int b = 0; b < TS; ++b) gw2 = fmaf(s_dy[b], sH1[b * LDU + tid], gw2);
    } else if (tid == 32) {
      for (int b = 0; b < TS; ++b) gb2 += s_dy[b];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const float h = sH1[(r0 + r) * LDU + u];
      sH1[(r0 + r) * LDU + u] = h > 0.f ? s_dy[r0 + r] * sw2[u] : 0.f;
    }
    __syncthreads();
    // ---- dh0 partials over the owned units for ALL inputs -> sP (slice-major), E3 -------------
    {
      float acc[R][4];
#pragma unroll
      for (int r = 0; r < R; ++r) acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.f;
      pol::mm_acc<R>(acc, sH1 + (warp * R) * LDU, LDU, sW1 + 4 * lane, LDH, UC);
      float* dst = sP + (lane >> 2) * SS + 4 * (lane & 3);
#pragma unroll
      for (int r = 0; r < R; ++r)
        *reinterpret_cast<float4*>(dst + (warp * R + r) * UC) = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid < CL && tid != c) bulk_to_peer(sRECV + c * SS, sP + tid * SS, TS * UC * 4, &barR[p], (uint32_t)tid);
    CL_STAMP(5);
    // ---- (overlaps E3) fc_1 gradient for the owned rows: 8 entries per thread; gb1 -------------
    float g1[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float gb1 = 0.f;
    {
      const float* ha = sH0 + (gi0 >> 4) * SS + (gi0 & 15);
      const float* hb = sH0 + ((gi0 + 64) >> 4) * SS + (gi0 & 15);
#pragma unroll 4
      for (int b = 0; b < TS; ++b) {
        const float d = sH1[b * LDU + gu];
        gb1 += d;      // fc_1 bias gradient of unit gu, same order over the batch as a separate loop (used by the gi0 == 0 thread)
        const float4 x = *reinterpret_cast<const float4*>(ha + b * UC);
        const float4 y = *reinterpret_cast<const float4*>(hb + b * UC);
        g1[0] = fmaf(d, x.x, g1[0]); g1[1] = fmaf(d, x.y, g1[1]); g1[2] = fmaf(d, x.z, g1[2]); g1[3] = fmaf(d, x.w, g1[3]);
        g1[4] = fmaf(d, y.x, g1[4]); g1[5] = fmaf(d, y.y, g1[5]); g1[6] = fmaf(d, y.z, g1[6]); g1[7] = fmaf(d, y.w, g1[7]);
      }
    }
    // ---- Adam on fc_1 rows (old rows were last read by the dh0 partials above) -----------------
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int o = gu * LDH + gi0 + 64 * hh;
      float4 w4 = *reinterpret_cast<float4*>(sW1 + o), m4 = *reinterpret_cast<float4*>(sM1 + o),
             v4 = *reinterpret_cast<float4*>(sV1 + o);
      w4.x = adam(w4.x, g1[4 * hh + 0], m4.x, v4.x, ac); w4.y = adam(w4.y, g1[4 * hh + 1], m4.y, v4.y, ac);
      w4.z = adam(w4.z, g1[4 * hh + 2], m4.z, v4.z, ac); w4.w = adam(w4.w, g1[4 * hh + 3], m4.w, v4.w, ac);
      *reinterpret_cast<float4*>(sW1 + o) = w4; *reinterpret_cast<float4*>(sM1 + o) = m4;
      *reinterpret_cast<float4*>(sV1 + o) = v4;
    }
    if (gi0 == 0) sb1[gu] = adam(sb1[gu], gb1, sb1m[gu], sb1v[gu], ac);
    if (tid < UC) {
      sw2[tid] = adam(sw2[tid], gw2, sw2m[tid], sw2v[tid], ac);
    } else if (tid == 32) {
      sb2[0] = adam(sb2[0], gb2, sb2[1], sb2[2], ac);
    }
    // ---- reduce the received dh0 partials (rank order), dz0 = dh0 * (h0 > 0) -----------------
    CL_STAMP(6);
    mbar_wait_parity(&barR[p], par);
    CL_STAMP(7);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int e = (r0 + r) * UC + u;
      float dsum = 0.f;
#pragma unroll
      for (int q = 0; q < CL; ++q) dsum += (q == c) ? sP[c * SS + e] : sRECV[q * SS + e];
      sDZ0[(r0 + r) * LDU + u] = sH0[c * SS + e] > 0.f ? dsum : 0.f;
    }
    __syncthreads();
    CL_STAMP(9);
    // ---- fc_0 gradient for the owned rows + Adam ------------------------------------------------
    for (int e = tid; e < UC * FP; e += PT) {
      const int uu = e / FP, k = e - uu * FP;
      if (k < F) {
        float g = 0.f, gb0 = 0.f;
#pragma unroll 8
        for (int b = 0; b < TS; ++b) {
          const float dz = sDZ0[b * LDU + uu];
          g = fmaf(dz, X[b * FP + k], g);
          gb0 += dz;     // fc_0 bias gradient of unit uu (used by the k == 0 thread), same order over the batch
        }
        const int o = uu * LD0 + k;
        sW0[o] = adam(sW0[o], g, sM0[o], sV0[o], ac);
        if (k == 0) sb0[uu] = adam(sb0[uu], gb0, sb0m[uu], sb0v[uu], ac);
      }
    }
    CL_STAMP(10);
    CL_STAMP(8);
    // the __syncthreads at the top of the next step orders these writes before its reads
  }
  if (a.phase_cycles && tid == 0 && c == 0)
    for (int q = 0; q < 12; ++q) a.phase_cycles[q] = ph[q];
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  // ---- write back the owned rows and moments -----------------------------------------------------
  for (int i = tid; i < UC * F; i += PT) {
    const int uu = i / F, k = i - uu * F, gi = O.oW0 + (u0 + uu) * F + k, o = uu * LD0 + k;
    a.params[gi] = sW0[o]; a.am[gi] = sM0[o]; a.av[gi] = sV0[o];
  }
  for (int i = tid; i < UC * H; i += PT) {
    const int uu = i >> 7, k = i & (H - 1), gi = O.oW1 + (u0 + uu) * H + k, o = uu * LDH + k;
    a.params[gi] = sW1[o]; a.am[gi] = sM1[o]; a.av[gi] = sV1[o];
  }
  if (tid < UC) {
    a.params[O.ob0 + u0 + tid] = sb0[tid]; a.am[O.ob0 + u0 + tid] = sb0m[tid]; a.av[O.ob0 + u0 + tid] = sb0v[tid];
    a.params[O.ob1 + u0 + tid] = sb1[tid]; a.am[O.ob1 + u0 + tid] = sb1m[tid]; a.av[O.ob1 + u0 + tid] = sb1v[tid];
    a.params[O.oW2 + u0 + tid] = sw2[tid]; a.am[O.oW2 + u0 + tid] = sw2m[tid]; a.av[O.oW2 + u0 + tid] = sw2v[tid];
  }
  if (c == 0 && tid == 0) { a.params[O.ob2] = sb2[0]; a.am[O.ob2] = sb2[1]; a.av[O.ob2] = sb2[2]; }
  cluster_sync_all();   // no CTA exits while a peer may still address its shared memory
}

}  // namespace mlpb

// =========================================================================================
// C-ABI
// =========================================================================================
extern "C" {

int64_t mjrl_mlpbase_param_count(int32_t n) {
  if (n < 1) return MJRL_E_ARG;
  return mlpb::Offsets(n).D;
}

int mjrl_mlpbase_features(mjrl_ctx* ctx, const float* obs, const double* returns,
                          const int32_t* path_offsets, int32_t P, int64_t N, int32_t n, float* feat,
                          float* y32, void* stream) {
  MJRL_REQUIRE(ctx && obs && path_offsets && feat, MJRL_E_ARG, "mlpbase_features: null argument");
  MJRL_REQUIRE(P >= 0 && N >= 0 && n >= 1, MJRL_E_ARG, "mlpbase_features: bad sizes");
  if (N == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = (int)std::min<int64_t>((N + 255) / 256, (int64_t)ctx->sm_count * 8);
  mlpb::features_kernel<<<grid, 256, 0, s>>>(obs, path_offsets, P, N, n, feat);
  int rc = mjrl_check_launch("mlpbase_features_kernel");
  if (rc) return rc;
  if (returns && y32) {
    mlpb::cast_y_kernel<<<grid, 256, 0, s>>>(returns, N, y32);
    rc = mjrl_check_launch("mlpbase_cast_y_kernel");
  }
  return rc;
}

int mjrl_mlpbase_predict(mjrl_ctx* ctx, const float* obs, const int32_t* path_offsets, int32_t P,
                         int64_t N, int32_t n, const float* params, double* pred, void* stream) {
  MJRL_REQUIRE(ctx && obs && path_offsets && params && pred, MJRL_E_ARG, "mlpbase_predict: null argument");
  MJRL_REQUIRE(P >= 0 && N >= 0 && n >= 1, MJRL_E_ARG, "mlpbase_predict: bad sizes");
  if (N == 0) return 0;
  const mlpb::Smem L(n + 4, 1);
  const size_t smem = sizeof(float) * (size_t)L.total;
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE,
               "mlpbase_predict: obs_dim=%d needs %zu B of shared memory (max %d)", n, smem, ctx->max_smem_optin);
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(mlpb::predict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         ctx->max_smem_optin);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr = true;
  }
  const int64_t ntiles = (N + mlpb::TS - 1) / mlpb::TS;
  const int grid = (int)std::min<int64_t>(ntiles, ctx->sm_count);
  mlpb::predict_kernel<<<grid, mlpb::PT, smem, (cudaStream_t)stream>>>(obs, path_offsets, P, N, n, params, pred);
  return mjrl_check_launch("mlpbase_predict_kernel");
}

int mjrl_mlpbase_sqerr(mjrl_ctx* ctx, const double* pred, const double* returns, int64_t N,
                       double* sums2, void* stream) {
  MJRL_REQUIRE(ctx && pred && returns && sums2, MJRL_E_ARG, "mlpbase_sqerr: null argument");
  MJRL_REQUIRE(N >= 0, MJRL_E_ARG, "mlpbase_sqerr: bad size");
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((N + 255) / 256, (int64_t)ctx->sm_count * 4));
  double* part = (double*)mjrl_scratch(ctx, sizeof(double) * 2 * grid);
  if (!part) return MJRL_E_STATE;
  mlpb::sqerr_partial_kernel<<<grid, 256, 0, s>>>(pred, returns, N, part);
  int rc = mjrl_check_launch("mlpbase_sqerr_partial_kernel");
  if (rc) return rc;
  mlpb::sqerr_final_kernel<<<1, 32, 0, s>>>(part, grid, sums2);
  return mjrl_check_launch("mlpbase_sqerr_final_kernel");
}

int mjrl_mlpbase_fit(mjrl_ctx* ctx, const float* feat, const float* y32, int64_t N, int32_t n,
                     float* params, float* adam_m, float* adam_v, int64_t step0, const int32_t* perm,
                     int64_t nsteps, int32_t batch, double lr, double beta1, double beta2, double eps,
                     double weight_decay, void* stream) {
  MJRL_REQUIRE(ctx && feat && y32 && params && adam_m && adam_v, MJRL_E_ARG, "mlpbase_fit: null argument");
  MJRL_REQUIRE(N >= 0 && n >= 1 && nsteps >= 0 && step0 >= 0, MJRL_E_ARG, "mlpbase_fit: bad sizes");
  MJRL_REQUIRE(batch >= 1 && batch <= mlpb::MAXB, MJRL_E_SHAPE,
               "mlpbase_fit: batch_size=%d not supported (1..%d)", batch, mlpb::MAXB);
  if (nsteps == 0) return 0;
  MJRL_REQUIRE(perm, MJRL_E_ARG, "mlpbase_fit: null permutation");
  const mlpb::Smem L(n + 4, 2);
  const size_t smem = sizeof(float) * (size_t)L.total;
  MJRL_REQUIRE(smem + 1024 <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE,
               "mlpbase_fit: obs_dim=%d needs %zu B of shared memory (max %d)", n, smem, ctx->max_smem_optin);
  static size_t attr = 0;   // fit_kernel also has ~1 KB of static shared memory
  if (smem > attr) {
    cudaError_t e = cudaFuncSetAttribute(mlpb::fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
    attr = smem;
  }
  mlpb::FitArgs a;
  a.feat = feat; a.y32 = y32; a.perm = perm; a.params = params; a.am = adam_m; a.av = adam_v;
  a.nsteps = nsteps; a.step0 = step0; a.n = n; a.bs = batch;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay;
  a.phase_cycles = nullptr;
  // Cluster kernel (8 CTAs, model parallel over hidden units) unless it does not fit or
  // MJRL_B200_MLPFIT=single asks for the one-CTA chain.
  {
    const mlpb::CSmem LC(n + 4);
    const size_t csmem = sizeof(float) * (size_t)LC.total;
    static const char* mode = getenv("MJRL_B200_MLPFIT");
    if (!(mode && mode[0] == 's') && csmem + 1024 <= (size_t)ctx->max_smem_optin) {
      static size_t cattr = 0;
      if (csmem > cattr) {
        cudaError_t e = cudaFuncSetAttribute(mlpb::fit_cluster_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csmem);
        if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return (int)e; }
        cattr = csmem;
      }
      static const bool cdbg = getenv("MJRL_B200_PHASE_CYCLES") != nullptr;
      if (cdbg) {
        long long* dev; cudaMalloc(&dev, sizeof(long long) * 12);
        a.phase_cycles = dev;
        mlpb::fit_cluster_kernel<<<mlpb::CL, mlpb::PT, csmem, (cudaStream_t)stream>>>(a);
        cudaStreamSynchronize((cudaStream_t)stream);
        long long h[12]; cudaMemcpy(h, dev, sizeof(h), cudaMemcpyDeviceToHost); cudaFree(dev);
        const char* nm[11] = {"top(wait,sync,prefetch)", "L0+E1 issue", "E1 wait", "L1+head+E2 issue", "E2 wait",
                              "dy,gw2,dz1,dgrad,E3 issue", "gW1+adam1", "E3 wait", "gb0+adam (tail)", "dz0 reduce+sync", "gW0+adam0"};
        for (int q = 0; q < 11; ++q) fprintf(stderr, "[cluster fit] %-28s %8.0f cycles/step\n", nm[q], (double)h[q] / nsteps);
        return mjrl_check_launch("mlpbase_fit_cluster_kernel(dbg)");
      }
      mlpb::fit_cluster_kernel<<<mlpb::CL, mlpb::PT, csmem, (cudaStream_t)stream>>>(a);
      return mjrl_check_launch("mlpbase_fit_cluster_kernel");
    }
  }
  static const bool dbg = getenv("MJRL_B200_PHASE_CYCLES") != nullptr;
  if (dbg) {
    long long* dev; cudaMalloc(&dev, sizeof(long long) * 12);
    a.phase_cycles = dev;
    mlpb::fit_kernel<<<1, mlpb::PT, smem, (cudaStream_t)stream>>>(a);
    cudaStreamSynchronize((cudaStream_t)stream);
    long long h[12]; cudaMemcpy(h, dev, sizeof(h), cudaMemcpyDeviceToHost); cudaFree(dev);
    const char* nm[10] = {"wait+sync", "prefetch", "L0", "L1", "head+gw2", "dz1", "gW1+gb1", "dgrad", "adam1", "gW0+adam0"};
    for (int q = 0; q < 10; ++q) fprintf(stderr, "[fit phase] %-10s %8.0f cycles/step\n", nm[q], (double)h[q] / nsteps);
    return mjrl_check_launch("mlpbase_fit_kernel(dbg)");
  }
  mlpb::fit_kernel<<<1, mlpb::PT, smem, (cudaStream_t)stream>>>(a);
  return mjrl_check_launch("mlpbase_fit_kernel");
}

}  // extern "C"
