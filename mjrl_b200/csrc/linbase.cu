// (b) LinearBaseline: fused feature build + fp64 Gram accumulation, on-device regularised
// Cholesky solve with the reference's NaN-retry ladder, and predict.
//
// Replaces mjrl/baselines/linear_baseline.py:10-18 (_features), :20-44 (fit), :46-49 (predict).
// Features per timestep: [clip(obs,-10,10) (n), al, al^2, al^3, 1], al = t_in_path/1000; F = n+4.
//
// Precision: features are built from fp32 obs widened to fp64, all products and sums in
// fp64 (TF32/fp32 accumulation breaks the 1e-6 tolerance).  The contraction is
// HBM-bound for small n (4n+8 B per timestep) and fp64-FMA-bound for n ~ 100.
#include "common.cuh"

namespace {

// ------------------------------------------------------------------ predict
constexpr int PRED_TS = 256;  // timesteps per tile == threads per CTA

__global__ void __launch_bounds__(PRED_TS)
linbase_predict_kernel(const float* __restrict__ obs, const int32_t* __restrict__ off, int P,
                       int64_t N, int n, const double* __restrict__ coeffs,
                       double* __restrict__ pred) {
  extern __shared__ unsigned char smem_raw[];
  const int ns = n | 1;  // odd row stride: thread-per-row reads are bank-conflict free
  double* sc = reinterpret_cast<double*>(smem_raw);           // [n+4]
  float* so = reinterpret_cast<float*>(sc + ((n + 4 + 1) & ~1));  // [PRED_TS][ns]
  for (int k = threadIdx.x; k < n + 4; k += blockDim.x) sc[k] = coeffs[k];
  const int64_t ntiles = (N + PRED_TS - 1) / PRED_TS;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t t0 = tile * PRED_TS;
    const int rows = (int)min((int64_t)PRED_TS, N - t0);
    __syncthreads();
    // contiguous, coalesced copy of rows*n floats
    const float* src = obs + t0 * n;
    const int cnt = rows * n;
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      int r = i / n, c = i - r * n;
      so[r * ns + c] = ld_stream(src + i);
    }
    __syncthreads();
    const int r = threadIdx.x;
    if (r < rows) {
      const int64_t t = t0 + r;
      const int p = find_path(off, P, t);
      const double al = (double)(t - off[p]) / 1000.0;
      double acc = 0.0;
      const float* row = so + r * ns;
      for (int k = 0; k < n; ++k) {
        double o = (double)fminf(fmaxf(row[k], -10.f), 10.f);
        acc = fma(o, sc[k], acc);
      }
      const double al2 = al * al;
      acc = fma(al, sc[n], acc);
      acc = fma(al2, sc[n + 1], acc);
      acc = fma(al2 * al, sc[n + 2], acc);
      acc += sc[n + 3];
      pred[t] = acc;
    }
  }
}

__global__ void fill_zero_kernel(double* __restrict__ x, int64_t N) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < N; i += stride) x[i] = 0.0;
}

// --------------------------------------------------------------------- Gram
// Symmetric (F+1)x(F+1) matrix S = Z^T Z with Z = [X | y]; only 4x4 tiles with tj >= ti are
// computed.  Each thread owns up to MAXT tiles (16 fp64 accumulators each) for the whole
// persistent loop; when there are fewer tiles than threads the timesteps of a smem tile are
// split over G thread groups.  Per-CTA partials go to scratch, a second kernel sums them in a
// fixed order and mirrors the triangle.
constexpr int GRAM_THREADS = 256;

template <int MAXT>
__global__ void __launch_bounds__(GRAM_THREADS)
linbase_gram_kernel(const float* __restrict__ obs, const double* __restrict__ ret,
                    const int32_t* __restrict__ off, int P, int64_t N, int n, int TS, int G,
                    int ntile_side, int ntiles_tri, double* __restrict__ partial) {
  extern __shared__ unsigned char smem_raw[];
  const int F1 = n + 5;               // features + y
  const int FP = (F1 + 3) & ~3;       // padded row length (multiple of 4)
  double* feat = reinterpret_cast<double*>(smem_raw);  // [TS][FP]
  const int tid = threadIdx.x;

  // tile ownership: slot s = tid % ntiles_tri (+ k*GRAM_THREADS for MAXT>1), group = tid / ntiles_tri
  int ti[MAXT], tj[MAXT];
  bool own[MAXT];
  int grp = 0;
#pragma unroll
  for (int k = 0; k < MAXT; ++k) {
    int s;
    if (MAXT == 1) { s = tid % ntiles_tri; grp = tid / ntiles_tri; own[k] = grp < G; }
    else { s = tid + k * GRAM_THREADS; grp = 0; own[k] = s < ntiles_tri; }
    // decode s -> (ti, tj) with tj >= ti, row-major over the upper triangle
    int a = 0, rem = own[k] ? s : 0, rowlen = ntile_side;
    while (rem >= rowlen) { rem -= rowlen; ++a; --rowlen; }
    ti[k] = a; tj[k] = a + rem;
  }
  double acc[MAXT][16];
#pragma unroll
  for (int k = 0; k < MAXT; ++k)
#pragma unroll
    for (int e = 0; e < 16; ++e) acc[k][e] = 0.0;

  const int64_t ntiles = (N + TS - 1) / TS;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t t0 = tile * TS;
    const int rows = (int)min((int64_t)TS, N - t0);
    __syncthreads();
    // obs block (coalesced) -> clipped fp64 features
    const float* src = obs + t0 * n;
    const int cnt = rows * n;
    for (int i = tid; i < cnt; i += GRAM_THREADS) {
      int r = i / n, c = i - r * n;
      feat[r * FP + c] = (double)fminf(fmaxf(ld_stream(src + i), -10.f), 10.f);
    }
    for (int r = tid; r < TS; r += GRAM_THREADS) {
      double* row = feat + r * FP;
      if (r < rows) {
        const int64_t t = t0 + r;
        const int p = find_path(off, P, t);
        const double al = (double)(t - off[p]) / 1000.0;
        row[n] = al; row[n + 1] = al * al; row[n + 2] = al * al * al; row[n + 3] = 1.0;
        row[n + 4] = ld_stream(ret + t);
        for (int c = F1; c < FP; ++c) row[c] = 0.0;
      } else {
        for (int c = 0; c < FP; ++c) row[c] = 0.0;  // zero rows contribute nothing
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < MAXT; ++k) {
      if (!own[k]) continue;
      const double* pa = feat + 4 * ti[k];
      const double* pb = feat + 4 * tj[k];
      for (int r = grp; r < TS; r += G) {
        const double2 a01 = *reinterpret_cast<const double2*>(pa + r * FP);
        const double2 a23 = *reinterpret_cast<const double2*>(pa + r * FP + 2);
        const double2 b01 = *reinterpret_cast<const double2*>(pb + r * FP);
        const double2 b23 = *reinterpret_cast<const double2*>(pb + r * FP + 2);
        const double a[4] = {a01.x, a01.y, a23.x, a23.y};
        const double b[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
          for (int y = 0; y < 4; ++y) acc[k][x * 4 + y] = fma(a[x], b[y], acc[k][x * 4 + y]);
      }
    }
  }
  // ---- reduce the G groups of this CTA in a fixed order, write the CTA partial ----
  __syncthreads();
  double* red = feat;  // reuse: [G][ntiles_tri][16] may exceed the feature tile -> sized by host
  const int S1 = FP;   // partial layout: dense [FP][FP] upper-tile entries
  double* out = partial + (size_t)blockIdx.x * S1 * S1;
  if (MAXT == 1) {
    if (own[0]) {
      double* dst = red + ((size_t)grp * ntiles_tri + (tid % ntiles_tri)) * 16;
#pragma unroll
      for (int e = 0; e < 16; ++e) dst[e] = acc[0][e];
    }
    __syncthreads();
    for (int i = tid; i < ntiles_tri * 16; i += GRAM_THREADS) {
      double s = 0.0;
      for (int g = 0; g < G; ++g) s += red[(size_t)g * ntiles_tri * 16 + i];
      int slot = i >> 4, e = i & 15;
      int a = 0, rem = slot, rowlen = ntile_side;
      while (rem >= rowlen) { rem -= rowlen; ++a; --rowlen; }
      int b = a + rem;
      out[(4 * a + (e >> 2)) * S1 + 4 * b + (e & 3)] = s;
    }
  } else {
#pragma unroll
    for (int k = 0; k < MAXT; ++k) {
      if (!own[k]) continue;
#pragma unroll
      for (int e = 0; e < 16; ++e)
        out[(4 * ti[k] + (e >> 2)) * S1 + 4 * tj[k] + (e & 3)] = acc[k][e];
    }
  }
}

// gram[i*(F+1)+j] = sum_cta S[i][j] (upper) mirrored; gram[F*(F+1)] = y^T y
__global__ void linbase_gram_reduce_kernel(const double* __restrict__ partial, int nparts, int F,
                                           int FP, double* __restrict__ gram) {
  const int F1 = F + 1;
  const int total = F * F1 + 1;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    int i, j;
    if (idx == F * F1) { i = F; j = F; } else { i = idx / F1; j = idx - i * F1; }
    int a = min(i, j), b = max(i, j);
    // entry (a,b) lives in tile (a/4, b/4) which is stored iff b/4 >= a/4 -- always true
    double s = 0.0;
    for (int c = 0; c < nparts; ++c) s += partial[(size_t)c * FP * FP + a * FP + b];
    gram[idx] = s;
  }
}

// -------------------------------------------------------------------- solve
// One CTA.  A = X^T X + lam I in shared memory (F <= 160), right-looking Cholesky, two
// triangular solves.  Retry ladder: if any coefficient is NaN, lam *= 10, up to 10 tries
// (linear_baseline.py:30-38).  Errors from the Gram: sum (y - Xc)^2 = yy - 2 c.b + c^T G c.
__global__ void __launch_bounds__(256)
linbase_solve_kernel(const double* __restrict__ gram, int F, double reg, const double* __restrict__ cprev,
                     double* __restrict__ coeffs, double* __restrict__ info) {
  extern __shared__ unsigned char smem_raw[];
  double* A = reinterpret_cast<double*>(smem_raw);  // [F][F+1] (padded stride)
  const int ld = F + 1;
  double* x = A + (size_t)F * ld;                   // [F]
  __shared__ double red[32];
  __shared__ int s_bad;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int F1 = F + 1;
  const double yy = gram[F * F1];

  auto quad_err = [&](const double* c) -> double {  // (yy - 2 c.b + c^T G c)/yy, warp 0 valid
    double acc = 0.0;
    for (int i = tid; i < F; i += nt) {
      double gi = 0.0;
      for (int j = 0; j < F; ++j) gi = fma(gram[i * F1 + j], c[j], gi);
      acc += c[i] * (gi - 2.0 * gram[i * F1 + F]);
    }
    double r = block_sum(acc, red);
    return (yy + r) / yy;
  };

  if (cprev) {
    double e = quad_err(cprev);
    if (tid == 0) info[2] = e;
  } else if (tid == 0) {
    info[2] = 1.0;  // zero predictions: sum y^2 / sum y^2
  }

  double lam = reg;
  int tries = 0;
  for (; tries < 10; ++tries) {
    __syncthreads();
    for (int idx = tid; idx < F * F; idx += nt) {
      int i = idx / F, j = idx - i * F;
      A[i * ld + j] = gram[i * F1 + j] + (i == j ? lam : 0.0);
    }
    for (int i = tid; i < F; i += nt) x[i] = gram[i * F1 + F];
    __syncthreads();
    for (int k = 0; k < F; ++k) {
      const double d = sqrt(A[k * ld + k]);  // NaN if the pivot is negative
      __syncthreads();
      for (int i = k + tid; i < F; i += nt) A[i * ld + k] = (i == k) ? d : A[i * ld + k] / d;
      __syncthreads();
      // trailing update of the lower triangle: rows i > k, cols k < j <= i
      const int m = F - k - 1;
      for (int idx = tid; idx < m * m; idx += nt) {
        int i = k + 1 + idx / m, j = k + 1 + idx % m;
        if (j <= i) A[i * ld + j] = fma(-A[i * ld + k], A[j * ld + k], A[i * ld + j]);
      }
      __syncthreads();
    }
    // forward substitution L z = b
    for (int k = 0; k < F; ++k) {
      if (tid == 0) x[k] = x[k] / A[k * ld + k];
      __syncthreads();
      const double zk = x[k];
      for (int i = k + 1 + tid; i < F; i += nt) x[i] = fma(-A[i * ld + k], zk, x[i]);
      __syncthreads();
    }
    // back substitution L^T c = z
    for (int k = F - 1; k >= 0; --k) {
      if (tid == 0) x[k] = x[k] / A[k * ld + k];
      __syncthreads();
      const double ck = x[k];
      for (int i = tid; i < k; i += nt) x[i] = fma(-A[k * ld + i], ck, x[i]);
      __syncthreads();
    }
    if (tid == 0) s_bad = 0;
    __syncthreads();
    for (int i = tid; i < F; i += nt)
      if (isnan(x[i])) s_bad = 1;
    __syncthreads();
    if (!s_bad) break;
    lam *= 10.0;
  }
  for (int i = tid; i < F; i += nt) coeffs[i] = x[i];
  __syncthreads();
  double e = quad_err(x);
  if (tid == 0) {
    info[0] = (tries < 10) ? lam : lam / 10.0;
    info[1] = (double)(tries < 10 ? tries : 9);
    info[3] = e;
  }
}

}  // namespace

extern "C" {

int mjrl_linbase_predict(mjrl_ctx* ctx, const float* obs, const int32_t* off, int32_t P, int64_t N,
                         int32_t n, const double* coeffs, double* pred, void* stream) {
  MJRL_REQUIRE(ctx && N >= 0 && n > 0 && P >= 0, MJRL_E_ARG, "linbase_predict: bad ctx/N/n/P");
  if (N == 0) return 0;
  MJRL_REQUIRE(pred, MJRL_E_ARG, "linbase_predict: null pred");
  cudaStream_t s = (cudaStream_t)stream;
  if (!coeffs) {  // never fit: zeros (linear_baseline.py:46-48)
    fill_zero_kernel<<<ctx->sm_count * 4, 256, 0, s>>>(pred, N);
    return mjrl_check_launch("linbase_predict(zero)");
  }
  MJRL_REQUIRE(obs && off, MJRL_E_ARG, "linbase_predict: null pointer");
  size_t smem = sizeof(double) * ((n + 4 + 1) & ~1) + sizeof(float) * PRED_TS * (n | 1);
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE, "linbase_predict: n=%d too large", n);
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(linbase_predict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int64_t ntiles = (N + PRED_TS - 1) / PRED_TS;
  int grid = (int)(ntiles < (int64_t)ctx->sm_count * 8 ? ntiles : (int64_t)ctx->sm_count * 8);
  linbase_predict_kernel<<<grid, PRED_TS, smem, s>>>(obs, off, P, N, n, coeffs, pred);
  return mjrl_check_launch("linbase_predict");
}

int mjrl_linbase_gram(mjrl_ctx* ctx, const float* obs, const double* returns, const int32_t* off,
                      int32_t P, int64_t N, int32_t n, double* gram, void* stream) {
  MJRL_REQUIRE(ctx && obs && returns && off && gram && N > 0 && n > 0 && P > 0, MJRL_E_ARG,
               "linbase_gram: bad argument");
  const int F = n + 4, F1 = F + 1, FP = (F1 + 3) & ~3;
  const int side = FP / 4, ntri = side * (side + 1) / 2;
  MJRL_REQUIRE(ntri <= 2 * GRAM_THREADS, MJRL_E_SHAPE,
               "linbase_gram: n=%d gives %d Gram tiles (> %d)", n, ntri, 2 * GRAM_THREADS);
  const int maxt = ntri <= GRAM_THREADS ? 1 : 2;
  int TS = FP <= 32 ? 128 : 32;
  int G = 1;
  if (maxt == 1) { G = GRAM_THREADS / ntri; if (G > TS) G = TS; }
  size_t smem_feat = sizeof(double) * (size_t)TS * FP;
  size_t smem_red = maxt == 1 ? sizeof(double) * (size_t)G * ntri * 16 : 0;
  size_t smem = smem_feat > smem_red ? smem_feat : smem_red;
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE, "linbase_gram: smem %zu", smem);
  int64_t ntiles = (N + TS - 1) / TS;
  int grid = (int)(ntiles < (int64_t)ctx->sm_count * 2 ? ntiles : (int64_t)ctx->sm_count * 2);
  double* partial = (double*)mjrl_scratch2(ctx, sizeof(double) * (size_t)grid * FP * FP);
  if (!partial) return MJRL_E_STATE;
  cudaStream_t s = (cudaStream_t)stream;
  // partials are written sparsely (upper tiles only); entries never written are never read
  if (maxt == 1) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(linbase_gram_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    linbase_gram_kernel<1><<<grid, GRAM_THREADS, smem, s>>>(obs, returns, off, P, N, n, TS, G, side,
                                                            ntri, partial);
  } else {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(linbase_gram_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    linbase_gram_kernel<2><<<grid, GRAM_THREADS, smem, s>>>(obs, returns, off, P, N, n, TS, G, side,
                                                            ntri, partial);
  }
  int rc = mjrl_check_launch("linbase_gram");
  if (rc) return rc;
  int total = F * F1 + 1;
  linbase_gram_reduce_kernel<<<(total + 127) / 128, 128, 0, s>>>(partial, grid, F, FP, gram);
  return mjrl_check_launch("linbase_gram_reduce");
}

int mjrl_linbase_solve(mjrl_ctx* ctx, const double* gram, int32_t F, double reg_coeff,
                       const double* coeffs_prev, double* coeffs, double* info4, void* stream) {
  MJRL_REQUIRE(ctx && gram && coeffs && info4 && F > 0, MJRL_E_ARG, "linbase_solve: bad argument");
  size_t smem = sizeof(double) * ((size_t)F * (F + 1) + F);
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE,
               "linbase_solve: F=%d needs %zu B of shared memory", F, smem);
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(linbase_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  linbase_solve_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(gram, F, reg_coeff, coeffs_prev, coeffs, info4);
  return mjrl_check_launch("linbase_solve");
}

}  // extern "C"
