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
// One thread per timestep, 256 timesteps per tile.  The obs slab of a tile is contiguous in HBM
// (256*n floats): one thread issues a single bulk copy (TMA) into shared memory, double-buffered,
// so the next tile streams in while this one is reduced and no thread spends instructions on
// staging.  The path of the tile's first timestep is found once per warp (32-ary search).
constexpr int PRED_TS = 256;  // timesteps per tile == threads per CTA

// sum_k clip(row[k]) * sc[k] in ascending k for every vector width (identical rounding)
__device__ __forceinline__ double clipd(float x) { return (double)fminf(fmaxf(x, -10.f), 10.f); }
template <int VEC>
__device__ __forceinline__ double pred_dot(const float* __restrict__ row, const double* __restrict__ sc,
                                           int n) {
  double acc = 0.0;
  int k = 0;
  for (; k + 4 <= n; k += 4) {
    float x0, x1, x2, x3;
    if (VEC == 4) {
      const float4 v = *reinterpret_cast<const float4*>(row + k);
      x0 = v.x; x1 = v.y; x2 = v.z; x3 = v.w;
    } else if (VEC == 2) {
      const float2 u = *reinterpret_cast<const float2*>(row + k);
      const float2 v = *reinterpret_cast<const float2*>(row + k + 2);
      x0 = u.x; x1 = u.y; x2 = v.x; x3 = v.y;
    } else {
      x0 = row[k]; x1 = row[k + 1]; x2 = row[k + 2]; x3 = row[k + 3];
    }
    const double2 c01 = *reinterpret_cast<const double2*>(sc + k);
    const double2 c23 = *reinterpret_cast<const double2*>(sc + k + 2);
    acc = fma(clipd(x0), c01.x, acc);
    acc = fma(clipd(x1), c01.y, acc);
    acc = fma(clipd(x2), c23.x, acc);
    acc = fma(clipd(x3), c23.y, acc);
  }
  for (; k < n; ++k) acc = fma(clipd(row[k]), sc[k], acc);
  return acc;
}

__global__ void __launch_bounds__(PRED_TS)
linbase_predict_kernel(const float* __restrict__ obs, const int32_t* __restrict__ off, int P,
                       int64_t N, int n, int nbuf, const double* __restrict__ coeffs,
                       double* __restrict__ pred) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar[2];
  __shared__ int sTP[2][4];
  double* sc = reinterpret_cast<double*>(smem_raw);                  // [n+4]
  float* tiles = reinterpret_cast<float*>(sc + ((n + 4 + 1) & ~1));  // [nbuf][PRED_TS*n], dense rows
  const int tid = threadIdx.x;
  const int ntiles = (int)((N + PRED_TS - 1) / PRED_TS);
  const int tile_elems = PRED_TS * n;
  const bool base_ok = (reinterpret_cast<uintptr_t>(obs) & 15) == 0;
  // a tile goes through the bulk path iff its byte count is a multiple of 16 (all but maybe the last)
  auto tile_rows = [&](int tile) { return tile + 1 < ntiles ? PRED_TS : (int)(N - (int64_t)tile * PRED_TS); };
  auto bulk_ok = [&](int tile) { return base_ok && ((tile_rows(tile) * n) & 3) == 0; };
  const double inv = (double)P / (double)N;
  auto start_copy = [&](int tile, int buf) {  // one thread
    if (bulk_ok(tile)) {
      const uint32_t bytes = (uint32_t)(tile_rows(tile) * n) * 4u;
      tma::mbar_expect(&bar[buf], bytes);
      tma::bulk_g2s(tiles + (size_t)buf * tile_elems, obs + (int64_t)tile * tile_elems, bytes, &bar[buf]);
    }
  };
  auto publish = [&](int slot, const TilePath& tp) {
    sTP[slot][0] = tp.p0; sTP[slot][1] = tp.s0; sTP[slot][2] = tp.s1;
  };
  for (int k = tid; k < n + 4; k += PRED_TS) sc[k] = coeffs[k];
  if (tid == 0) { tma::mbar_init(&bar[0], 1); tma::mbar_init(&bar[1], 1); tma::mbar_init_fence(); }
  __syncthreads();
  int tile = blockIdx.x;
  if (tid < 32 && tile < ntiles) {
    if (tid == 0 && nbuf == 2) start_copy(tile, 0);
    const TilePath tp = tile_path(off, P, (int64_t)tile * PRED_TS, inv);
    if (tid == 0) publish(0, tp);
  }
  __syncthreads();
  for (int it = 0; tile < ntiles; tile += gridDim.x, ++it) {
    const int slot = it & 1, buf = nbuf == 2 ? slot : 0;
    // warp 0 looks one tile ahead: the copy starts now, the path lookup's loads are in flight while
    // the warp reduces its own rows and are consumed just before the closing barrier
    const int next = tile + gridDim.x;
    const bool ahead = tid < 32 && next < ntiles;
    TilePath tpn;
    if (tid == 0) {
      if (nbuf == 2) { if (next < ntiles) start_copy(next, buf ^ 1); }
      else start_copy(tile, 0);
    }
    if (ahead) tpn = tile_path(off, P, (int64_t)next * PRED_TS, inv);
    const int64_t t0 = (int64_t)tile * PRED_TS;
    const int rows = tile_rows(tile);
    TilePath tp;
    tp.p0 = sTP[slot][0]; tp.s0 = sTP[slot][1]; tp.s1 = sTP[slot][2];
    float* so = tiles + (size_t)buf * tile_elems;
    if (bulk_ok(tile)) {
      tma::mbar_wait(&bar[buf], nbuf == 2 ? ((it >> 1) & 1) : (it & 1));
    } else {
      stream_rows<PRED_TS>(obs + t0 * n, rows * n, n, tid, [&](int r, int c, float v) { so[r * n + c] = v; });
      __syncthreads();
    }
    if (tid < rows) {
      const int64_t t = t0 + tid;
      const double al = (double)t_in_path(off, P, tp, t) / 1000.0;
      const float* row = so + tid * n;
      double acc = (n & 3) == 0 ? pred_dot<4>(row, sc, n)
                 : (n & 1) == 0 ? pred_dot<2>(row, sc, n) : pred_dot<1>(row, sc, n);
      const double al2 = al * al;
      acc = fma(al, sc[n], acc);
      acc = fma(al2, sc[n + 1], acc);
      acc = fma(al2 * al, sc[n + 2], acc);
      acc += sc[n + 3];
      pred[t] = acc;
    }
    if (ahead && tid == 0) publish(slot ^ 1, tpn);
    __syncthreads();  // every thread is done with `so` and sTP[slot] before they are refilled
  }
}

// QuadraticBaseline predict (quadratic_baseline.py:66-68): one thread per timestep; its o row is
// widened once into shared memory, then linear + upper-triangle products + bias + time powers.
constexpr int QPRED_TS = 128;
__global__ void __launch_bounds__(QPRED_TS)
quadbase_predict_kernel(const float* __restrict__ obs, const int32_t* __restrict__ off, int P,
                        int64_t N, int n, int F, const double* __restrict__ coeffs,
                        double* __restrict__ pred) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sc = reinterpret_cast<double*>(smem_raw);            // [F]
  double* so = sc + ((F + 1) & ~1);                            // [QPRED_TS][n | 1]
  const int ns = n | 1, tid = threadIdx.x;
  for (int k = tid; k < F; k += QPRED_TS) sc[k] = coeffs[k];
  const int64_t ntiles = (N + QPRED_TS - 1) / QPRED_TS;
  const double inv = (double)P / (double)N;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t t0 = tile * QPRED_TS;
    const int rows = (int)min((int64_t)QPRED_TS, N - t0);
    __syncthreads();
    const TilePath tp = tile_path(off, P, t0, inv);
    stream_rows<QPRED_TS>(obs + t0 * n, rows * n, n, tid, [&](int r, int c, float v) {
      so[r * ns + c] = (double)fminf(fmaxf(v, -10.f), 10.f) / 10.0;
    });
    __syncthreads();
    if (tid < rows) {
      const int64_t t = t0 + tid;
      const double* o = so + tid * ns;
      double acc = 0.0;
      for (int k = 0; k < n; ++k) acc = fma(o[k], sc[k], acc);
      int qq = n;
      for (int i = 0; i < n; ++i) {
        const double oi = o[i];
        for (int j = i; j < n; ++j, ++qq) acc = fma(oi * o[j], sc[qq], acc);
      }
      const double al = (double)t_in_path(off, P, tp, t) / 1000.0, al2 = al * al;
      acc += sc[qq];
      acc = fma(al, sc[qq + 1], acc);
      acc = fma(al2, sc[qq + 2], acc);
      acc = fma(al2 * al, sc[qq + 3], acc);
      acc = fma(al2 * al2, sc[qq + 4], acc);
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
// Symmetric (F+1)x(F+1) matrix S = Z^T Z with Z = [X | y], on the fp64 tensor-core path
// (mma.sync m8n8k4 f64: the only fp64 MMA sm_100a has; tcgen05 has no fp64 kind).  The matrix is
// cut into 8x8 tiles, only tiles with tj >= ti are computed.  One MMA contracts 4 timesteps:
//   A fragment (8 features x 4 timesteps, row-major):  lane holds Z[t0 + lane%4][8*ti + lane/4]
//   B fragment (4 timesteps x 8 features, col-major):  lane holds Z[t0 + lane%4][8*tj + lane/4]
// i.e. both come from the SAME shared-memory feature tile with the same index map, one 8-byte load
// each.  Row stride FS = 4 mod 16 doubles makes those loads bank-conflict free.
//   SPLIT (few tiles, small n): every warp owns all NS*(NS+1)/2 tiles and takes every 8th group
//     of 4 timesteps; the 8 warp partials are summed in a fixed order at the end.
//   otherwise: warp w owns a contiguous run of <= MAXT tiles and walks all timesteps.
// Per-CTA partials go to scratch, a second kernel sums them in a fixed order and mirrors.
constexpr int GRAM_THREADS = 256;
constexpr int GRAM_WARPS = GRAM_THREADS / 32;

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// fills feat[TS][FS] for the tile starting at timestep t0 (rows valid, the rest zero)
__device__ __forceinline__ void gram_stage_tile(const float* __restrict__ obs,
                                                const double* __restrict__ ret,
                                                const int32_t* __restrict__ off, int P, double inv,
                                                int64_t t0, int rows, int n, int TS, int FP8, int FS,
                                                double* __restrict__ feat, int tid) {
  const int F1 = n + 5;
  const TilePath tp = tile_path(off, P, t0, inv);
  // time features / target / padding first: the lookup latency overlaps the obs copy below
  for (int r = tid; r < TS; r += GRAM_THREADS) {
    double* row = feat + r * FS;
    if (r < rows) {
      const int64_t t = t0 + r;
      const double y = ldg_stream(ret + t);
      const double al = (double)t_in_path(off, P, tp, t) / 1000.0;
      row[n] = al; row[n + 1] = al * al; row[n + 2] = al * al * al; row[n + 3] = 1.0;
      row[n + 4] = y;
      for (int c = F1; c < FP8; ++c) row[c] = 0.0;
    } else {
      for (int c = 0; c < FP8; ++c) row[c] = 0.0;  // zero rows contribute nothing
    }
  }
  stream_rows<GRAM_THREADS>(obs + t0 * n, rows * n, n, tid, [&](int r, int c, float v) {
    feat[r * FS + c] = (double)fminf(fmaxf(v, -10.f), 10.f);
  });
}

// QuadraticBaseline features (quadratic_baseline.py:10-37): [o (n) | o_i o_j, i <= j row-major (Q) | 1 |
// al, al^2, al^3, al^4 | y | zero padding], o = clip(obs,-10,10)/10.  `pairs` = the (i | j << 8) table
// of the Q products in shared memory.  Contains a CTA barrier: the products read o of other threads.
__device__ __forceinline__ void gram_stage_tile_quad(const float* __restrict__ obs,
                                                     const double* __restrict__ ret,
                                                     const int32_t* __restrict__ off, int P, double inv,
                                                     int64_t t0, int rows, int n, int Q, int TS, int FP8,
                                                     int FS, const int* __restrict__ pairs,
                                                     double* __restrict__ feat, int tid) {
  const int c1 = n + Q;          // column of the constant 1
  const int F1 = c1 + 6;         // 1, four powers of al, y
  const TilePath tp = tile_path(off, P, t0, inv);
  for (int r = tid; r < TS; r += GRAM_THREADS) {
    double* row = feat + r * FS;
    if (r < rows) {
      const int64_t t = t0 + r;
      const double y = ldg_stream(ret + t);
      const double al = (double)t_in_path(off, P, tp, t) / 1000.0;
      const double al2 = al * al;
      row[c1] = 1.0; row[c1 + 1] = al; row[c1 + 2] = al2; row[c1 + 3] = al2 * al; row[c1 + 4] = al2 * al2;
      row[c1 + 5] = y;
      for (int c = F1; c < FP8; ++c) row[c] = 0.0;
    } else {
      for (int c = 0; c < FP8; ++c) row[c] = 0.0;  // zero rows contribute nothing
    }
  }
  stream_rows<GRAM_THREADS>(obs + t0 * n, rows * n, n, tid, [&](int r, int c, float v) {
    feat[r * FS + c] = (double)fminf(fmaxf(v, -10.f), 10.f) / 10.0;
  });
  __syncthreads();
  for (int idx = tid; idx < rows * Q; idx += GRAM_THREADS) {
    const int r = idx / Q, qq = idx - r * Q;
    const int pr = pairs[qq];
    const double* row = feat + r * FS;
    feat[r * FS + n + qq] = row[pr & 255] * row[pr >> 8];
  }
}

// SPLIT variant (n <= 19).  The obs slab of a tile arrives by one bulk copy (TMA), double-buffered,
// and stays fp32 in shared memory.  Columns are cut into 8-wide blocks: the first n/8 blocks are
// pure obs columns (clipped + widened when a fragment is loaded); everything after them -- the
// last n%8 obs columns, [al, al^2, al^3, 1, y] and the zero padding -- is written once per timestep
// as fp64 into `ex` by one thread per timestep, so those fragments are a single 8-byte load.
// Warp 0 looks one tile ahead (path lookup + bulk copy) while the others compute.
template <int NS>
__global__ void __launch_bounds__(GRAM_THREADS)
linbase_gram_split_kernel(const float* __restrict__ obs, const double* __restrict__ ret,
                          const int32_t* __restrict__ off, int P, int64_t N, int n, int TS, int EXS,
                          double* __restrict__ partial) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar[2];
  __shared__ int sTP[2][4];
  constexpr int FP8 = 8 * NS, NTRI = NS * (NS + 1) / 2;
  const int tile_elems = TS * n;
  float* tiles = reinterpret_cast<float*>(smem_raw);                               // [2][TS*n]
  double* ex = reinterpret_cast<double*>(smem_raw + (size_t)2 * tile_elems * 4);   // [TS][EXS]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int nb_obs = n >> 3, tail = n & 7;   // pure-obs blocks; obs columns that live in `ex`
  const int exw = 8 * (NS - nb_obs);         // fp64 columns per timestep (8 or 16)
  double acc[NTRI][2];
#pragma unroll
  for (int k = 0; k < NTRI; ++k) acc[k][0] = acc[k][1] = 0.0;

  const int ntiles = (int)((N + TS - 1) / TS);
  const int nq = TS >> 2;
  const bool base_ok = (reinterpret_cast<uintptr_t>(obs) & 15) == 0;
  auto tile_rows = [&](int tile) { return tile + 1 < ntiles ? TS : (int)(N - (int64_t)tile * TS); };
  auto bulk_ok = [&](int tile) { return base_ok && ((tile_rows(tile) * n) & 3) == 0; };
  const double inv = (double)P / (double)N;
  auto start_copy = [&](int tile, int buf) {  // one thread
    if (bulk_ok(tile)) {
      const uint32_t bytes = (uint32_t)(tile_rows(tile) * n) * 4u;
      tma::mbar_expect(&bar[buf], bytes);
      tma::bulk_g2s(tiles + (size_t)buf * tile_elems, obs + (int64_t)tile * tile_elems, bytes, &bar[buf]);
    }
  };
  auto publish = [&](int slot, const TilePath& tp) {
    sTP[slot][0] = tp.p0; sTP[slot][1] = tp.s0; sTP[slot][2] = tp.s1;
  };
  if (tid == 0) { tma::mbar_init(&bar[0], 1); tma::mbar_init(&bar[1], 1); tma::mbar_init_fence(); }
  __syncthreads();
  int tile = blockIdx.x;
  if (tid < 32 && tile < ntiles) {
    if (tid == 0) start_copy(tile, 0);
    const TilePath tp0 = tile_path(off, P, (int64_t)tile * TS, inv);
    if (tid == 0) publish(0, tp0);
  }
  __syncthreads();
  for (int it = 0; tile < ntiles; tile += gridDim.x, ++it) {
    const int buf = it & 1;
    // warp 0 looks one tile ahead: copy now, path lookup consumed just before the closing barrier
    const int next = tile + gridDim.x;
    const bool ahead = tid < 32 && next < ntiles;
    TilePath tpn;
    if (tid == 0 && next < ntiles) start_copy(next, buf ^ 1);
    if (ahead) tpn = tile_path(off, P, (int64_t)next * TS, inv);
    const int64_t t0 = (int64_t)tile * TS;
    const int rows = tile_rows(tile);
    TilePath tp;
    tp.p0 = sTP[buf][0]; tp.s0 = sTP[buf][1]; tp.s1 = sTP[buf][2];
    // per-timestep scalars first: their global loads overlap the wait for the tile
    double y = 0.0, al = 0.0;
    if (tid < rows) {  // TS <= GRAM_THREADS
      y = ldg_stream(ret + t0 + tid);
      al = (double)t_in_path(off, P, tp, t0 + tid) / 1000.0;
    }
    float* of = tiles + (size_t)buf * tile_elems;
    if (bulk_ok(tile)) {
      tma::mbar_wait(&bar[buf], (it >> 1) & 1);
    } else {
      stream_rows<GRAM_THREADS>(obs + t0 * n, rows * n, n, tid, [&](int r, int c, float v) { of[r * n + c] = v; });
      __syncthreads();
    }
    if (rows < TS)  // last tile: rows past the end must read as zero
      for (int i = rows * n + tid; i < tile_elems; i += GRAM_THREADS) of[i] = 0.f;
    if (tid < TS) {
      double* e = ex + tid * EXS;
      if (tid < rows) {
        const float* orow = of + tid * n + 8 * nb_obs;
        for (int c = 0; c < tail; ++c) e[c] = (double)fminf(fmaxf(orow[c], -10.f), 10.f);
        e[tail] = al; e[tail + 1] = al * al; e[tail + 2] = al * al * al; e[tail + 3] = 1.0; e[tail + 4] = y;
        for (int c = tail + 5; c < exw; ++c) e[c] = 0.0;
      } else {
        for (int c = 0; c < exw; ++c) e[c] = 0.0;  // zero rows contribute nothing
      }
    }
    __syncthreads();
    const float* orow = of + (4 * warp + q) * n + g;
    const double* erow = ex + (4 * warp + q) * EXS + g;
#pragma unroll 2
    for (int qd = warp; qd < nq; qd += GRAM_WARPS) {
      double f[NS];
#pragma unroll
      for (int a = 0; a < NS; ++a) {
        if (a < nb_obs) f[a] = (double)fminf(fmaxf(orow[8 * a], -10.f), 10.f);  // warp-uniform branch
        else f[a] = erow[8 * (a - nb_obs)];
      }
      orow += 4 * GRAM_WARPS * n;
      erow += 4 * GRAM_WARPS * EXS;
      int k = 0;
#pragma unroll
      for (int a = 0; a < NS; ++a)
#pragma unroll
        for (int b = a; b < NS; ++b, ++k) dmma_m8n8k4(acc[k][0], acc[k][1], f[a], f[b]);
    }
    if (ahead && tid == 0) publish(buf ^ 1, tpn);
    __syncthreads();  // tile, ex and sTP[buf] are free again
  }
  // ---- sum the 8 warp partials in a fixed order, write the CTA partial [FP8][FP8] ----
  double* red = reinterpret_cast<double*>(smem_raw);  // [GRAM_WARPS][NTRI][64] (host sizes for both uses)
#pragma unroll
  for (int k = 0; k < NTRI; ++k) {
    double* dst = red + ((size_t)warp * NTRI + k) * 64 + g * 8 + 2 * q;
    dst[0] = acc[k][0]; dst[1] = acc[k][1];
  }
  __syncthreads();
  double* out = partial + (size_t)blockIdx.x * FP8 * FP8;
  for (int i = tid; i < NTRI * 64; i += GRAM_THREADS) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < GRAM_WARPS; ++w) s += red[(size_t)w * NTRI * 64 + i];
    int slot = i >> 6, e = i & 63;
    int a = 0, rem = slot, rowlen = NS;
    while (rem >= rowlen) { rem -= rowlen; ++a; --rowlen; }
    const int b = a + rem;
    out[(8 * a + (e >> 3)) * FP8 + 8 * b + (e & 7)] = s;
  }
}

template <int MAXT>
__global__ void __launch_bounds__(GRAM_THREADS)
linbase_gram_tiles_kernel(const float* __restrict__ obs, const double* __restrict__ ret,
                          const int32_t* __restrict__ off, int P, int64_t N, int n, int TS, int FS,
                          int nside, int ntri, int Q /* > 0: quadratic features */,
                          double* __restrict__ partial) {
  extern __shared__ unsigned char smem_raw[];
  __shared__ int pairs[160];
  if (Q > 0) {  // (i, j) of product q, i <= j row-major
    for (int qq = threadIdx.x; qq < Q; qq += GRAM_THREADS) {
      int i = 0, rem = qq, rowlen = n;
      while (rem >= rowlen) { rem -= rowlen; ++i; --rowlen; }
      pairs[qq] = i | ((i + rem) << 8);
    }
  }
  const int FP8 = 8 * nside;
  double* feat = reinterpret_cast<double*>(smem_raw);  // [TS][FS]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  // tile ownership: contiguous row-major run of the upper triangle -> the A fragment is shared
  const int per = (ntri + GRAM_WARPS - 1) / GRAM_WARPS;
  const int s_begin = warp * per;
  const int nown = max(0, min(per, ntri - s_begin));
  int ti[MAXT], tj[MAXT];
  {
    int a = 0, rem = nown > 0 ? s_begin : 0, rowlen = nside;
    while (rem >= rowlen) { rem -= rowlen; ++a; --rowlen; }
    int b = a + rem;
#pragma unroll
    for (int k = 0; k < MAXT; ++k) {
      ti[k] = a; tj[k] = b;
      if (++b == nside) { ++a; b = a; }
      if (a >= nside) { a = nside - 1; b = nside - 1; }  // past the end: never used (k >= nown)
    }
  }
  unsigned newrow = 1u;       // bit k: tile k starts a new tile row (its A fragment differs from tile k - 1's)
#pragma unroll
  for (int k = 1; k < MAXT; ++k) newrow |= (ti[k] != ti[k - 1]) ? (1u << k) : 0u;
  double acc[MAXT][2];
#pragma unroll
  for (int k = 0; k < MAXT; ++k) acc[k][0] = acc[k][1] = 0.0;

  const int64_t ntiles = (N + TS - 1) / TS;
  const int nq = TS >> 2;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t t0 = tile * TS;
    const int rows = (int)min((int64_t)TS, N - t0);
    __syncthreads();
    if (Q > 0)
      gram_stage_tile_quad(obs, ret, off, P, (double)P / (double)N, t0, rows, n, Q, TS, FP8, FS, pairs, feat, tid);
    else
      gram_stage_tile(obs, ret, off, P, (double)P / (double)N, t0, rows, n, TS, FP8, FS, feat, tid);
    __syncthreads();
    // branch-free inner loop (ncu, obs_dim 111: two thirds of the stall samples sat on the two fragment loads below, issued one
    // at a time behind warp-uniform branches and per-tile index arithmetic; DMMA pipe 24 % busy): all B fragments of a quad of
    // timesteps are requested first, the A fragment is reloaded only where the tile row changes (a predicated load on a
    // precomputed bit), slots k >= nown multiply harmless in-bounds data into accumulators that are never written out.
    for (int qd = 0; qd < nq; ++qd) {
      const double* fr = feat + (4 * qd + q) * FS + g;
      constexpr int CH = MAXT < 16 ? MAXT : 16;      // B fragments in flight at once
      double fa = 0.0;
#pragma unroll
      for (int k0 = 0; k0 < MAXT; k0 += CH) {
        double fb[CH];
#pragma unroll
        for (int k = 0; k < CH; ++k) fb[k] = fr[8 * tj[k0 + k]];
#pragma unroll
        for (int k = 0; k < CH; ++k) {
          if ((newrow >> (k0 + k)) & 1u) fa = fr[8 * ti[k0 + k]];
          dmma_m8n8k4(acc[k0 + k][0], acc[k0 + k][1], fa, fb[k]);
        }
      }
    }
  }
  double* out = partial + (size_t)blockIdx.x * FP8 * FP8;
#pragma unroll
  for (int k = 0; k < MAXT; ++k) {
    if (k < nown) {
      double* dst = out + (size_t)(8 * ti[k] + g) * FP8 + 8 * tj[k] + 2 * q;
      dst[0] = acc[k][0]; dst[1] = acc[k][1];
    }
  }
}

// gram[i*(F+1)+j] = sum_cta S[i][j] (upper) mirrored; gram[F*(F+1)] = y^T y.  One 64-thread CTA per
// entry: thread k adds partials k, k+64, ... then a fixed-order shuffle + 2-slot sum (deterministic).
__global__ void __launch_bounds__(64)
linbase_gram_reduce_kernel(const double* __restrict__ partial, int nparts, int F, int FP,
                           double* __restrict__ gram) {
  __shared__ double sw[2];
  const int F1 = F + 1;
  const int idx = blockIdx.x;
  int i, j;
  if (idx == F * F1) { i = F; j = F; } else { i = idx / F1; j = idx - i * F1; }
  const int a = min(i, j), b = max(i, j);
  // entry (a,b) lives in tile (a/8, b/8) which is stored iff b/8 >= a/8 -- always true
  const double* src = partial + (size_t)a * FP + b;
  double s = 0.0;
  for (int c = threadIdx.x; c < nparts; c += 64) s += src[(size_t)c * FP * FP];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) gram[idx] = sw[0] + sw[1];
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
  const size_t coef_b = sizeof(double) * ((n + 4 + 1) & ~1), tile_b = sizeof(float) * PRED_TS * (size_t)n;
  const int nbuf = coef_b + 2 * tile_b <= (size_t)100 * 1024 ? 2 : 1;  // double-buffer while 2 CTAs/SM fit
  size_t smem = coef_b + nbuf * tile_b;
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE, "linbase_predict: n=%d too large", n);
  if (smem > 40 * 1024)
    cudaFuncSetAttribute(linbase_predict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int per_sm = 1;  // persistent grid = exactly the resident CTAs (a larger grid leaves a half-empty second wave)
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, linbase_predict_kernel, PRED_TS, smem);
  per_sm = per_sm < 1 ? 1 : per_sm;
  int64_t ntiles = (N + PRED_TS - 1) / PRED_TS;
  int grid = (int)(ntiles < (int64_t)ctx->sm_count * per_sm ? ntiles : (int64_t)ctx->sm_count * per_sm);
  linbase_predict_kernel<<<grid, PRED_TS, smem, s>>>(obs, off, P, N, n, nbuf, coeffs, pred);
  return mjrl_check_launch("linbase_predict");
}

// Q = 0: LinearBaseline features (F = n + 4); Q = n(n+1)/2: QuadraticBaseline (F = n + Q + 5)
static int gram_common(mjrl_ctx* ctx, const float* obs, const double* returns, const int32_t* off,
                       int32_t P, int64_t N, int32_t n, int Q, double* gram, void* stream) {
  MJRL_REQUIRE(ctx && obs && returns && off && gram && N > 0 && n > 0 && P > 0, MJRL_E_ARG,
               "baseline gram: bad argument");
  MJRL_REQUIRE(Q <= 160 && n < 256, MJRL_E_SHAPE, "quadbase_gram: obs_dim %d too large", n);
  const int F = Q > 0 ? n + Q + 5 : n + 4, F1 = F + 1, FP = (F1 + 7) & ~7;
  const int nside = FP / 8, ntri = nside * (nside + 1) / 2;
  const int per = (ntri + GRAM_WARPS - 1) / GRAM_WARPS;
  const bool split = Q == 0 && nside <= 3;
  MJRL_REQUIRE(split || per <= 32, MJRL_E_SHAPE,
               "baseline gram: n=%d gives %d Gram tiles (> %d)", n, ntri, 32 * GRAM_WARPS);
  int FS = FP;  // row stride in doubles, = 4 (mod 16): the fragment loads hit 16 distinct bank pairs
  while ((FS & 15) != 4) ++FS;
  const int TS = FS <= 24 ? 256 : (FS <= 52 ? 128 : 64);
  size_t smem = sizeof(double) * (size_t)TS * FS;
  // split variant: two fp32 obs tiles + the fp64 columns (row stride 9 doubles: the per-timestep
  // writes and the fragment reads are both conflict-free; 18 when two 8-wide blocks live there:
  // 2-way); reused for the warp partials
  const int EXS = (nside - n / 8) == 1 ? 9 : 18;
  if (split) {
    smem = (size_t)2 * TS * n * sizeof(float) + sizeof(double) * (size_t)TS * EXS;
    size_t smem_red = sizeof(double) * (size_t)GRAM_WARPS * ntri * 64;
    if (smem_red > smem) smem = smem_red;
  }
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE, "linbase_gram: smem %zu", smem);
  const void* kern = nullptr;
  if (split && nside == 1) kern = (const void*)linbase_gram_split_kernel<1>;
  else if (split && nside == 2) kern = (const void*)linbase_gram_split_kernel<2>;
  else if (split && nside == 3) kern = (const void*)linbase_gram_split_kernel<3>;
  else if (per <= 4) kern = (const void*)linbase_gram_tiles_kernel<4>;
  else if (per <= 8) kern = (const void*)linbase_gram_tiles_kernel<8>;
  else if (per <= 16) kern = (const void*)linbase_gram_tiles_kernel<16>;
  else kern = (const void*)linbase_gram_tiles_kernel<32>;
  if (smem > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int per_sm = 1;  // persistent grid = exactly the resident CTAs
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, GRAM_THREADS, smem);
  per_sm = per_sm < 1 ? 1 : per_sm;
  if (!split && per_sm > 2) per_sm = 2;  // bounds the scratch of the per-CTA partials
  int64_t ntiles = (N + TS - 1) / TS;
  int grid = (int)(ntiles < (int64_t)ctx->sm_count * per_sm ? ntiles : (int64_t)ctx->sm_count * per_sm);
  double* partial = (double*)mjrl_scratch2(ctx, sizeof(double) * (size_t)grid * FP * FP);
  if (!partial) return MJRL_E_STATE;
  cudaStream_t s = (cudaStream_t)stream;
  // partials are written sparsely (upper tiles only); entries never written are never read
#define GRAM_SPLIT(NS)                                                                              \
  do {                                                                                              \
    if (smem > 40 * 1024)                                                                           \
      cudaFuncSetAttribute(linbase_gram_split_kernel<NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                           (int)smem);                                                              \
    linbase_gram_split_kernel<NS><<<grid, GRAM_THREADS, smem, s>>>(obs, returns, off, P, N, n, TS,  \
                                                                   EXS, partial);                   \
  } while (0)
#define GRAM_TILES(MT)                                                                              \
  do {                                                                                              \
    if (smem > 40 * 1024)                                                                           \
      cudaFuncSetAttribute(linbase_gram_tiles_kernel<MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                           (int)smem);                                                              \
    linbase_gram_tiles_kernel<MT><<<grid, GRAM_THREADS, smem, s>>>(obs, returns, off, P, N, n, TS, FS, \
                                                                   nside, ntri, Q, partial);        \
  } while (0)
  if (split && nside == 1) GRAM_SPLIT(1);
  else if (split && nside == 2) GRAM_SPLIT(2);
  else if (split && nside == 3) GRAM_SPLIT(3);
  else if (per <= 4) GRAM_TILES(4);
  else if (per <= 8) GRAM_TILES(8);
  else if (per <= 16) GRAM_TILES(16);
  else GRAM_TILES(32);
#undef GRAM_SPLIT
#undef GRAM_TILES
  int rc = mjrl_check_launch("linbase_gram");
  if (rc) return rc;
  int total = F * F1 + 1;
  linbase_gram_reduce_kernel<<<total, 64, 0, s>>>(partial, grid, F, FP, gram);
  return mjrl_check_launch("linbase_gram_reduce");
}

int mjrl_linbase_gram(mjrl_ctx* ctx, const float* obs, const double* returns, const int32_t* off,
                      int32_t P, int64_t N, int32_t n, double* gram, void* stream) {
  return gram_common(ctx, obs, returns, off, P, N, n, 0, gram, stream);
}

int mjrl_quadbase_gram(mjrl_ctx* ctx, const float* obs, const double* returns, const int32_t* off,
                       int32_t P, int64_t N, int32_t n, double* gram, void* stream) {
  MJRL_REQUIRE(n > 0 && n <= 16, MJRL_E_SHAPE,
               "quadbase_gram: obs_dim %d > 16 (n + n(n+1)/2 + 5 features exceed the solve kernel)", n);
  return gram_common(ctx, obs, returns, off, P, N, n, n * (n + 1) / 2, gram, stream);
}

int mjrl_quadbase_predict(mjrl_ctx* ctx, const float* obs, const int32_t* off, int32_t P, int64_t N,
                          int32_t n, const double* coeffs, double* pred, void* stream) {
  MJRL_REQUIRE(ctx && N >= 0 && n > 0 && n <= 16 && P >= 0, MJRL_E_ARG, "quadbase_predict: bad ctx/N/n/P");
  if (N == 0) return 0;
  MJRL_REQUIRE(pred, MJRL_E_ARG, "quadbase_predict: null pred");
  cudaStream_t s = (cudaStream_t)stream;
  if (!coeffs) {  // never fit: zeros (quadratic_baseline.py:66-67)
    fill_zero_kernel<<<ctx->sm_count * 4, 256, 0, s>>>(pred, N);
    return mjrl_check_launch("quadbase_predict(zero)");
  }
  MJRL_REQUIRE(obs && off, MJRL_E_ARG, "quadbase_predict: null pointer");
  const int F = n + n * (n + 1) / 2 + 5;
  const size_t smem = sizeof(double) * (((F + 1) & ~1) + (size_t)QPRED_TS * (n | 1));
  int64_t ntiles = (N + QPRED_TS - 1) / QPRED_TS;
  int grid = (int)(ntiles < (int64_t)ctx->sm_count * 8 ? ntiles : (int64_t)ctx->sm_count * 8);
  quadbase_predict_kernel<<<grid, QPRED_TS, smem, s>>>(obs, off, P, N, n, F, coeffs, pred);
  return mjrl_check_launch("quadbase_predict");
}

int mjrl_linbase_solve(mjrl_ctx* ctx, const double* gram, int32_t F, double reg_coeff,
                       const double* coeffs_prev, double* coeffs, double* info4, void* stream) {
  MJRL_REQUIRE(ctx && gram && coeffs && info4 && F > 0, MJRL_E_ARG, "linbase_solve: bad argument");
  size_t smem = sizeof(double) * ((size_t)F * (F + 1) + F);
  MJRL_REQUIRE(smem <= (size_t)ctx->max_smem_optin, MJRL_E_SHAPE,
               "linbase_solve: F=%d needs %zu B of shared memory", F, smem);
  if (smem > 40 * 1024)
    cudaFuncSetAttribute(linbase_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  linbase_solve_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(gram, F, reg_coeff, coeffs_prev, coeffs, info4);
  return mjrl_check_launch("linbase_solve");
}

}  // extern "C"
