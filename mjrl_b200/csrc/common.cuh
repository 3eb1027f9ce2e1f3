// Shared helpers for the mjrl_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mjrl_b200.h"

struct mjrl_ctx {
  int device;
  int sm_count;
  int max_smem_optin;
  // scratch for per-CTA partial sums (fp64 / fp32 views of the same block)
  void* scratch;
  size_t scratch_bytes;
  // second scratch (Gram partials can be large)
  void* scratch2;
  size_t scratch2_bytes;
};

void mjrl_set_error(const char* fmt, ...);
int mjrl_check_launch(const char* what);
// grow-on-demand scratch; returns nullptr (and sets error) on failure
void* mjrl_scratch(mjrl_ctx* ctx, size_t bytes);
void* mjrl_scratch2(mjrl_ctx* ctx, size_t bytes);

#define MJRL_REQUIRE(cond, code, ...) \
  do {                                \
    if (!(cond)) {                    \
      mjrl_set_error(__VA_ARGS__);    \
      return (code);                  \
    }                                 \
  } while (0)

// ---- device helpers -------------------------------------------------------

// index p of the path containing timestep i: offsets[p] <= i < offsets[p+1].
// Empty paths (offsets[p] == offsets[p+1]) are skipped by construction.
__device__ __forceinline__ int find_path(const int32_t* __restrict__ offsets, int P, int64_t i) {
  int lo = 0, hi = P;  // invariant: offsets[lo] <= i < offsets[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if ((int64_t)__ldg(offsets + mid) <= i) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide fp64 sum, fixed order (deterministic).  `red` is >= 32 doubles of smem.
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    r = (l < nw) ? red[l] : 0.0;
    r = warp_sum(r);
  }
  return r;  // valid in warp 0
}

// streaming loads/stores that do not pollute L1
__device__ __forceinline__ double ld_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_stream(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
// Same loads without `volatile`: the compiler may hoist and batch them, so a thread keeps several
// requests in flight (only for buffers no thread writes during the kernel).
__device__ __forceinline__ double ldg_stream(const double* p) {
  double v;
  asm("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ldg_stream(const float* p) {
  float v;
  asm("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 v;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
      : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// Coalesced read of `cnt` consecutive floats viewed as rows of `n`: put(row, col, value) is called
// once per element.  16-byte loads, four per thread in flight before the first use; (row, col) of a
// thread's next 16 bytes advance by loop-invariant constants, so there is no division in the loop.
// NT = threads per CTA.
template <int NT, class Put>
__device__ __forceinline__ void stream_rows(const float* __restrict__ src, int cnt, int n, int tid,
                                            Put&& put) {
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    const int nv = cnt >> 2;
    const int dq = (4 * NT) / n, dr = 4 * NT - dq * n;
    int r = (4 * tid) / n, c = 4 * tid - r * n;
    auto put4 = [&](const float4& v) {  // consumes (r, c), then advances them by 4*NT elements
      int rr = r, cc = c;
      put(rr, cc, v.x); if (++cc == n) { cc = 0; ++rr; }
      put(rr, cc, v.y); if (++cc == n) { cc = 0; ++rr; }
      put(rr, cc, v.z); if (++cc == n) { cc = 0; ++rr; }
      put(rr, cc, v.w);
      r += dq; c += dr;
      if (c >= n) { c -= n; ++r; }
    };
    int i = tid;
    for (; i + 3 * NT < nv; i += 4 * NT) {
      const float4 v0 = ldg_stream(s4 + i), v1 = ldg_stream(s4 + i + NT);
      const float4 v2 = ldg_stream(s4 + i + 2 * NT), v3 = ldg_stream(s4 + i + 3 * NT);
      put4(v0); put4(v1); put4(v2); put4(v3);
    }
    float4 v[3];
    int k = 0;
#pragma unroll
    for (int j = 0; j < 3; ++j)
      if (i + j * NT < nv) { v[j] = ldg_stream(s4 + i + j * NT); k = j + 1; }
#pragma unroll
    for (int j = 0; j < 3; ++j)
      if (j < k) put4(v[j]);
    for (int e = (nv << 2) + tid; e < cnt; e += NT) { int rr = e / n; put(rr, e - rr * n, ldg_stream(src + e)); }
  } else {
#pragma unroll 4
    for (int e = tid; e < cnt; e += NT) { int rr = e / n; put(rr, e - rr * n, ldg_stream(src + e)); }
  }
}

// ---- path lookup for a tile of consecutive timesteps -----------------------------------------
// All 32 lanes of a warp search together (32 probes per round: 3 rounds for 16k paths instead of
// 14 dependent loads); every lane returns p with off[p] <= t < off[p+1].  Call with a warp-uniform t.
__device__ __forceinline__ int find_path_warp(const int32_t* __restrict__ off, int P, int64_t t) {
  const int lane = threadIdx.x & 31;
  int lo = 0, hi = P;
  while (hi - lo > 1) {
    const int step = (hi - lo + 31) >> 5;
    const int probe = lo + (lane + 1) * step;
    const bool le = probe < hi && (int64_t)__ldg(off + probe) <= t;
    const int cnt = __popc(__ballot_sync(0xffffffffu, le));
    lo += cnt * step;
    hi = min(hi, lo + step);
  }
  return lo;
}
struct TilePath { int p0, s0, s1; };  // path holding the tile's first timestep, its start, the next start
// Interpolation guess first: with near-equal path lengths p = t0*P/N is right (or off by a few), so
// the common case is ONE round of two independent loads; otherwise the warp searches the side
// the guess points to.  `inv` = P / N.  All 32 lanes call it with the same arguments.
__device__ __forceinline__ TilePath tile_path(const int32_t* __restrict__ off, int P, int64_t t0,
                                              double inv) {
  TilePath tp;
  int p = (int)((double)t0 * inv);
  p = max(0, min(P - 1, p));
  int s0 = __ldg(off + p), s1 = __ldg(off + p + 1);
  if (!(s0 <= t0 && t0 < s1)) {
    p = find_path_warp(off, P, t0);
    s0 = __ldg(off + p); s1 = __ldg(off + p + 1);
  }
  tp.p0 = p; tp.s0 = s0; tp.s1 = s1;
  return tp;
}
// index of timestep t >= t0 inside its path: no search unless a path boundary precedes t in the tile
__device__ __forceinline__ int t_in_path(const int32_t* __restrict__ off, int P, const TilePath& tp,
                                         int64_t t) {
  if (t < tp.s1) return (int)(t - tp.s0);
  int lo = tp.p0 + 1, step = 1;  // off[lo] <= t; gallop, then bisect
  while (lo + step < P && (int64_t)__ldg(off + lo + step) <= t) { lo += step; step <<= 1; }
  int hi = min(P, lo + step);
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((int64_t)__ldg(off + mid) <= t) lo = mid; else hi = mid;
  }
  return (int)(t - __ldg(off + lo));
}

// ---- bulk global -> shared copies (TMA, 1-D) completing on an mbarrier -------------------------
namespace tma {
__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(saddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// one arrival + `bytes` expected from bulk copies
__device__ __forceinline__ void mbar_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(saddr(bar)), "r"(bytes) : "memory");
}
// size and both addresses multiples of 16 bytes
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(saddr(dst)), "l"(src), "r"(bytes), "r"(saddr(bar)) : "memory");
}
// bounded wait: a protocol bug traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = saddr(bar);
  uint32_t ok = 0;
  for (long long it = 0; !ok; ++it) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (it > (1ll << 26)) __trap();
  }
}
}  // namespace tma
