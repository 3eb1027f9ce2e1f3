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
