// Micro-probe: per-SM throughput of legacy mma.sync TF32 (m16n8k8) vs FFMA on sm_100a.
// Decides whether a 3xTF32 mma.sync path can beat the fp32-FMA kernels (DESIGN.md "FVP roadmap").
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) mma_tf32(float* out, int iters) {
  float c[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 * 3, b1 = a0 * 5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) mma_bf16(float* out, int iters) {
  float c[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 * 3, b1 = a0 * 5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) ffma(float* out, int iters, float x, float y) {
  float c[16];
  for (int i = 0; i < 16; ++i) c[i] = threadIdx.x * 0.001f + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = fmaf(c[i], x, y);
  }
  float s = 0;
  for (int i = 0; i < 16; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount;
  float* out; cudaMalloc(&out, sizeof(float) * sms * 8 * 256);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int warps_cfg = 0; warps_cfg < 3; ++warps_cfg) {
    int ctas = sms * (warps_cfg == 0 ? 1 : warps_cfg == 1 ? 2 : 4);
    int iters = 20000;
    float ms;
    mma_tf32<<<ctas, 256>>>(out, 100);
    cudaEventRecord(e0); mma_tf32<<<ctas, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double fl = 2.0 * 16 * 8 * 8 * 8.0 * iters * 8 * ctas;
    printf("mma.sync tf32 m16n8k8 : %d CTAs x 8 warps: %.3f ms  %.1f TFLOP/s  (%.1f cyc/mma/SMSP at %d MHz)\n", ctas, ms,
           fl / ms / 1e9, ms * 1e-3 * p.clockRate * 1e3 / (8.0 * iters * 8 * ctas / sms / 4), p.clockRate / 1000);
    mma_bf16<<<ctas, 256>>>(out, 100);
    cudaEventRecord(e0); mma_bf16<<<ctas, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    fl = 2.0 * 16 * 8 * 16 * 8.0 * iters * 8 * ctas;
    printf("mma.sync bf16 m16n8k16: %d CTAs x 8 warps: %.3f ms  %.1f TFLOP/s\n", ctas, ms, fl / ms / 1e9);
    ffma<<<ctas, 256>>>(out, 100, 1.0001f, 0.5f);
    cudaEventRecord(e0); ffma<<<ctas, 256>>>(out, iters * 4, 1.0001f, 0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    fl = 2.0 * 16.0 * iters * 4 * 256 * ctas;
    printf("ffma                  : %d CTAs x 8 warps: %.3f ms  %.1f TFLOP/s\n", ctas, ms, fl / ms / 1e9);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
