// Bring-up probe for the tcgen05 layer in mjrl_b200/csrc/umma.cuh (run on the B200 box):
//   case 0: D[128][N] = A[128][K] * B[N][K]^T, A and B K-major views of [chunks][rows][4] buffers
//   case 1: D[M=128 cols of A][N cols of B] = sum_t A[t][.] * B[t][.]  (both MN-major, K = rows)
//   case 2: like 0 with A staged in tensor memory (tcgen05.st) instead of shared memory
// Each with 3xTF32 error compensation, checked against fp64 on the host; then a throughput loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I mjrl_b200/csrc profiles/probes/umma_probe.cu -o gpurun_out/umma_probe
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "umma.cuh"

using namespace umma;

struct P {
  const float* A; const float* B; float* D; long long* cyc;
  int mode, N, K, reps, variant, batches;
};

// A: [128][K] row-major (mode 0/2) or [K][128] (mode 1: rows = k).  B: [N][K] or [K][N].
__global__ void __launch_bounds__(128, 1) probe(P p) {
  extern __shared__ __align__(1024) unsigned char smraw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = p.N, K = p.K;
  // buffers: Ahi, Alo, Bhi, Blo
  const int Arows = p.mode == 1 ? K : 128, Acols = p.mode == 1 ? 128 : K;
  const int Brows = p.mode == 1 ? K : N, Bcols = p.mode == 1 ? N : K;
  float* Ahi = (float*)smraw; float* Alo = Ahi + Arows * Acols;
  float* Bhi = Alo + Arows * Acols; float* Blo = Bhi + Brows * Bcols;
  for (int i = tid; i < Arows * Acols; i += 128) {
    int r = i / Acols, c = i % Acols;
    float hi, lo; split_tf32(p.A[i], hi, lo);
    int o = (c >> 2) * Arows * 4 + r * 4 + (c & 3);
    Ahi[o] = hi; Alo[o] = lo;
  }
  for (int i = tid; i < Brows * Bcols; i += 128) {
    int r = i / Bcols, c = i % Bcols;
    float hi, lo; split_tf32(p.B[i], hi, lo);
    int o = (c >> 2) * Brows * 4 + r * 4 + (c & 3);
    Bhi[o] = hi; Blo[o] = lo;
  }
  if (warp == 0) tmem_alloc<512>(&tslot);
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tb = tslot;
  const uint32_t lane_base = (uint32_t)(32 * (warp & 3)) << 16;
  if (p.mode == 2) {  // stage A hi/lo into TMEM columns [256, 256+K) and [384, 384+K)
    for (int c0 = 0; c0 < K; c0 += 8) {
      uint32_t h[8], l[8];
      for (int j = 0; j < 8; ++j) {
        float hi, lo; split_tf32(p.A[tid * K + c0 + j], hi, lo);
        h[j] = __float_as_uint(hi); l[j] = __float_as_uint(lo);
      }
      tmem_st8(tb + lane_base + 256 + c0, h);
      tmem_st8(tb + lane_base + 384 + c0, l);
    }
    tmem_wait_st();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
  }
  const uint32_t aH = smem_u32(Ahi), aL = smem_u32(Alo), bH = smem_u32(Bhi), bL = smem_u32(Blo);
  long long t0 = 0, t1 = 0;
  uint32_t parity = 0;
  for (int rep = 0; rep < p.reps; ++rep) {
    if (tid == 0) {
      t0 = clock64();
      for (int bt = 0; bt < p.batches; ++bt)
      if (p.mode == 0 || p.mode == 2) {
        const uint32_t id = idesc_tf32(128, N, 0, 0);
        for (int k0 = 0; k0 < K; k0 += 8) {
          const int kc = k0 >> 2;
          uint64_t dbh = desc_kmajor(bH, N, 0, kc), dbl = desc_kmajor(bL, N, 0, kc);
          if (p.mode == 0) {
            uint64_t dah = desc_kmajor(aH, 128, 0, kc), dal = desc_kmajor(aL, 128, 0, kc);
            mma_ss(tb, dal, dbh, id, (k0 | bt) > 0);
            mma_ss(tb, dah, dbl, id, 1);
            mma_ss(tb, dah, dbh, id, 1);
          } else {
            mma_ts(tb, tb + 384 + k0, dbh, id, (k0 | bt) > 0);
            mma_ts(tb, tb + 256 + k0, dbl, id, 1);
            mma_ts(tb, tb + 256 + k0, dbh, id, 1);
          }
        }
      } else {
        const uint32_t id = idesc_tf32(128, N, 1, 1);
        for (int k0 = 0; k0 < K; k0 += 8) {
          uint64_t dah, dal, dbh, dbl;
          if (p.variant == 0) {
            dah = desc_mnmajor(aH, K, k0, 0); dal = desc_mnmajor(aL, K, k0, 0);
            dbh = desc_mnmajor(bH, K, k0, 0); dbl = desc_mnmajor(bL, K, k0, 0);
          } else {  // LBO <-> SBO swapped
            dah = smem_desc(aH + k0 * 16, K * 16, 128); dal = smem_desc(aL + k0 * 16, K * 16, 128);
            dbh = smem_desc(bH + k0 * 16, K * 16, 128); dbl = smem_desc(bL + k0 * 16, K * 16, 128);
          }
          mma_ss(tb, dal, dbh, id, (k0 | bt) > 0);
          mma_ss(tb, dah, dbl, id, 1);
          mma_ss(tb, dah, dbh, id, 1);
        }
      }
      commit(&bar);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    if (tid == 0) { t1 = clock64(); if (rep == p.reps - 1) p.cyc[0] = t1 - t0; }
    fence_after_sync();
  }
  // read back D: thread = row
  for (int c0 = 0; c0 < N; c0 += 8) {
    float v[8];
    tmem_ld8(tb + lane_base + c0, v);
    tmem_wait_ld();
    for (int j = 0; j < 8; ++j) p.D[tid * N + c0 + j] = v[j];
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}

static double run(int mode, int N, int K, int reps, int variant = 0, int batches = 1) {
  const int Arows = mode == 1 ? K : 128, Acols = mode == 1 ? 128 : K;
  const int Brows = mode == 1 ? K : N, Bcols = mode == 1 ? N : K;
  std::vector<float> A(Arows * Acols), B(Brows * Bcols), D(128 * N);
  srand(1234 + mode);
  for (auto& x : A) x = (float)rand() / RAND_MAX * 2.f - 1.f;
  for (auto& x : B) x = (float)rand() / RAND_MAX * 2.f - 1.f;
  float *dA, *dB, *dD; long long* dC;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
  cudaMalloc(&dC, 8);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, D.size() * 4);
  size_t smem = 2 * 4 * (A.size() + B.size()) + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  P p{dA, dB, dD, dC, mode, N, K, reps, variant, batches};
  probe<<<1, 128, smem>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("mode %d N %d K %d: CUDA error %s\n", mode, N, K, cudaGetErrorString(e)); exit(1); }
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  long long cyc; cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost);
  double num = 0, den = 0;
  for (int i = 0; i < 128; ++i)
    for (int j = 0; j < N; ++j) {
      double s = 0;
      for (int k = 0; k < K; ++k) {
        double a = mode == 1 ? A[k * 128 + i] : A[i * K + k];
        double b = mode == 1 ? B[k * N + j] : B[j * K + k];
        s += a * b;
      }
      s *= batches;
      if (i < 2 && j < 4 && variant >= 0 && mode == 1) printf("   D[%d][%d] = %g expected %g\n", i, j, D[i * N + j], s);
      num += (s - D[i * N + j]) * (s - D[i * N + j]); den += s * s;
    }
  double rel = sqrt(num / den);
  printf("mode %d v%d b%d N %3d  K %3d : rel err %.3e   last batch %lld cycles for %d MMAs (%.1f cyc/MMA)\n", mode, variant, batches, N, K,
         rel, cyc, batches * 3 * K / 8, (double)cyc / (batches * 3 * K / 8));
  cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dC);
  return rel;
}

int main() {
  int bad = 0;
  const int cases[][3] = {{0, 64, 64}, {0, 128, 16}, {0, 16, 64}, {0, 72, 64}, {0, 32, 32},
                          {1, 64, 128}, {1, 72, 128}, {1, 16, 128}, {2, 64, 64}, {2, 16, 64}, {2, 128, 64}};
  for (auto& c : cases) bad += run(c[0], c[1], c[2], 3) > 1e-5;
  run(1, 64, 128, 1, 1);
  // throughput: long batches (accumulating; results scaled by `batches`)
  run(0, 64, 64, 2, 0, 16); run(0, 16, 64, 2, 0, 16); run(0, 128, 64, 2, 0, 16); run(0, 256, 64, 2, 0, 16);
  run(2, 64, 64, 2, 0, 16); run(2, 16, 64, 2, 0, 16); run(2, 128, 64, 2, 0, 16);
  run(1, 64, 128, 2, 0, 8); run(1, 64, 128, 2, 1, 8);
  printf(bad ? "PROBE FAILED (%d cases)\n" : "PROBE OK\n", bad);
  return bad;
}
