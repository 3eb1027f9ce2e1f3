// Round-2 bring-up probe for the tcgen05 path of the Fisher-vector product (csrc/fvp_tc.cu):
// swizzled (32/64/128 B) shared-memory operands in both majors, A from tensor memory, M = 64 / 128,
// 3xTF32 with the hi plane left as RAW fp32 (does the tensor core truncate?), and the issue rate of
// kind::tf32 for the shapes the kernel uses.  Every case is checked against fp64 on the host.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I mjrl_b200/csrc profiles/probes/umma2_probe.cu -o gpurun_out/umma2_probe
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- swizzled plane: R rows x C fp32 columns in blocks of S/4 columns; block = R rows of S bytes -------
struct Plane { uint32_t base; int R, S; };   // S = 32, 64, 128 (bytes per row inside a block)
// S = 129 selects 128-byte rows with the 32-byte-atom swizzle (SWIZZLE_128B_BASE32B, Swizzle<2,5,2> on byte addresses):
// the only layout kind::tf32 takes for MN-major operands.
__host__ __device__ inline int row_bytes(int S) { return S == 129 ? 128 : S; }
__host__ __device__ inline uint32_t plane_off(int R, int S, int r, int c) {
  const int rb = row_bytes(S), cb = rb / 4, blk = c / cb, cc = c % cb;
  uint32_t off = (uint32_t)blk * R * rb + (uint32_t)r * rb + cc * 4;
  if (S == 129) off ^= ((off >> 7) & 3u) << 5;
  else off ^= ((off >> 7) & (uint32_t)(S / 16 - 1)) << 4;
  return off;
}
__device__ __forceinline__ uint64_t desc_bits(uint32_t start, uint32_t lbo, uint32_t sbo, int S) {
  const uint64_t lt = S == 129 ? 1 : (S == 128 ? 2 : (S == 64 ? 4 : (S == 32 ? 6 : 0)));
  uint64_t d = (uint64_t)((start >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= lt << 61;
  return d;
}
// K-major view: MMA rows = plane rows (from r0), K = plane columns (from k0, 8 per instruction)
__device__ __forceinline__ uint64_t desc_k(const Plane& p, int r0, int k0) {
  const int cb = p.S / 4;
  return desc_bits(p.base + (k0 / cb) * p.R * p.S + r0 * p.S + (k0 % cb) * 4, 16u, 8u * p.S, p.S);
}
// MN-major view: MMA rows = plane columns (from c0), K = plane rows (from r0, 8 per instruction)
__device__ __forceinline__ uint64_t desc_mn(const Plane& p, int c0, int r0) {
  const int rb = row_bytes(p.S), cb = rb / 4;
  // LBO: next block of 32 MN values; SBO: next K atom (4 rows with the 32-byte-atom swizzle, else 8)
  return desc_bits(p.base + (c0 / cb) * p.R * rb + r0 * rb, (uint32_t)p.R * rb, (p.S == 129 ? 4u : 8u) * rb, p.S);
}
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
               :: "r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
               :: "r"(d), "r"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = smem_u32(bar);
  uint32_t ok = 0;
  for (long long it = 0; !ok; ++it) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (it > (1ll << 24)) __trap();
  }
}
__device__ __forceinline__ void tld8(uint32_t t, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(t) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tst8(uint32_t t, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "r"(t), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}

__device__ __forceinline__ uint32_t elect_one_sync() {
  uint32_t pred = 0, laneid = 0;
  asm volatile("{\n\t.reg .b32 %%rx;\n\t.reg .pred %%px;\n\telect.sync %%rx|%%px, %2;\n\t@%%px mov.s32 %1, 1;\n\tmov.s32 %0, %%rx;\n\t}\n"
               : "+r"(laneid), "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred;
}

struct P {
  const float* A; const float* B; float* D; long long* cyc;
  int M, N, K;        // MMA M (64/128), N, total K
  int amode;          // 0 = A smem K-major, 1 = A smem MN-major, 2 = A in TMEM
  int bmn;            // 0 = B K-major ([N][K] plane), 1 = B MN-major ([K][N] plane)
  int S;              // swizzle kind of the A plane
  int SB;             // swizzle kind of the B plane
  int round_hi;       // 1: hi plane rounded to tf32 (rna); 0: raw fp32 (relies on hardware truncation)
  int terms;          // 3 = 3xTF32, 1 = plain
  int timing_mmas;    // > 0: after the numerics pass, issue this many MMAs round-robin over `nacc` accumulators
  int nacc;
};

__device__ __forceinline__ void split(float x, int round_hi, float& hi, float& lo) {
  uint32_t h;
  if (round_hi) asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  else h = __float_as_uint(x) & 0xffffe000u;
  lo = x - __uint_as_float(h);
  hi = round_hi ? __uint_as_float(h) : x;      // raw: the low 13 bits stay in the word
}

__global__ void __launch_bounds__(128, 1) probe(P p) {
  extern __shared__ __align__(1024) unsigned char smraw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t sbase = (smem_u32(smraw) + 1023u) & ~1023u;
  unsigned char* sm = smraw + (sbase - smem_u32(smraw));
  const int M = p.M, N = p.N, K = p.K, S = p.S, SB = p.SB;
  // A plane: amode 0: rows = M, cols = K;  amode 1: rows = K, cols = M.   B plane: bmn 0: [N][K]; 1: [K][N].
  const int aR = p.amode == 1 ? K : M, aC = p.amode == 1 ? M : K;
  const int bR = p.bmn ? K : N, bC = p.bmn ? N : K;
  auto plane_bytes = [&](int R, int C, int SS) { const int rb = row_bytes(SS); return (uint32_t)((C + rb / 4 - 1) / (rb / 4)) * R * rb; };
  const uint32_t aB = (plane_bytes(aR, aC, S) + 1023u) & ~1023u, bB = (plane_bytes(bR, bC, SB) + 1023u) & ~1023u;
  Plane Ah{sbase, aR, S}, Al{sbase + aB, aR, S}, Bh{sbase + 2 * aB, bR, SB}, Bl{sbase + 2 * aB + bB, bR, SB};
  for (uint32_t i = tid; i < (2 * aB + 2 * bB) / 4; i += 128) ((float*)sm)[i] = 0.f;
  __syncthreads();
  if (p.amode != 2)
    for (int i = tid; i < aR * aC; i += 128) {
      const int r = i / aC, c = i % aC;
      float hi, lo; split(p.A[i], p.round_hi, hi, lo);
      *(float*)(sm + plane_off(aR, S, r, c)) = hi;
      *(float*)(sm + aB + plane_off(aR, S, r, c)) = lo;
    }
  for (int i = tid; i < bR * bC; i += 128) {
    const int r = i / bC, c = i % bC;
    float hi, lo; split(p.B[i], p.round_hi, hi, lo);
    *(float*)(sm + 2 * aB + plane_off(bR, SB, r, c)) = hi;
    *(float*)(sm + 2 * aB + bB + plane_off(bR, SB, r, c)) = lo;
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tslot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tslot;
  const uint32_t lane_base = (uint32_t)(32 * warp) << 16;
  const int AH = 256, AL = 384;    // TMEM columns of the A planes (amode 2)
  if (p.amode == 2) {
    for (int c0 = 0; c0 < K; c0 += 8) {
      uint32_t h[8], l[8];
      for (int j = 0; j < 8; ++j) {
        float hi, lo; split(p.A[tid * K + c0 + j], p.round_hi, hi, lo);
        h[j] = __float_as_uint(hi); l[j] = __float_as_uint(lo);
      }
      tst8(tb + lane_base + AH + c0, h);
      tst8(tb + lane_base + AL + c0, l);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  const uint32_t id = idesc_tf32(M, N, p.amode == 1, p.bmn);
  uint32_t parity = 0;
  auto issue_step = [&](uint32_t dcol, int k0, int term, uint32_t acc) {
    // term 0: hi*hi, 1: lo*hi, 2: hi*lo
    const Plane& bp = term == 2 ? Bl : Bh;
    const uint64_t db = p.bmn ? desc_mn(bp, 0, k0) : desc_k(bp, 0, k0);
    if (p.amode == 2) {
      mma_ts(tb + dcol, tb + (term == 1 ? AL : AH) + k0, db, id, acc);
    } else {
      const Plane& ap = term == 1 ? Al : Ah;
      const uint64_t da = p.amode == 1 ? desc_mn(ap, 0, k0) : desc_k(ap, 0, k0);
      mma_ss(tb + dcol, da, db, id, acc);
    }
  };
  if (tid == 0) {
    for (int k0 = 0; k0 < K; k0 += 8)
      for (int term = 0; term < p.terms; ++term) issue_step(0, k0, term, (k0 | term) > 0);
    commit(&bar);
  }
  mbar_wait(&bar, parity); parity ^= 1;
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // dump all 128 lanes x N columns (for M = 64 this shows where the rows land)
  for (int c0 = 0; c0 < N; c0 += 8) {
    uint32_t v[8];
    tld8(tb + lane_base + c0, v);
    for (int j = 0; j < 8; ++j) p.D[tid * N + c0 + j] = __uint_as_float(v[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (p.timing_mmas > 0) {
    long long t0 = 0;
    if (warp == 0) {          // warp-uniform role, one elected lane issues (no per-instruction elect loop in SASS)
      const uint64_t db0 = p.bmn ? desc_mn(Bh, 0, 0) : desc_k(Bh, 0, 0), db1 = p.bmn ? desc_mn(Bl, 0, 0) : desc_k(Bl, 0, 0);
      const uint64_t da0 = p.amode == 1 ? desc_mn(Ah, 0, 0) : desc_k(Ah, 0, 0), da1 = p.amode == 1 ? desc_mn(Al, 0, 0) : desc_k(Al, 0, 0);
      const uint32_t nacc = p.nacc, dstep = (uint32_t)N;
      const int amode = p.amode, n = p.timing_mmas;
      t0 = clock64();
      if (elect_one_sync()) {
        uint32_t dcol = tb;
#pragma unroll 1
        for (int i = 0; i < n; i += 3) {
          if (amode == 2) {
            mma_ts(dcol, tb + AH, db0, id, 1); mma_ts(dcol, tb + AL, db0, id, 1); mma_ts(dcol, tb + AH, db1, id, 1);
          } else {
            mma_ss(dcol, da0, db0, id, 1); mma_ss(dcol, da1, db0, id, 1); mma_ss(dcol, da0, db1, id, 1);
          }
          dcol += dstep; if (dcol >= tb + nacc * dstep) dcol = tb;
        }
        p.cyc[1] = clock64() - t0;    // issue-side time
        commit(&bar);
      }
      __syncwarp();
    }
    mbar_wait(&bar, parity); parity ^= 1;
    if (tid == 0) p.cyc[0] = clock64() - t0;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tb) : "memory");
}

static int run(const char* name, int M, int N, int K, int amode, int bmn, int S, int round_hi, int terms,
               int timing = 0, int nacc = 1, bool show = false, int SB = 0) {
  if (SB == 0) SB = S;
  const int aR = amode == 1 ? K : M, aC = amode == 1 ? M : K, bR = bmn ? K : N, bC = bmn ? N : K;
  std::vector<float> A(aR * aC), B(bR * bC), D(128 * N);
  srand(99 + M + N + K + amode * 7 + bmn * 3 + S);
  for (auto& x : A) x = ((float)rand() / RAND_MAX * 2.f - 1.f) * (1.f + (rand() % 5));
  for (auto& x : B) x = (float)rand() / RAND_MAX * 2.f - 1.f;
  float *dA, *dB, *dD; long long* dC;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4); cudaMalloc(&dC, 16);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, D.size() * 4); cudaMemset(dC, 0, 16);
  auto pb = [&](int R, int C, int SS) { const int rb = row_bytes(SS); return (size_t)((((C + rb / 4 - 1) / (rb / 4)) * R * rb + 1023) & ~1023); };
  size_t smem = 2 * pb(aR, aC, S) + 2 * pb(bR, bC, SB) + 2048;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  P p{dA, dB, dD, dC, M, N, K, amode, bmn, S, SB, round_hi, terms, timing, nacc};
  probe<<<1, 128, smem>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); exit(2); }
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  long long cyc[2]; cudaMemcpy(cyc, dC, 16, cudaMemcpyDeviceToHost);
  // expected: E[i][j] = sum_k A(i,k) B(j,k)
  std::vector<double> E(M * N);
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      double s = 0;
      for (int k = 0; k < K; ++k) {
        const double a = amode == 1 ? A[k * M + i] : A[i * K + k];
        const double b = bmn ? B[k * N + j] : B[j * K + k];
        s += a * b;
      }
      E[i * N + j] = s;
    }
  // candidate row -> lane maps: identity (M = 128); for M = 64 try lane = r, lane = r + 16*(r/16) (16 rows per 32-lane quadrant)
  double best = 1e30; int bestmap = -1;
  for (int map = 0; map < 3; ++map) {
    double num = 0, den = 0;
    for (int i = 0; i < M; ++i) {
      int lane = map == 0 ? i : (map == 1 ? (i % 16) + 32 * (i / 16) : (i % 32) + 64 * (i / 32));
      if (lane >= 128) { num = 1e30; break; }
      for (int j = 0; j < N; ++j) { double d = E[i * N + j] - D[lane * N + j]; num += d * d; den += E[i * N + j] * E[i * N + j]; }
    }
    double rel = sqrt(num / den);
    if (rel < best) { best = rel; bestmap = map; }
  }
  printf("%-34s M%3d N%3d K%3d S%3d %s terms %d: rel err %.3e (row map %d)", name, M, N, K, S, round_hi ? "rna" : "raw", terms, best, bestmap);
  if (timing) printf("   %d MMAs over %d acc: %.1f cyc/MMA total, %.1f issue", timing, nacc, (double)cyc[0] / timing, (double)cyc[1] / timing);
  printf("\n");
  if (show) for (int i = 0; i < 2; ++i) printf("    D[%d][0..3] = %g %g %g %g   expected %g %g %g %g\n", i, D[i * N], D[i * N + 1], D[i * N + 2], D[i * N + 3], E[i * N], E[i * N + 1], E[i * N + 2], E[i * N + 3]);
  cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dC);
  const double tol = terms == 3 ? 3e-6 : 2e-3;
  return best > tol;
}

int main() {
  int bad = 0;
  // numerics: K-major SS with each swizzle, raw vs rounded hi plane
  bad += run("SS kmajor", 128, 64, 64, 0, 0, 128, 1, 3, 0, 1, true);
  bad += run("SS kmajor raw-hi", 128, 64, 64, 0, 0, 128, 0, 3);
  bad += run("SS kmajor 1 term raw", 128, 64, 64, 0, 0, 128, 0, 1);
  bad += run("SS kmajor 1 term rna", 128, 64, 64, 0, 0, 128, 1, 1);
  bad += run("SS kmajor sw64 K16", 128, 64, 16, 0, 0, 64, 0, 3);
  bad += run("SS kmajor sw32 K8", 128, 64, 8, 0, 0, 32, 0, 3);
  bad += run("SS kmajor N128", 128, 128, 64, 0, 0, 128, 0, 3);
  bad += run("SS kmajor N8", 128, 8, 64, 0, 0, 128, 0, 3);
  bad += run("SS kmajor N16", 128, 16, 64, 0, 0, 128, 0, 3);
  bad += run("SS kmajor N256 K32", 128, 256, 32, 0, 0, 128, 0, 3);
  // A from TMEM
  bad += run("TS kmajor", 128, 64, 64, 2, 0, 128, 0, 3);
  bad += run("TS kmajor N128", 128, 128, 64, 2, 0, 128, 0, 3);
  bad += run("TS B mn-major (32B-atom swizzle)", 128, 64, 64, 2, 1, 129, 0, 3);
  // weight gradients: A MN-major (K = rows = timesteps), B MN-major or K-major (transposed small operand)
  bad += run("SS mn/mn (wgrad) M128", 128, 64, 96, 1, 1, 129, 0, 3, 0, 1, true);
  bad += run("SS mn/mn (wgrad) M128 N128", 128, 128, 64, 1, 1, 129, 0, 3);
  bad += run("SS mn/mn (wgrad) M64", 64, 64, 96, 1, 1, 129, 0, 3);
  bad += run("SS A mn / B kmajor N32 K96", 128, 32, 96, 1, 0, 129, 0, 3, 0, 1, false, 128);
  bad += run("SS A mn / B kmajor N16 K96", 128, 16, 96, 1, 0, 129, 0, 3, 0, 1, false, 128);
  bad += run("SS kmajor M64", 64, 64, 64, 0, 0, 128, 0, 3);
  // issue rates (descriptors precomputed; 3 MMAs per accumulator visit)
  const int T = 960;
  run("rate SS k/k N64", 128, 64, 64, 0, 0, 128, 0, 3, T, 4);
  run("rate SS k/k N128", 128, 128, 64, 0, 0, 128, 0, 3, T, 2);
  run("rate SS k/k N256", 128, 256, 32, 0, 0, 128, 0, 3, T, 1);
  run("rate SS k/k N16", 128, 16, 64, 0, 0, 128, 0, 3, T, 4);
  run("rate SS k/k N8", 128, 8, 64, 0, 0, 128, 0, 3, T, 4);
  run("rate SS k/k N64 one acc", 128, 64, 64, 0, 0, 128, 0, 3, T, 1);
  run("rate TS N64", 128, 64, 64, 2, 0, 128, 0, 3, T, 2);
  run("rate TS N64 one acc", 128, 64, 64, 2, 0, 128, 0, 3, T, 1);
  run("rate TS N128", 128, 128, 64, 2, 0, 128, 0, 3, T, 1);
  run("rate TS N32", 128, 32, 64, 2, 0, 128, 0, 3, T, 4);
  run("rate TS N16", 128, 16, 64, 2, 0, 128, 0, 3, T, 4);
  run("rate TS N8", 128, 8, 64, 2, 0, 128, 0, 3, T, 4);
  run("rate TS B-mn N64", 128, 64, 64, 2, 1, 129, 0, 3, T, 2);
  run("rate SS mn/mn M128 N64", 128, 64, 96, 1, 1, 129, 0, 3, T, 4);
  run("rate SS mn/mn M128 N128", 128, 128, 64, 1, 1, 129, 0, 3, T, 2);
  run("rate SS mn/mn M128 N256", 128, 256, 32, 1, 1, 129, 0, 3, T, 1);
  run("rate SS mn/mn M64 N64", 64, 64, 96, 1, 1, 129, 0, 3, T, 4);
  run("rate SS A mn/B k N32", 128, 32, 96, 1, 0, 129, 0, 3, T, 4, false, 128);
  run("rate SS A mn/B k N16", 128, 16, 96, 1, 0, 129, 0, 3, T, 4, false, 128);
  run("rate SS k/k M64 N64", 64, 64, 64, 0, 0, 128, 0, 3, T, 4);
  printf(bad ? "PROBE: %d numerics cases off\n" : "PROBE OK\n", bad);
  return 0;
}
