// Pieces shared by the tcgen05 Fisher-vector-product kernels (fvp_tc.cu: the fused narrow-observation kernel;
// fvp_tc_wide.cu: the three-pass form for wide observations / action vectors): PTX wrappers, shared-memory
// descriptors and the plane layouts of DESIGN.md 3.1.2.
#pragma once
#include "common.cuh"

namespace pol {
namespace tc {

constexpr int ROWS = 96;            // timesteps per tile (smem bound: four 24 KB planes next to 104 KB of weights)
constexpr int H = 64;               // hidden width
constexpr uint32_t PLANE = ROWS * 256;              // 96 rows x 64 fp32 = 2 blocks of [96][128 B]
constexpr uint32_t BLK = ROWS * 128;                // one block of 32 columns

// ---- PTX wrappers ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0, laneid = 0;
  asm volatile("{\n\t.reg .b32 %%rx;\n\t.reg .pred %%px;\n\telect.sync %%rx|%%px, %2;\n\t@%%px mov.s32 %1, 1;\n\tmov.s32 %0, %%rx;\n\t}\n"
               : "+r"(laneid), "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred;
}
// layout: 1 = 128 B swizzle on 32 B atoms (MN-major tf32), 2 = 128 B swizzle, 4 = 64 B swizzle
__device__ __forceinline__ uint64_t make_desc(uint32_t start, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = (uint64_t)((start >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
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
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
#define TC_R16(v, o) "=r"(v[o + 0]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]), "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7]), \
                     "=r"(v[o + 8]), "=r"(v[o + 9]), "=r"(v[o + 10]), "=r"(v[o + 11]), "=r"(v[o + 12]), "=r"(v[o + 13]), "=r"(v[o + 14]), "=r"(v[o + 15])
#define TC_I16(v, o) "r"(v[o + 0]), "r"(v[o + 1]), "r"(v[o + 2]), "r"(v[o + 3]), "r"(v[o + 4]), "r"(v[o + 5]), "r"(v[o + 6]), "r"(v[o + 7]), \
                     "r"(v[o + 8]), "r"(v[o + 9]), "r"(v[o + 10]), "r"(v[o + 11]), "r"(v[o + 12]), "r"(v[o + 13]), "r"(v[o + 14]), "r"(v[o + 15])
// 32x32b shape: lane L of warp w <-> TMEM lane 32 w + L; register j <-> column col0 + j
__device__ __forceinline__ void tld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
               "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
               : TC_R16(v, 0), TC_R16(v, 16) : "r"(taddr) : "memory");
}
// load + wait in ONE statement: the registers are only valid after tcgen05.wait::ld, and nothing else ties their
// uses to a separate wait instruction
__device__ __forceinline__ void tld32w(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
               "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n\t"
               "tcgen05.wait::ld.sync.aligned;"
               : TC_R16(v, 0), TC_R16(v, 16) : "r"(taddr));
}
__device__ __forceinline__ void tld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : TC_R16(v, 0) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tst32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
               "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
               :: "r"(taddr), TC_I16(v, 0), TC_I16(v, 16));
}
__device__ __forceinline__ void tst16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               :: "r"(taddr), TC_I16(v, 0));
}
// shared memory -> global bulk copy (bulk async-group completion, per issuing thread)
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// lo plane of the compensated product: x - trunc_tf32(x) (exact)
__device__ __forceinline__ float lo_of(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// ---- plane addressing ----------------------------------------------------------------------------------
// MN-major planes ([96 rows = timesteps][64 columns]): 2 blocks of 32 columns, rows of 128 B, 32-byte atoms
// swizzled by the row index mod 4.  A thread owns one row; `q` = 32-byte atom 0..7 of the 64 columns.
__device__ __forceinline__ uint32_t mn_atom(int r, int q) {
  return (uint32_t)(q >> 2) * BLK + (uint32_t)r * 128u + (uint32_t)(((q & 3) ^ (r & 3)) << 5);
}
// K-major plane, 128 B swizzle: [R rows][K columns], blocks of 32 columns
__host__ __device__ inline uint32_t k128_off(int R, int r, int k) {
  return (uint32_t)(k >> 5) * R * 128u + (uint32_t)r * 128u + (uint32_t)((((k & 31) >> 2) ^ (r & 7)) << 4) + (uint32_t)(k & 3) * 4u;
}
// K-major plane, 64 B swizzle: [R rows][16 columns]
__host__ __device__ inline uint32_t k64_off(int r, int k) {
  return (uint32_t)r * 64u + (uint32_t)((((k & 15) >> 2) ^ ((r >> 1) & 3)) << 4) + (uint32_t)(k & 3) * 4u;
}

// Row access without bank conflicts: rows r and r + 4 share a swizzle pattern, so in a 128-bit access two of any
// eight consecutive lanes would hit the same 16 bytes.  Lanes with bit 2 of the row set take the two halves of
// every 32-byte atom in the opposite order (the values are put back in column order by selects).
__device__ __forceinline__ void load_row32(const unsigned char* plane, int r, int blk, float (&v)[32]) {
  const int sel = (r >> 2) & 1;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const unsigned char* p = plane + mn_atom(r, blk * 4 + q);
    const float4 a = *reinterpret_cast<const float4*>(p + 16 * sel);
    const float4 b = *reinterpret_cast<const float4*>(p + 16 * (sel ^ 1));
    v[8 * q + 0] = sel ? b.x : a.x; v[8 * q + 1] = sel ? b.y : a.y; v[8 * q + 2] = sel ? b.z : a.z; v[8 * q + 3] = sel ? b.w : a.w;
    v[8 * q + 4] = sel ? a.x : b.x; v[8 * q + 5] = sel ? a.y : b.y; v[8 * q + 6] = sel ? a.z : b.z; v[8 * q + 7] = sel ? a.w : b.w;
  }
}
__device__ __forceinline__ void store_row32(unsigned char* plane, int r, int blk, const float (&v)[32]) {
  const int sel = (r >> 2) & 1;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    unsigned char* p = plane + mn_atom(r, blk * 4 + q);
    const float4 lo4 = make_float4(v[8 * q + 0], v[8 * q + 1], v[8 * q + 2], v[8 * q + 3]);
    const float4 hi4 = make_float4(v[8 * q + 4], v[8 * q + 5], v[8 * q + 6], v[8 * q + 7]);
    *reinterpret_cast<float4*>(p + 16 * sel) = sel ? hi4 : lo4;
    *reinterpret_cast<float4*>(p + 16 * (sel ^ 1)) = sel ? lo4 : hi4;
  }
}

}  // namespace tc
}  // namespace pol
