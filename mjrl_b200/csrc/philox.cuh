// Philox4x32-10 counter-based generator and the Box-Muller pair used by the action samplers
// (policy.cu: mjrl_policy_act, rollout.cu: mjrl_rollout_act).
#pragma once
#include <cstdint>

namespace pol {

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) { philox_round(c, k0, k1); k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
}

// four standard normals from one Philox block (two Box-Muller pairs)
__device__ __forceinline__ void philox_normal4(const uint32_t (&c)[4], float (&e)[4]) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const float u1 = ((float)c[2 * h] + 0.5f) * 2.3283064365386963e-10f;   // (0,1)
    const float u2 = ((float)c[2 * h + 1] + 0.5f) * 2.3283064365386963e-10f;
    const float rad = sqrtf(-2.f * logf(u1));
    float sn, cs;
    sincospif(2.f * u2, &sn, &cs);
    e[2 * h] = rad * cs; e[2 * h + 1] = rad * sn;
  }
}

}  // namespace pol
