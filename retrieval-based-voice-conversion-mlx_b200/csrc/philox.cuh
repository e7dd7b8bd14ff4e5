// Philox4x32-10 counter-based generator + Box-Muller normals (shared by the harmonic source and the z_p sampling).
#pragma once
#include <stdint.h>

namespace rvcdec {

__device__ __forceinline__ uint32_t mulhilo32(uint32_t a, uint32_t b, uint32_t* hi) {
  *hi = __umulhi(a, b);
  return a * b;
}

// Philox4x32-10 (Salmon et al., SC'11), counter = (c0, c1, 0, 0), key = seed.
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t k0, uint32_t k1) {
  uint32_t c2 = 0, c3 = 0;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0, hi1;
    const uint32_t lo0 = mulhilo32(0xD2511F53u, c0, &hi0);
    const uint32_t lo1 = mulhilo32(0xCD9E8D57u, c2, &hi1);
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float philox_normal(uint64_t seed, uint64_t idx) {
  const uint64_t ctr = idx >> 2;
  const uint4 r = philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)seed, (uint32_t)(seed >> 32));
  const int lane = (int)(idx & 3);
  const uint32_t ua = (lane < 2) ? r.x : r.z, ub = (lane < 2) ? r.y : r.w;
  const float u1 = ((float)ua + 1.0f) * 2.3283064365386963e-10f;  // (0,1]
  const float u2 = (float)ub * 2.3283064365386963e-10f;
  const float rad = sqrtf(-2.0f * __logf(u1));
  float sn, cs;
  sincospif(2.0f * u2, &sn, &cs);
  return (lane & 1) ? rad * sn : rad * cs;
}

}  // namespace rvcdec
