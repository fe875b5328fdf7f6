// Philox4x32-10 counter-based generator + float transforms (K-RNG).
// Host replica with the identical stream definition: nanofed_b200/ops/philox.py.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nfb {

struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  static constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;

  // 4 uint32 for 64-bit counter `group` under 64-bit key `seed`.
  __device__ __forceinline__ static uint4 draw(uint64_t group, uint64_t seed) {
    uint32_t c0 = static_cast<uint32_t>(group), c1 = static_cast<uint32_t>(group >> 32);
    uint32_t c2 = 0u, c3 = 0u;
    uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
      const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
      const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
      c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
      k0 += W0; k1 += W1;
    }
    return make_uint4(c0, c1, c2, c3);
  }

  // (0,1): ((x >> 8) + 0.5) * 2^-24
  __device__ __forceinline__ static float unit(uint32_t x) {
    return (static_cast<float>(x >> 8) + 0.5f) * 5.9604644775390625e-8f;
  }

  // 4 standard normals: lanes (0,1) and (2,3) are Box-Muller (cos, sin) pairs.
  __device__ __forceinline__ static float4 normal4(uint64_t group, uint64_t seed) {
    const uint4 w = draw(group, seed);
    const float r0 = sqrtf(-2.0f * logf(unit(w.x)));
    const float r1 = sqrtf(-2.0f * logf(unit(w.z)));
    float s0, c0, s1, c1;
    sincosf(6.283185307179586f * unit(w.y), &s0, &c0);
    sincosf(6.283185307179586f * unit(w.w), &s1, &c1);
    return make_float4(r0 * c0, r0 * s0, r1 * c1, r1 * s1);
  }

  __device__ __forceinline__ static float4 uniform4(uint64_t group, uint64_t seed) {
    const uint4 w = draw(group, seed);
    return make_float4(unit(w.x), unit(w.y), unit(w.z), unit(w.w));
  }

  // Laplace(0,1) by inverse CDF: -sgn(u-.5) * log1p(-2|u-.5|)
  __device__ __forceinline__ static float laplace1(float u) {
    const float d = u - 0.5f;
    const float m = -log1pf(fmaxf(-2.0f * fabsf(d), -1.0f + 1e-7f));
    return d > 0.f ? m : (d < 0.f ? -m : 0.f);
  }
  __device__ __forceinline__ static float4 laplace4(uint64_t group, uint64_t seed) {
    const float4 u = uniform4(group, seed);
    return make_float4(laplace1(u.x), laplace1(u.y), laplace1(u.z), laplace1(u.w));
  }
};

}  // namespace nfb
