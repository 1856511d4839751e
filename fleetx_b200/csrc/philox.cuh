// philox.cuh — counter-based dropout masks (Philox4x32-10).
//
// The reference draws dropout masks from Paddle's per-device Philox stream inside
// get_rng_state_tracker().rng_state('global_seed' | 'local_seed') (hybrid_model.py:280, 546,
// 556; seeds from ppfleetx/utils/env.py:27-46). Paddle's stream consumption cannot be matched
// without Paddle, so this library defines the mask as a pure function instead:
//
//   key     = (seed lo32, seed hi32)
//   counter = (group lo32, group hi32, offset lo32, offset hi32)
//   one call -> 4 x 32 random bits -> 8 x 16-bit lanes -> 8 consecutive elements
//   keep element j  <=>  lane_j >= round(p * 65536)
//
// `group` is the index of the 8-element group, `offset` the per-call counter kept by the host
// RNG tracker. oracle/philox.py restates exactly this, so masks are compared bit for bit.
#pragma once
#include <stdint.h>

namespace fx {

struct Philox8 {
  uint32_t lanes[4];  // each holds two 16-bit draws: element 2i = low half, 2i+1 = high half
};

__host__ __device__ __forceinline__ void philox_mulhilo(uint32_t a, uint32_t b, uint32_t& hi,
                                                        uint32_t& lo) {
#ifdef __CUDA_ARCH__
  lo = a * b;
  hi = __umulhi(a, b);
#else
  const uint64_t p = static_cast<uint64_t>(a) * b;
  lo = static_cast<uint32_t>(p);
  hi = static_cast<uint32_t>(p >> 32);
#endif
}

__host__ __device__ __forceinline__ Philox8 philox4x32_10(uint64_t seed, uint64_t offset,
                                                          uint64_t group) {
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
  uint32_t c0 = static_cast<uint32_t>(group), c1 = static_cast<uint32_t>(group >> 32);
  uint32_t c2 = static_cast<uint32_t>(offset), c3 = static_cast<uint32_t>(offset >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0, lo0, hi1, lo1;
    philox_mulhilo(0xD2511F53u, c0, hi0, lo0);
    philox_mulhilo(0xCD9E8D57u, c2, hi1, lo1);
    const uint32_t n0 = hi1 ^ c1 ^ k0;
    const uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  Philox8 out;
  out.lanes[0] = c0; out.lanes[1] = c1; out.lanes[2] = c2; out.lanes[3] = c3;
  return out;
}

__host__ __device__ __forceinline__ uint32_t dropout_threshold(float p) {
  float t = p * 65536.0f + 0.5f;
  if (t < 0.f) t = 0.f;
  if (t > 65536.f) t = 65536.f;
  return static_cast<uint32_t>(t);
}

// keep flag of element j (0..7) of a group
__host__ __device__ __forceinline__ bool philox_keep(const Philox8& r, int j, uint32_t thresh) {
  const uint32_t w = r.lanes[j >> 1];
  const uint32_t v = (j & 1) ? (w >> 16) : (w & 0xFFFFu);
  return v >= thresh;
}

}  // namespace fx
