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
//   keep element j  <=>  (lane_j & 0x7fff) >= round(p * 32768)
//   (15-bit draws leave each lane's top bit free as a guard, so two lanes are compared with one
//    32-bit subtract and turned into bf16x2 / fp32 bit masks with one PRMT — see keep_flags2)
//
// `group` is the index of the 8-element group, `offset` the per-call counter kept by the host
// RNG tracker. oracle/philox.py restates exactly this, so masks are compared bit for bit.
#pragma once
#include <stdint.h>

namespace fx {

struct Philox8 {
  uint32_t lanes[4];  // each holds two 16-bit draws: element 2i = low half, 2i+1 = high half
};

__host__ __device__ __forceinline__ Philox8 philox4x32_10(uint64_t seed, uint64_t offset,
                                                          uint64_t group) {
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
  uint32_t c0 = static_cast<uint32_t>(group), c1 = static_cast<uint32_t>(group >> 32);
  uint32_t c2 = static_cast<uint32_t>(offset), c3 = static_cast<uint32_t>(offset >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = static_cast<uint64_t>(0xD2511F53u) * c0;  // one IMAD.WIDE each
    const uint64_t p1 = static_cast<uint64_t>(0xCD9E8D57u) * c2;
    const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k1;
    c0 = n0; c1 = static_cast<uint32_t>(p1); c2 = n2; c3 = static_cast<uint32_t>(p0);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  Philox8 out;
  out.lanes[0] = c0; out.lanes[1] = c1; out.lanes[2] = c2; out.lanes[3] = c3;
  return out;
}

// Same function with the key schedule precomputed on the host (rk[2r] = k0 + r*W0, rk[2r+1] =
// k1 + r*W1): when rk lives in kernel-parameter (constant) memory the round keys cost no
// instructions, leaving 2 IMAD.WIDE + 2 LOP3 per round.
struct PhiloxKeys {
  uint32_t rk[20];
};
__host__ __device__ __forceinline__ PhiloxKeys philox_round_keys(uint64_t seed) {
  PhiloxKeys k;
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
  for (int r = 0; r < 10; ++r) {
    k.rk[2 * r] = k0;
    k.rk[2 * r + 1] = k1;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return k;
}
__device__ __forceinline__ Philox8 philox4x32_10_rk(const PhiloxKeys& k, uint32_t off_lo, uint32_t off_hi,
                                                    uint64_t group) {
  uint32_t c0 = static_cast<uint32_t>(group), c1 = static_cast<uint32_t>(group >> 32);
  uint32_t c2 = off_lo, c3 = off_hi;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = static_cast<uint64_t>(0xD2511F53u) * c0;
    const uint64_t p1 = static_cast<uint64_t>(0xCD9E8D57u) * c2;
    const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k.rk[2 * r];
    const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k.rk[2 * r + 1];
    c0 = n0; c1 = static_cast<uint32_t>(p1); c2 = n2; c3 = static_cast<uint32_t>(p0);
  }
  Philox8 out;
  out.lanes[0] = c0; out.lanes[1] = c1; out.lanes[2] = c2; out.lanes[3] = c3;
  return out;
}

__host__ __device__ __forceinline__ uint32_t dropout_threshold(float p) {
  float t = p * 32768.0f + 0.5f;
  if (t < 0.f) t = 0.f;
  if (t > 32767.f) t = 32767.f;
  return static_cast<uint32_t>(t);
}

// keep flag of element j (0..7) of a group
__host__ __device__ __forceinline__ bool philox_keep(const Philox8& r, int j, uint32_t thresh) {
  const uint32_t w = r.lanes[j >> 1];
  const uint32_t v = ((j & 1) ? (w >> 16) : w) & 0x7FFFu;
  return v >= thresh;
}

// Both lanes of a word at once: bit 15 (elements 2i) and bit 31 (2i+1) of the result are the keep
// flags. thresh2 = thresh | thresh << 16. Per lane (d | 0x8000) - T lies in [1, 0xFFFF], so the
// low lane never borrows from the high one.
__host__ __device__ __forceinline__ uint32_t keep_flags2(uint32_t w, uint32_t thresh2) {
  return ((w & 0x7FFF7FFFu) | 0x80008000u) - thresh2;
}

}  // namespace fx
