// rng.h — counter-based input generator shared by host and device.
// The reference fills inputs with rand()&0xFF (reduction.cpp:698-705) and per-rank-seeded MT19937
// (mpi/externalfunctions.h:45-179, mpi/reduce.c:38-57). MT19937 is sequential; here element i of stream
// `seed` is a pure function splitmix64(seed, i), so any rank/GPU/thread can produce any element and the host
// mirror used by the oracle is bit-identical to what the fill kernel writes.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B2R_HD __host__ __device__ __forceinline__
#else
#define B2R_HD inline
#endif

namespace b2r {

B2R_HD uint64_t splitmix64(uint64_t seed, uint64_t index) {
  uint64_t z = seed * 0xD1342543DE82EF95ull + (index + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// bf16 <-> f32 bit helpers usable on both sides (round-to-nearest-even, NaN kept quiet).
B2R_HD uint16_t f32_to_bf16_bits(float f) {
  union { float f; uint32_t u; } v;
  v.f = f;
  uint32_t u = v.u;
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u);  // NaN
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}
B2R_HD float bf16_bits_to_f32(uint16_t b) {
  union { float f; uint32_t u; } v;
  v.u = ((uint32_t)b) << 16;
  return v.f;
}

enum FillMode { kFillSdk = 0, kFillFull = 1, kFillSigned = 2 };

B2R_HD int32_t gen_i32(uint64_t seed, uint64_t i, int mode) {
  uint64_t r = splitmix64(seed, i);
  if (mode == kFillSdk) return (int32_t)(r & 0xFF);
  return (int32_t)(uint32_t)(r >> 32);
}
B2R_HD float gen_f32(uint64_t seed, uint64_t i, int mode) {
  uint64_t r = splitmix64(seed, i);
  if (mode == kFillSdk) return (float)(r & 0xFF) / 2147483647.0f;
  float u = (float)(r >> 40) * (1.0f / 16777216.0f);  // 24 bits in [0,1)
  return mode == kFillSigned ? 2.0f * u - 1.0f : u;
}
B2R_HD double gen_f64(uint64_t seed, uint64_t i, int mode) {
  uint64_t r = splitmix64(seed, i);
  if (mode == kFillSdk) return (double)(r & 0xFF) / 2147483647.0;
  double u = (double)(r >> 11) * (1.0 / 9007199254740992.0);  // 53 bits in [0,1)
  return mode == kFillSigned ? 2.0 * u - 1.0 : u;
}
B2R_HD uint16_t gen_bf16(uint64_t seed, uint64_t i, int mode) {
  uint64_t r = splitmix64(seed, i);
  float u = (float)(r & 0xFF) * (1.0f / 256.0f);  // exactly representable in bf16
  if (mode == kFillSigned) u = 2.0f * u - 1.0f;
  return f32_to_bf16_bits(u);
}

}  // namespace b2r
