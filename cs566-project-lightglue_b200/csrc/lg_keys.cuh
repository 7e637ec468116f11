// Order-preserving 64-bit (value, index) keys shared by the match filter and the assignment kernel:
// atomicMax on  orderable(value) << 32 | ~index  picks the largest value and, on ties, the LOWEST index -- exactly
// torch.max's choice on the reference's score tensor (lightglue.py:323-324).
#pragma once
#include "lg_common.cuh"

namespace lg {

__device__ __forceinline__ uint32_t orderable_f32(float v) {  // monotone float -> uint32 map
  v += 0.0f;  // -0 -> +0 so that signed zeros tie like torch
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ unsigned long long make_key(float v, uint32_t idx) {
  return (static_cast<unsigned long long>(orderable_f32(v)) << 32) | (0xFFFFFFFFu - idx);
}
__device__ __forceinline__ void decode_key(unsigned long long k, float& v, uint32_t& idx) {
  uint32_t u = static_cast<uint32_t>(k >> 32);
  u = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
  v = __uint_as_float(u);
  idx = 0xFFFFFFFFu - static_cast<uint32_t>(k & 0xFFFFFFFFu);
}

}  // namespace lg
