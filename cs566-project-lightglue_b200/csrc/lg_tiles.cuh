// Split-tile activation format helpers shared by the linear / FFN kernels (see lg_linear.cu for the layout).
#pragma once
#include "lg_common.cuh"

namespace lg {

constexpr int kSplitAtom = 32768;   // [128 rows][64 K] as hi 16 KB | lo 16 KB, each a SWIZZLE_128B bf16 tile image
constexpr int kHalfAtom = 16384;

template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// 8 fp32 -> (hi, lo) bf16x8 chunks
__device__ __forceinline__ void split8(const float* x, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(x[2 * i]), h1 = __float2bfloat16_rn(x[2 * i + 1]);
    const float r0 = x[2 * i] - __bfloat162float(h0), r1 = x[2 * i + 1] - __bfloat162float(h1);
    h[i] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
    l[i] = pack_bf16x2(r0, r1);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}
// the same on 4 packed pairs: 1 F2FP + 2 bit ops + 1 FADD2 + 1 F2FP per pair
__device__ __forceinline__ void split8(const float2* x, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    h[i] = pack_bf16x2(x[i].x, x[i].y);
    const float2 r = __fadd2_rn(x[i], make_float2(-__uint_as_float(h[i] << 16), -__uint_as_float(h[i] & 0xFFFF0000u)));
    l[i] = pack_bf16x2(r.x, r.y);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }


}  // namespace lg
