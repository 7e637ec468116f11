// lg_pack_* : producer of the attention operand tiles.
//
// Replaces, for SelfBlock (reference gluefactory/models/matchers/lightglue.py:185-188), the
// unflatten/transpose of the interleaved Wqkv output plus apply_cached_rotary_emb on q and k
// (lightglue.py:36-43), and for CrossBlock (lightglue.py:225-228) the head split of to_qk / to_v.
// Output: bf16 operand tiles laid out in HBM as the exact byte image of a SWIZZLE_128B shared
// memory tile, so the attention kernel stages them with one flat bulk (TMA) copy each:
//   Q / K tile  : [128 points][64 dh]            16 KB, K-major rows of 128 B
//   V^T tile    : 2 x [64 dh][64 points]         16 KB, K-major (points are the MMA K dim)
// Buffers are [B][H][tiles_pad][16 KB], tiles_pad even; rows >= valid length are zero.
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

struct PackParams {
  const float* src_qk;   // self: qkv (B,N,768) interleaved; cross: to_qk output (B,N,256)
  const float* src_v;    // cross only: to_v output (B,N,256)
  const float* cos_t;    // (B,N,32) or null (no rotary)
  const float* sin_t;
  const int* len;        // per-image valid point count or null
  uint8_t* q_tiles;      // self: rotated q ; cross: qk
  uint8_t* k_tiles;      // self: rotated k ; cross: unused (null)
  uint8_t* v_tiles;      // transposed v
  int B, N, H, tiles_pad;
  int self_mode;
};

// grid (tiles_pad, H, B), block 1024 = 128 rows x 8 chunks of 8 head-dims
__global__ void __launch_bounds__(1024, 1) pack_kernel(PackParams p) {
  __shared__ __align__(16) __nv_bfloat16 vt[64][128 + 8];  // +8: break bank alignment of the transposed writes
  const int tile = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int c = threadIdx.x & 7;    // 16-byte chunk = head dims 8c..8c+7
  const int r = threadIdx.x >> 3;   // row inside the tile
  const int n = tile * kTile + r;
  const int len = p.len ? min(p.len[b], p.N) : p.N;
  const bool valid = n < len;

  float q[8], k[8], v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) q[i] = k[i] = v[i] = 0.f;

  if (valid) {
    if (p.self_mode) {
      // f = h*192 + d*3 + {q,k,v}; this thread owns d = 8c..8c+7 -> 24 contiguous floats
      const float4* src = reinterpret_cast<const float4*>(p.src_qk + (static_cast<size_t>(b) * p.N + n) * (p.H * 192) +
                                                          h * 192 + c * 24);
      float t[24];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        float4 x = __ldg(src + i);
        t[4 * i] = x.x; t[4 * i + 1] = x.y; t[4 * i + 2] = x.z; t[4 * i + 3] = x.w;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) { q[i] = t[3 * i]; k[i] = t[3 * i + 1]; v[i] = t[3 * i + 2]; }
      if (p.cos_t) {
        // frequency j = d/2 sits in lanes (2j, 2j+1): (x0,x1) -> (x0*c - x1*s, x1*c + x0*s)
        const size_t tb = (static_cast<size_t>(b) * p.N + n) * 32 + c * 4;
        const float4 cs = __ldg(reinterpret_cast<const float4*>(p.cos_t + tb));
        const float4 sn = __ldg(reinterpret_cast<const float4*>(p.sin_t + tb));
        const float cc[4] = {cs.x, cs.y, cs.z, cs.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // same rounding sequence as the reference: (t*cos) + (rot(t)*sin), no fma contraction
          float q0 = __fadd_rn(__fmul_rn(q[2 * j], cc[j]), __fmul_rn(-q[2 * j + 1], ss[j]));
          float q1 = __fadd_rn(__fmul_rn(q[2 * j + 1], cc[j]), __fmul_rn(q[2 * j], ss[j]));
          float k0 = __fadd_rn(__fmul_rn(k[2 * j], cc[j]), __fmul_rn(-k[2 * j + 1], ss[j]));
          float k1 = __fadd_rn(__fmul_rn(k[2 * j + 1], cc[j]), __fmul_rn(k[2 * j], ss[j]));
          q[2 * j] = q0; q[2 * j + 1] = q1; k[2 * j] = k0; k[2 * j + 1] = k1;
        }
      }
    } else {
      const size_t off = (static_cast<size_t>(b) * p.N + n) * (p.H * 64) + h * 64 + c * 8;
      const float4* sq = reinterpret_cast<const float4*>(p.src_qk + off);
      const float4* sv = reinterpret_cast<const float4*>(p.src_v + off);
      float4 a0 = __ldg(sq), a1 = __ldg(sq + 1), b0 = __ldg(sv), b1 = __ldg(sv + 1);
      q[0] = a0.x; q[1] = a0.y; q[2] = a0.z; q[3] = a0.w; q[4] = a1.x; q[5] = a1.y; q[6] = a1.z; q[7] = a1.w;
      v[0] = b0.x; v[1] = b0.y; v[2] = b0.z; v[3] = b0.w; v[4] = b1.x; v[5] = b1.y; v[6] = b1.z; v[7] = b1.w;
    }
  }

  const size_t tile_base = ((static_cast<size_t>(b) * p.H + h) * p.tiles_pad + tile) * 16384;
  {
    uint4 o;
    o.x = pack_bf16x2(q[0], q[1]); o.y = pack_bf16x2(q[2], q[3]);
    o.z = pack_bf16x2(q[4], q[5]); o.w = pack_bf16x2(q[6], q[7]);
    *reinterpret_cast<uint4*>(p.q_tiles + tile_base + sw128_offset(r, c)) = o;
    if (p.k_tiles) {
      o.x = pack_bf16x2(k[0], k[1]); o.y = pack_bf16x2(k[2], k[3]);
      o.z = pack_bf16x2(k[4], k[5]); o.w = pack_bf16x2(k[6], k[7]);
      *reinterpret_cast<uint4*>(p.k_tiles + tile_base + sw128_offset(r, c)) = o;
    }
  }
  // V transposed through shared memory: vt[d][point]
#pragma unroll
  for (int i = 0; i < 8; ++i) vt[c * 8 + i][r] = __float2bfloat16_rn(v[i]);
  __syncthreads();
  {
    // 64 dh rows x 16 chunks of 8 points = 1024 chunks: one per thread
    const int d = threadIdx.x >> 4, ch = threadIdx.x & 15;
    const uint4 o = *reinterpret_cast<const uint4*>(&vt[d][ch * 8]);
    const int atom = ch >> 3, cc = ch & 7;  // points 0-63 | 64-127
    *reinterpret_cast<uint4*>(p.v_tiles + tile_base + atom * 8192 + sw128_offset(d, cc)) = o;
  }
}

}  // namespace lg

static int check_pack(const void* a, const void* q, const void* v, int B, int N, int H, int tiles_pad) {
  if (!a || !q || !v) return LG_ERR_NULL;
  if (B <= 0 || N <= 0 || H <= 0) return LG_ERR_SHAPE;
  if (tiles_pad * lg::kTile < N || (tiles_pad & 1)) return LG_ERR_SHAPE;
  return 0;
}

extern "C" int lg_pack_self_qkv(const float* qkv, const float* cos_t, const float* sin_t, const int* len,
                                void* q_tiles, void* k_tiles, void* v_tiles, int B, int N, int H, int tiles_pad,
                                cudaStream_t stream) {
  if (int e = check_pack(qkv, q_tiles, v_tiles, B, N, H, tiles_pad)) return e;
  if (!k_tiles || (cos_t == nullptr) != (sin_t == nullptr)) return LG_ERR_NULL;
  lg::PackParams p{qkv, nullptr, cos_t, sin_t, len, static_cast<uint8_t*>(q_tiles), static_cast<uint8_t*>(k_tiles),
                   static_cast<uint8_t*>(v_tiles), B, N, H, tiles_pad, 1};
  lg::pack_kernel<<<dim3(tiles_pad, H, B), 1024, 0, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_pack_cross_qkv(const float* qk, const float* v, const int* len, void* qk_tiles, void* v_tiles, int B,
                                 int N, int H, int tiles_pad, cudaStream_t stream) {
  if (int e = check_pack(qk, qk_tiles, v_tiles, B, N, H, tiles_pad)) return e;
  if (!v) return LG_ERR_NULL;
  lg::PackParams p{qk, v, nullptr, nullptr, len, static_cast<uint8_t*>(qk_tiles), nullptr,
                   static_cast<uint8_t*>(v_tiles), B, N, H, tiles_pad, 0};
  lg::pack_kernel<<<dim3(tiles_pad, H, B), 1024, 0, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}
