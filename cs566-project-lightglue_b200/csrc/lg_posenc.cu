// lg_posenc : keypoint normalisation + count-conditioned Fourier positional encoding.
//
// Replaces normalize_keypoints (reference gluefactory/models/matchers/lightglue.py:21-33) and
// ConditionalLearnableFourierPositionalEncoding.forward (lightglue.py:63-77) as called at
// lightglue.py:448-456,490-494.  Writes the compact (B,N,32) cos / sin tables (frequency j is
// lanes 2j, 2j+1 of the reference's repeat_interleave'd (2,B,1,N,64) tensor); the tables are
// consumed by lg_pack_self_qkv, which applies the rotary while producing the Q/K tiles.
// Phases reach ~N radians (condition ~ w*N), so the fp32 rounding sequence of the reference
// is kept (projection rounded, then + condition rounded) and full-range sincosf is used.
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

// per-image (shift, scale) when no image size is given: size = 1 + max - min over the points
// lens (B) non-null: only the first lens[b] points of image b count (padded batches of unequal images)
__global__ void posenc_extent_kernel(const float* __restrict__ kpts, const int* __restrict__ lens,
                                     float* __restrict__ size_out, int N) {
  const int b = blockIdx.x;
  const int nv = lens ? min(max(lens[b], 0), N) : N;
  float mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    const float2 k = reinterpret_cast<const float2*>(kpts)[static_cast<size_t>(b) * N + i];
    mnx = fminf(mnx, k.x); mxx = fmaxf(mxx, k.x); mny = fminf(mny, k.y); mxy = fmaxf(mxy, k.y);
  }
  __shared__ float red[4][32];
  mnx = -warp_max(-mnx); mny = -warp_max(-mny); mxx = warp_max(mxx); mxy = warp_max(mxy);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { red[0][w] = mnx; red[1][w] = mny; red[2][w] = mxx; red[3][w] = mxy; }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    mnx = l < nw ? red[0][l] : INFINITY; mny = l < nw ? red[1][l] : INFINITY;
    mxx = l < nw ? red[2][l] : -INFINITY; mxy = l < nw ? red[3][l] : -INFINITY;
    mnx = -warp_max(-mnx); mny = -warp_max(-mny); mxx = warp_max(mxx); mxy = warp_max(mxy);
    if (l == 0) {  // an image without points gets a unit extent (nothing reads its encodings)
      size_out[2 * b] = nv > 0 ? __fsub_rn(__fadd_rn(1.f, mxx), mnx) : 1.f;
      size_out[2 * b + 1] = nv > 0 ? __fsub_rn(__fadd_rn(1.f, mxy), mny) : 1.f;
    }
  }
}

// one thread per (b, n, j): 32 consecutive threads write one point's 32 frequencies
// scales / oris (B,N) non-null: the 4-d positional input of add_scale_ori (lightglue.py:458-476), wr is (32,4)
__global__ void posenc_kernel(const float* __restrict__ kpts, const float* __restrict__ scales,
                              const float* __restrict__ oris, const float* __restrict__ size,
                              const float* __restrict__ wr, const float* __restrict__ wc,
                              const float* __restrict__ bc, float cond, const float* __restrict__ cond_b,
                              float* __restrict__ cos_t, float* __restrict__ sin_t, int B, int N) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<size_t>(B) * N * 32) return;
  const int j = idx & 31;
  const size_t pt = idx >> 5;
  const int b = static_cast<int>(pt / N);
  const float2 k = reinterpret_cast<const float2*>(kpts)[pt];
  const float w = size[2 * b], h = size[2 * b + 1];
  // (kpts - size/2) / (max(size)/2), IEEE division like torch
  const float scale = __fmul_rn(fmaxf(w, h), 0.5f);
  const float x = __fdiv_rn(__fsub_rn(k.x, __fmul_rn(w, 0.5f)), scale);
  const float y = __fdiv_rn(__fsub_rn(k.y, __fmul_rn(h, 0.5f)), scale);
  // Linear(2->32, no bias): x*w0 + y*w1 accumulated with fma like the CPU GEMV micro-kernel
  float proj;
  if (scales == nullptr) {
    proj = __fmaf_rn(y, wr[2 * j + 1], __fmul_rn(x, wr[2 * j]));
  } else {  // Linear(4->32): [x, y, scale, ori] . wr[j, :]
    proj = __fmaf_rn(y, wr[4 * j + 1], __fmul_rn(x, wr[4 * j]));
    proj = __fmaf_rn(scales[pt], wr[4 * j + 2], proj);
    proj = __fmaf_rn(oris[pt], wr[4 * j + 3], proj);
  }
  // Linear(1->32)(relu(count)): count*w + b
  // cond_b (B) non-null: one count per image (a padded batch of unequal images, each conditioned on its own count)
  const float modulated = __fmaf_rn(fmaxf(cond_b ? cond_b[b] : cond, 0.f), wc[j], bc[j]);
  const float phase = __fadd_rn(proj, modulated);
  float s, c;
  sincosf(phase, &s, &c);
  cos_t[idx] = c;
  sin_t[idx] = s;
}

}  // namespace lg

extern "C" int lg_posenc(const float* kpts, const float* size, const float* wr, const float* wc, const float* bc,
                         float cond, float* cos_t, float* sin_t, int B, int N, cudaStream_t stream) {
  if (!kpts || !wr || !wc || !bc || !cos_t || !sin_t) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  if (!size) return LG_ERR_NULL;  // size=None branch: call lg_keypoint_extent first
  const size_t total = static_cast<size_t>(B) * N * 32;
  lg::posenc_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(kpts, nullptr, nullptr, size, wr, wc,
                                                                                   bc, cond, nullptr, cos_t, sin_t, B, N);
  return static_cast<int>(cudaGetLastError());
}

// add_scale_ori = True (lightglue.py:458-476): the positional input is [x, y, scale, ori]; wr is (32, 4)
extern "C" int lg_posenc_scale_ori(const float* kpts, const float* scales, const float* oris, const float* size,
                                   const float* wr, const float* wc, const float* bc, float cond, float* cos_t,
                                   float* sin_t, int B, int N, cudaStream_t stream) {
  if (!kpts || !scales || !oris || !size || !wr || !wc || !bc || !cos_t || !sin_t) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  const size_t total = static_cast<size_t>(B) * N * 32;
  lg::posenc_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(kpts, scales, oris, size, wr, wc, bc,
                                                                                   cond, nullptr, cos_t, sin_t, B, N);
  return static_cast<int>(cudaGetLastError());
}

// size_out (B,2) f32 = 1 + max - min over each image's keypoints (lightglue.py:25-26)
extern "C" int lg_keypoint_extent(const float* kpts, float* size_out, int B, int N, cudaStream_t stream) {
  if (!kpts || !size_out) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  lg::posenc_extent_kernel<<<B, 256, 0, stream>>>(kpts, nullptr, size_out, N);
  return static_cast<int>(cudaGetLastError());
}

// The same over the first lens[b] points of every image (lens (B) int32, may be null = all N)
extern "C" int lg_keypoint_extent_valid(const float* kpts, const int* lens, float* size_out, int B, int N,
                                        cudaStream_t stream) {
  if (!kpts || !size_out) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  lg::posenc_extent_kernel<<<B, 256, 0, stream>>>(kpts, lens, size_out, N);
  return static_cast<int>(cudaGetLastError());
}

// lg_posenc / lg_posenc_scale_ori with ONE CONDITION PER IMAGE: cond_b (B) f32 replaces the scalar count, so that a
// padded batch of images with unequal keypoint counts gets, image by image, the encodings of the reference's own
// one-pair-per-call forward (condition = that image's count, lightglue.py:490-491).  scales / oris may be null.
extern "C" int lg_posenc_per_image(const float* kpts, const float* scales, const float* oris, const float* size,
                                   const float* wr, const float* wc, const float* bc, const float* cond_b,
                                   float* cos_t, float* sin_t, int B, int N, cudaStream_t stream) {
  if (!kpts || !size || !wr || !wc || !bc || !cond_b || !cos_t || !sin_t) return LG_ERR_NULL;
  if ((scales == nullptr) != (oris == nullptr)) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  const size_t total = static_cast<size_t>(B) * N * 32;
  lg::posenc_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(kpts, scales, oris, size, wr, wc, bc,
                                                                                   0.f, cond_b, cos_t, sin_t, B, N);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_version(void) { return 100; }
#ifndef LG_SOURCE_ID
#define LG_SOURCE_ID "unknown"
#endif
// the "LG_SOURCE_ID=" prefix lets the binding read the digest from the file before loading it
extern "C" const char* lg_source_id(void) { return "LG_SOURCE_ID=" LG_SOURCE_ID + 13; }
