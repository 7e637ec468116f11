// lg_nn : nearest-neighbour matching on a similarity matrix -- find_nn + mutual_check of the reference's
// NearestNeighborMatcher (gluefactory/models/matchers/nearest_neighbor_matcher.py:16-39,58-66).
//
//   find_nn(sim):  top-2 (or top-1) of every row, dist = 2 (1 - sim), keep the best index where
//                  dist0 <= ratio^2 * dist1 (ratio test, if set) and dist0 <= distance^2 (if set), else -1;
//   the same down the columns (sim^T); mutual_check keeps i <-> j only if both directions agree.
// Two streaming passes over sim (B, M, N) -- warp per row, thread per column (coalesced across the warp) -- each
// keeping (best, second best, best index; lowest index on ties), then one finalize kernel for both directions.
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

struct Top2 {
  float v1, v2;
  int i1;
};

__device__ __forceinline__ void top2_push(Top2& t, float v, int i) {
  if (v > t.v1 || t.i1 < 0) {
    t.v2 = t.i1 < 0 ? -INFINITY : t.v1;
    t.v1 = v;
    t.i1 = i;
  } else if (v > t.v2) {
    t.v2 = v;
  }
}
// merge with a candidate list that covers HIGHER-or-interleaved indices: ties go to the lower index
__device__ __forceinline__ void top2_merge(Top2& a, const Top2& b) {
  if (b.i1 < 0) return;
  if (a.i1 < 0) { a = b; return; }
  if (b.v1 > a.v1 || (b.v1 == a.v1 && b.i1 < a.i1)) {
    a.v2 = fmaxf(a.v1, b.v2);
    a.v1 = b.v1;
    a.i1 = b.i1;
  } else {
    a.v2 = fmaxf(a.v2, b.v1);
  }
}

// warp per row; out: best[b][r] = (v1, v2), idx[b][r] = i1 (-1 for rows beyond the valid count)
__global__ void __launch_bounds__(256) nn_rows_kernel(const float* __restrict__ sim, const int* __restrict__ m_len,
                                                      const int* __restrict__ n_len, float2* __restrict__ best,
                                                      int* __restrict__ idx, int M, int N) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= M) return;
  const int mv = m_len ? min(m_len[b], M) : M, nv = n_len ? min(n_len[b], N) : N;
  Top2 t{-INFINITY, -INFINITY, -1};
  if (r < mv) {
    const float* row = sim + (static_cast<int64_t>(b) * M + r) * N;
    for (int c = lane; c < nv; c += 32) top2_push(t, __ldcs(row + c), c);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      Top2 u{__shfl_xor_sync(0xffffffffu, t.v1, o), __shfl_xor_sync(0xffffffffu, t.v2, o),
             __shfl_xor_sync(0xffffffffu, t.i1, o)};
      top2_merge(t, u);
    }
  }
  if (lane == 0) {
    best[static_cast<int64_t>(b) * M + r] = make_float2(t.v1, t.v2);
    idx[static_cast<int64_t>(b) * M + r] = t.i1;
  }
}

// thread per column, rows split over blockDim.y slices merged through shared memory
__global__ void __launch_bounds__(256) nn_cols_kernel(const float* __restrict__ sim, const int* __restrict__ m_len,
                                                      const int* __restrict__ n_len, float2* __restrict__ best,
                                                      int* __restrict__ idx, int M, int N) {
  __shared__ Top2 part[8][32];
  const int b = blockIdx.y, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  const int mv = m_len ? min(m_len[b], M) : M, nv = n_len ? min(n_len[b], N) : N;
  Top2 t{-INFINITY, -INFINITY, -1};
  if (c < nv) {
    const float* col = sim + static_cast<int64_t>(b) * M * N + c;
    for (int r = ty; r < mv; r += 8) top2_push(t, __ldcs(col + static_cast<int64_t>(r) * N), r);
  }
  part[ty][tx] = t;
  __syncthreads();
  if (ty == 0 && c < N) {
    for (int k = 1; k < 8; ++k) top2_merge(t, part[k][tx]);
    best[static_cast<int64_t>(b) * N + c] = make_float2(t.v1, t.v2);
    idx[static_cast<int64_t>(b) * N + c] = t.i1;
  }
}

__device__ __forceinline__ int nn_pick(float2 v, int i, float ratio2, float dist2, bool use_ratio, bool use_dist) {
  if (i < 0) return -1;
  const float d0 = 2.f * (1.f - v.x), d1 = 2.f * (1.f - v.y);
  bool ok = true;
  if (use_ratio) ok = ok && (d0 <= ratio2 * d1);
  if (use_dist) ok = ok && (d0 <= dist2);
  return ok ? i : -1;
}

// one thread per point of image 0 and image 1
__global__ void nn_finalize_kernel(const float2* __restrict__ best0, const int* __restrict__ idx0,
                                   const float2* __restrict__ best1, const int* __restrict__ idx1, float ratio2,
                                   float dist2, int use_ratio, int use_dist, int mutual, int64_t* __restrict__ m0,
                                   int64_t* __restrict__ m1, float* __restrict__ ms0, float* __restrict__ ms1, int M,
                                   int N) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const float2* b0 = best0 + static_cast<int64_t>(b) * M;
  const float2* b1 = best1 + static_cast<int64_t>(b) * N;
  const int* i0 = idx0 + static_cast<int64_t>(b) * M;
  const int* i1 = idx1 + static_cast<int64_t>(b) * N;
  if (i < M) {
    int m = nn_pick(b0[i], i0[i], ratio2, dist2, use_ratio, use_dist);
    if (mutual && m >= 0 && nn_pick(b1[m], i1[m], ratio2, dist2, use_ratio, use_dist) != i) m = -1;
    m0[static_cast<int64_t>(b) * M + i] = m;
    ms0[static_cast<int64_t>(b) * M + i] = m >= 0 ? 1.f : 0.f;
  }
  if (i < N) {
    int m = nn_pick(b1[i], i1[i], ratio2, dist2, use_ratio, use_dist);
    if (mutual && m >= 0 && nn_pick(b0[m], i0[m], ratio2, dist2, use_ratio, use_dist) != i) m = -1;
    m1[static_cast<int64_t>(b) * N + i] = m;
    ms1[static_cast<int64_t>(b) * N + i] = m >= 0 ? 1.f : 0.f;
  }
}

}  // namespace lg

extern "C" size_t lg_nn_ws_bytes(int B, int M, int N) {
  if (B <= 0 || M <= 0 || N <= 0) return 0;
  return static_cast<size_t>(B) * (static_cast<size_t>(M) + N) * 12 + 512;
}

extern "C" int lg_nn_match(const float* sim, const int* m_len, const int* n_len, float ratio_thresh,
                           float distance_thresh, int mutual, int64_t* m0, int64_t* m1, float* ms0, float* ms1, void* ws,
                           int B, int M, int N, cudaStream_t stream) {
  if (!sim || !m0 || !m1 || !ms0 || !ms1 || !ws) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0 || B > 65535) return LG_ERR_SHAPE;
  uint8_t* w = static_cast<uint8_t*>(ws);
  float2* best0 = reinterpret_cast<float2*>(w); w += static_cast<size_t>(B) * M * 8;
  float2* best1 = reinterpret_cast<float2*>(w); w += static_cast<size_t>(B) * N * 8;
  int* idx0 = reinterpret_cast<int*>(w); w += static_cast<size_t>(B) * M * 4;
  int* idx1 = reinterpret_cast<int*>(w);
  lg::nn_rows_kernel<<<dim3((M + 7) / 8, B), 256, 0, stream>>>(sim, m_len, n_len, best0, idx0, M, N);
  lg::nn_cols_kernel<<<dim3((N + 31) / 32, B), 256, 0, stream>>>(sim, m_len, n_len, best1, idx1, M, N);
  const int pts = M > N ? M : N;
  lg::nn_finalize_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(
      best0, idx0, best1, idx1, ratio_thresh * ratio_thresh, distance_thresh * distance_thresh, ratio_thresh > 0.f,
      distance_thresh > 0.f, mutual, m0, m1, ms0, ms1, M, N);
  return static_cast<int>(cudaGetLastError());
}
