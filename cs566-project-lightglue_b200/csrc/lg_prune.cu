// lg_prune : adaptive depth / width -- token confidence, matchability, keep mask,
// order-preserving stream compaction.
//
// Replaces, for the eval-mode adaptive path of the reference
// (gluefactory/models/matchers/lightglue.py:527-547):
//   TokenConfidence.forward (:102-107)            sigmoid(Linear(256->1)(desc))
//   MatchAssignment.get_matchability (:317-318)   sigmoid(matchability(desc))
//   get_pruning_mask (:586-593)                   keep = match > 1-width_conf  |  conf <= thr_i
//   check_if_stop (:595-606)                      count(conf < thr_i) over the kept tokens
//   torch.where + index_select x3 + prune[ind]+=1 (:536-547)
// All HBM-bound: one read of desc for the two dot products, one read + one write of the kept
// rows for the compaction.  The compaction is order preserving (ascending positions), as
// torch.where(mask)[1] is.
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

__device__ __forceinline__ float sigmoidf(float x) { return 1.f / (1.f + expf(-x)); }

// warp per 4 tokens: two 256-long dot products from ONE read of each row, all eight 16-byte loads of the lane
// in flight before the first use (the pass is a pure HBM stream of the descriptors)
// grid (ceil(N/32), B), block 256
constexpr int kRowsPerWarp = 4;
__global__ void __launch_bounds__(256) token_scores_kernel(const float* __restrict__ desc,
                                                           const float* __restrict__ w_tok, const float* __restrict__ b_tok,
                                                           const float* __restrict__ w_match,
                                                           const float* __restrict__ b_match, const int* __restrict__ len,
                                                           float* __restrict__ conf, float* __restrict__ match, int N) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const int n0 = (blockIdx.x * 8 + (threadIdx.x >> 5)) * kRowsPerWarp;
  const int nv = len ? min(len[b], N) : N;
  if (n0 >= N) return;
  float4 x[kRowsPerWarp][2];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    const bool ok = n0 + r < nv;
    const float4* row = reinterpret_cast<const float4*>(desc + (static_cast<size_t>(b) * N + (ok ? n0 + r : n0)) * 256);
#pragma unroll
    for (int h = 0; h < 2; ++h) x[r][h] = ok ? __ldcs(row + h * 32 + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float4 wm[2], wt[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    wm[h] = __ldg(reinterpret_cast<const float4*>(w_match) + h * 32 + lane);
    wt[h] = w_tok ? __ldg(reinterpret_cast<const float4*>(w_tok) + h * 32 + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    float sm = 0.f, st = 0.f;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      sm += x[r][h].x * wm[h].x + x[r][h].y * wm[h].y + x[r][h].z * wm[h].z + x[r][h].w * wm[h].w;
      st += x[r][h].x * wt[h].x + x[r][h].y * wt[h].y + x[r][h].z * wt[h].z + x[r][h].w * wt[h].w;
    }
    sm = warp_sum(sm);
    st = warp_sum(st);
    const int n = n0 + r;
    if (lane == 0 && n < N) {
      const size_t o = static_cast<size_t>(b) * N + n;
      match[o] = n < nv ? sigmoidf(sm + b_match[0]) : 0.f;
      if (conf) conf[o] = (n < nv && w_tok) ? sigmoidf(st + b_tok[0]) : 0.f;
    }
  }
}

// warp per 4 rows: one 256-long dot product + bias (matchability logits for the assignment, lightglue.py:313-314)
__global__ void __launch_bounds__(256) logit_rows_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                         const float* __restrict__ bias, float* __restrict__ out,
                                                         int64_t rows) {
  const int lane = threadIdx.x & 31;
  const int64_t n0 = (static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5)) * kRowsPerWarp;
  if (n0 >= rows) return;
  float4 v[kRowsPerWarp][2];
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    const float4* row = reinterpret_cast<const float4*>(x + min(n0 + r, rows - 1) * 256);
#pragma unroll
    for (int h = 0; h < 2; ++h) v[r][h] = __ldcs(row + h * 32 + lane);
  }
  float4 ww[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) ww[h] = __ldg(reinterpret_cast<const float4*>(w) + h * 32 + lane);
#pragma unroll
  for (int r = 0; r < kRowsPerWarp; ++r) {
    float acc = 0.f;
#pragma unroll
    for (int h = 0; h < 2; ++h) acc += v[r][h].x * ww[h].x + v[r][h].y * ww[h].y + v[r][h].z * ww[h].z + v[r][h].w * ww[h].w;
    acc = warp_sum(acc);
    if (lane == 0 && n0 + r < rows) out[n0 + r] = acc + bias[0];
  }
}

// one block per image: keep flags -> exclusive scan -> ascending list of kept positions.
// counts[b] = {kept, number of tokens with conf < thr (0 if conf == null)}
__global__ void __launch_bounds__(1024) prune_plan_kernel(const float* __restrict__ conf, const float* __restrict__ match,
                                                          const int* __restrict__ len, float thr, float min_match,
                                                          int* __restrict__ keep_idx, int* __restrict__ counts, int N) {
  __shared__ int warp_tot[32];
  __shared__ int carry, unconf_total;
  const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nv = len ? min(len[b], N) : N;
  if (threadIdx.x == 0) { carry = 0; unconf_total = 0; }
  __syncthreads();
  int unconf = 0;
  for (int base = 0; base < nv; base += 1024) {
    const int i = base + threadIdx.x;
    bool keep = false;
    if (i < nv) {
      const size_t o = static_cast<size_t>(b) * N + i;
      keep = match[o] > min_match;
      if (conf) {
        const float c = conf[o];
        keep = keep || (c <= thr);
        unconf += (c < thr) ? 1 : 0;
      }
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
    const int in_warp = __popc(ballot & ((1u << lane) - 1u));
    if (lane == 0) warp_tot[warp] = __popc(ballot);
    __syncthreads();
    int before = carry;
    for (int w = 0; w < warp; ++w) before += warp_tot[w];
    if (keep) keep_idx[static_cast<size_t>(b) * N + before + in_warp] = i;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += warp_tot[w];
      carry += t;
    }
    __syncthreads();
  }
  unconf = static_cast<int>(warp_sum(static_cast<float>(unconf)) + 0.5f);
  if (lane == 0 && unconf) atomicAdd(&unconf_total, unconf);
  __syncthreads();
  if (threadIdx.x == 0) { counts[2 * b] = carry; counts[2 * b + 1] = unconf_total; }
}

// gather the kept rows: desc (256 f32), cos / sin (32 f32 each), ind; bump prune[ind].
// grid (ceil(N/4), B), block 256: 64 threads per row
__global__ void __launch_bounds__(256) prune_gather_kernel(const float* __restrict__ desc, const float* __restrict__ cos_t,
                                                           const float* __restrict__ sin_t, const int64_t* __restrict__ ind,
                                                           const int* __restrict__ keep_idx, const int* __restrict__ counts,
                                                           float* __restrict__ desc_o, float* __restrict__ cos_o,
                                                           float* __restrict__ sin_o, int64_t* __restrict__ ind_o,
                                                           int64_t* __restrict__ prune, int N, int N_out, int N_orig) {
  const int b = blockIdx.y;
  const int j = blockIdx.x * 4 + (threadIdx.x >> 6), t = threadIdx.x & 63;
  if (j >= counts[2 * b]) return;
  const int src = keep_idx[static_cast<size_t>(b) * N + j];
  const size_t si = static_cast<size_t>(b) * N + src, di = static_cast<size_t>(b) * N_out + j;
  reinterpret_cast<float4*>(desc_o + di * 256)[t] = __ldg(reinterpret_cast<const float4*>(desc + si * 256) + t);
  if (t < 8) reinterpret_cast<float4*>(cos_o + di * 32)[t] = __ldg(reinterpret_cast<const float4*>(cos_t + si * 32) + t);
  else if (t < 16) reinterpret_cast<float4*>(sin_o + di * 32)[t - 8] = __ldg(reinterpret_cast<const float4*>(sin_t + si * 32) + t - 8);
  else if (t == 16) {
    const int64_t orig = ind[si];
    ind_o[di] = orig;
    prune[static_cast<size_t>(b) * N_orig + orig] += 1;  // kept positions are distinct: no race
  }
}

}  // namespace lg

extern "C" int lg_token_scores(const float* desc, const float* w_tok, const float* b_tok, const float* w_match,
                               const float* b_match, const int* len, float* conf, float* match, int B, int N,
                               cudaStream_t stream) {
  if (!desc || !w_match || !b_match || !match) return LG_ERR_NULL;
  if ((w_tok == nullptr) != (b_tok == nullptr) || (w_tok && !conf)) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  lg::token_scores_kernel<<<dim3((N + 31) / 32, B), 256, 0, stream>>>(desc, w_tok, b_tok, w_match, b_match, len, conf,
                                                                    match, N);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_prune_plan(const float* conf, const float* match, const int* len, float thr, float min_match,
                             int* keep_idx, int* counts, int B, int N, cudaStream_t stream) {
  if (!match || !keep_idx || !counts) return LG_ERR_NULL;
  if (B <= 0 || N <= 0) return LG_ERR_SHAPE;
  lg::prune_plan_kernel<<<B, 1024, 0, stream>>>(conf, match, len, thr, min_match, keep_idx, counts, N);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_prune_gather(const float* desc, const float* cos_t, const float* sin_t, const int64_t* ind,
                               const int* keep_idx, const int* counts, float* desc_o, float* cos_o, float* sin_o,
                               int64_t* ind_o, int64_t* prune, int B, int N, int N_out, int N_orig,
                               cudaStream_t stream) {
  if (!desc || !cos_t || !sin_t || !ind || !keep_idx || !counts || !desc_o || !cos_o || !sin_o || !ind_o || !prune)
    return LG_ERR_NULL;
  if (B <= 0 || N <= 0 || N_out <= 0 || N_orig < N) return LG_ERR_SHAPE;
  lg::prune_gather_kernel<<<dim3((N + 3) / 4, B), 256, 0, stream>>>(desc, cos_t, sin_t, ind, keep_idx, counts, desc_o,
                                                                    cos_o, sin_o, ind_o, prune, N, N_out, N_orig);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_logit_rows(const float* x, const float* w, const float* bias, float* out, int64_t rows,
                             cudaStream_t stream) {
  if (!x || !w || !bias || !out) return LG_ERR_NULL;
  if (rows <= 0 || rows > (static_cast<int64_t>(1) << 33)) return LG_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) & 15) return LG_ERR_ALIGN;
  lg::logit_rows_kernel<<<static_cast<unsigned>((rows + 31) / 32), 256, 0, stream>>>(x, w, bias, out, rows);
  return static_cast<int>(cudaGetLastError());
}
