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

// ---------------------------------------------------------------------------------------------------------------
// lg_score_prune_tiles: the adaptive step of one layer for a whole batch in ONE launch, directly on the engine's split
// tiles (north_star (5): "fused confidence-MLP plus prefix-sum stream-compaction kernel"):
//   token confidence + matchability of every valid token (TokenConfidence.forward :102-107, get_matchability :317-318)
//   keep mask (get_pruning_mask :586-593), the count check_if_stop needs (:595-606),
//   exclusive scan, and the order-preserving compaction of the kept tokens (:536-547): descriptor row
//   (hi | lo bf16 split, 1 KB), rotary row (cos | sin, 256 B) and original index move to their new row of the OUTPUT
//   tiles; prune[ind] + 1 goes to a copy of the counters (a pair that exits at this layer keeps the old ones: the
//   reference breaks before pruning).  Rows behind the kept ones are zero-filled (padding must stay finite: the PV MMA
//   multiplies masked keys by P = 0).
// No fp32 rows, no tiles -> rows -> tiles round trip, no index_select.
// Work item = one ROW TILE (128 tokens) of one image, one CTA of 256 threads each; the grid covers every tile of the
// batch, so one pair per call spreads over 32 CTAs and 32 pairs over 1024.  Items are taken in ticket order (atomic
// counter), which makes the prefix exchange below safe on any dispatch order.  A CTA scores its 128 rows (1 KB read
// per token), publishes (kept, unconfident) of its tile in the workspace, sums the published counts of the earlier
// tiles of its image (at most tp - 1 words, all produced concurrently: no serial chain), and moves its kept rows to
// their final position; the re-read of the tile it has just scored comes from L1 / L2.  The image's last valid tile
// writes the counts and zero-fills the rows between the kept count and its own end; tiles behind the valid length
// zero-fill themselves.
namespace lg {

constexpr int kSpThreads = 256;
constexpr int kSpWarps = kSpThreads / 32;
constexpr unsigned long long kSpReady = 1ull << 63;

struct ScorePruneParams {
  const uint8_t* x;      // split tiles [B*tp0 + B*tp1][4][32 KB]
  const uint8_t* rot;    // [B*tp0 + B*tp1][cos 16 KB | sin 16 KB]
  uint8_t* x_out;
  uint8_t* rot_out;
  const int64_t* ind[2];     // (B, n_cur) original index of every row; null = identity
  int64_t* ind_out[2];
  const int64_t* prune[2];   // (B, n_orig)
  int64_t* prune_out[2];     // holds a copy of prune when the kernel starts (the entry point copies)
  const int* len[2];         // valid rows per image, null = n_cur
  int* len_out[2];
  int* counts;               // (B, 4): kept0, unconfident0, kept1, unconfident1
  unsigned long long* ws;    // [0] ticket, [1 + tile] = ready | unconfident << 32 | kept of the tile; zero on entry
  const float* w_tok; const float* b_tok; const float* w_match; const float* b_match;
  float thr, min_match;
  int do_prune, B, n_cur[2], tp[2], n_orig[2];
};

__device__ __forceinline__ void sp_zero_rows(uint8_t* xo, uint8_t* roto, int64_t* indo, int d0, int d1, int n_cur,
                                             int warp, int lane) {
  const int atom = lane >> 3, chunk = lane & 7;
  for (int d = d0 + warp; d < d1; d += kSpWarps) {
    const size_t dof = static_cast<size_t>(d >> 7) * 4 * 32768 + atom * 32768 + sw128_offset(d & 127, chunk);
    *reinterpret_cast<uint4*>(xo + dof) = make_uint4(0u, 0u, 0u, 0u);
    *reinterpret_cast<uint4*>(xo + dof + 16384) = make_uint4(0u, 0u, 0u, 0u);
    if (lane < 16)
      *reinterpret_cast<uint4*>(roto + static_cast<size_t>(d >> 7) * 32768 + (lane >> 3) * 16384 +
                                sw128_offset(d & 127, lane & 7)) = make_uint4(0u, 0u, 0u, 0u);
    else if (lane == 16 && d < n_cur) indo[d] = 0;
  }
}

__global__ void __launch_bounds__(kSpThreads) score_prune_tiles_kernel(const ScorePruneParams p) {
  __shared__ int s_item, s_keep[128], s_wsum[kSpWarps], s_wunc[kSpWarps], s_base, s_unc_before;
  __shared__ unsigned s_ballot[4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_item = static_cast<int>(atomicAdd(p.ws, 1ull));
  __syncthreads();
  const int item = s_item;
  const int tiles0 = p.B * p.tp[0];
  const int side = item >= tiles0;
  const int tp = p.tp[side];
  const int b = (item - side * tiles0) / tp, tile = (item - side * tiles0) % tp;
  const int n_cur = p.n_cur[side];
  const int nv = p.len[side] ? min(p.len[side][b], n_cur) : n_cur;
  const int nvt = max(1, (nv + 127) >> 7);  // tiles that take part in the prefix exchange
  const size_t tile0 = static_cast<size_t>(item - tile);  // first row tile of the image in the buffers
  const uint8_t* x = p.x + tile0 * 4 * 32768;
  uint8_t* xo = p.do_prune ? p.x_out + tile0 * 4 * 32768 : nullptr;
  uint8_t* roto = p.do_prune ? p.rot_out + tile0 * 32768 : nullptr;
  int64_t* indo = p.do_prune ? p.ind_out[side] + static_cast<size_t>(b) * n_cur : nullptr;
  if (tile >= nvt) {  // behind the valid rows: nothing to score, the tile of the output is padding
    if (p.do_prune) sp_zero_rows(xo, roto, indo, tile * 128, tile * 128 + 128, n_cur, warp, lane);
    return;
  }

  // lane -> (K atom, 16-byte chunk): always the same 8 features, their weights live in registers
  const int atom = lane >> 3, chunk = lane & 7;
  float wm[8], wt[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    wm[k] = __ldg(p.w_match + atom * 64 + chunk * 8 + k);
    wt[k] = p.w_tok ? __ldg(p.w_tok + atom * 64 + chunk * 8 + k) : 0.f;
  }
  const float bm = __ldg(p.b_match), bt = p.b_tok ? __ldg(p.b_tok) : 0.f;
  const int rows = min(128, nv - tile * 128);  // valid rows of this tile (0 only for an empty image)
  const uint8_t* xt = x + static_cast<size_t>(tile) * 4 * 32768 + atom * 32768;
  int unconf = 0;
  // warp w scores rows w, w + 8, ...: four rows (eight 16-byte loads per lane) in flight
  for (int r0 = warp; r0 < 128; r0 += 4 * kSpWarps) {
    uint4 hi[4], lo[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * kSpWarps;
      if (r < rows) {
        const uint8_t* src = xt + sw128_offset(r, chunk);
        hi[u] = __ldg(reinterpret_cast<const uint4*>(src));
        lo[u] = __ldg(reinterpret_cast<const uint4*>(src + 16384));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * kSpWarps;
      if (r >= 128) continue;
      int keep = 0;
      if (r < rows) {  // warp-uniform
        const uint32_t hh[4] = {hi[u].x, hi[u].y, hi[u].z, hi[u].w}, ll[4] = {lo[u].x, lo[u].y, lo[u].z, lo[u].w};
        float sm = 0.f, st = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float x0 = __uint_as_float(hh[k] << 16) + __uint_as_float(ll[k] << 16);
          const float x1 = __uint_as_float(hh[k] & 0xFFFF0000u) + __uint_as_float(ll[k] & 0xFFFF0000u);
          sm = fmaf(x0, wm[2 * k], sm); sm = fmaf(x1, wm[2 * k + 1], sm);
          st = fmaf(x0, wt[2 * k], st); st = fmaf(x1, wt[2 * k + 1], st);
        }
        sm = warp_sum(sm);
        st = warp_sum(st);
        const float match = sigmoidf(sm + bm);
        bool kp = match > p.min_match;
        if (p.w_tok) {
          const float conf = sigmoidf(st + bt);
          kp = kp || (conf <= p.thr);
          unconf += conf < p.thr ? 1 : 0;   // (every lane counts the same rows; lane 0's total is used)
        }
        keep = (kp || !p.do_prune) ? 1 : 0;
      }
      if (lane == 0) s_keep[r] = keep;
    }
  }
  if (lane == 0) s_wunc[warp] = unconf;
  __syncthreads();

  // local exclusive scan of the 128 keep flags (ballots of the first four warps) and the tile totals
  int flag = 0, excl = 0;
  if (threadIdx.x < 128) {
    flag = s_keep[threadIdx.x];
    const unsigned bal = __ballot_sync(0xffffffffu, flag != 0);
    if (lane == 0) s_ballot[warp] = bal;
    excl = __popc(bal & ((1u << lane) - 1u));
  }
  __syncthreads();
  int before_w = 0, cnt = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const int c = __popc(s_ballot[w]);
    if (w < warp) before_w += c;
    cnt += c;
  }
  int unc = 0;
#pragma unroll
  for (int w = 0; w < kSpWarps; ++w) unc += s_wunc[w];
  unsigned long long* slot = p.ws + 1 + tile0;  // one word per tile of this image
  if (threadIdx.x == 0) {
    __threadfence();
    atomicExch(slot + tile, kSpReady | (static_cast<unsigned long long>(unc) << 32) | static_cast<unsigned long long>(cnt));
  }
  // kept / unconfident tokens in the earlier tiles of the image: every earlier tile has an earlier ticket, so it is
  // running or done -- waiting for its word cannot deadlock
  if (warp == 0) {
    int base = 0, ub = 0;
    for (int t = lane; t < tile; t += 32) {
      unsigned long long v;
      do {
        v = *reinterpret_cast<volatile unsigned long long*>(slot + t);
      } while (!(v & kSpReady));
      base += static_cast<int>(v & 0xffffffffull);
      ub += static_cast<int>((v >> 32) & 0x7fffffffull);
    }
    base = static_cast<int>(warp_sum(static_cast<float>(base)));  // counts < 2^24: exact in fp32
    ub = static_cast<int>(warp_sum(static_cast<float>(ub)));
    if (lane == 0) { s_base = base; s_unc_before = ub; }
  }
  __syncthreads();
  const int base = s_base;
  if (threadIdx.x < 128) s_keep[threadIdx.x] = flag ? base + before_w + excl : -1;  // destination row or -1
  __syncthreads();
  const bool last = tile == nvt - 1;
  const int kept = base + cnt;
  if (last && threadIdx.x == 0) {
    p.counts[4 * b + 2 * side] = kept;
    p.counts[4 * b + 2 * side + 1] = s_unc_before + unc;
    if (p.len_out[side]) p.len_out[side][b] = kept;
  }
  if (!p.do_prune) return;

  const uint8_t* rot = p.rot + tile0 * 32768;
  const int n_orig = p.n_orig[side];
  const int64_t* pr = p.prune[side] + static_cast<size_t>(b) * n_orig;
  int64_t* pro = p.prune_out[side] + static_cast<size_t>(b) * n_orig;
  const int64_t* ind = p.ind[side] ? p.ind[side] + static_cast<size_t>(b) * n_cur : nullptr;
  for (int r = warp; r < rows; r += kSpWarps) {
    const int d = s_keep[r];
    if (d < 0) continue;
    const int i = tile * 128 + r;
    const uint8_t* src = xt + sw128_offset(r, chunk);
    const size_t dof = static_cast<size_t>(d >> 7) * 4 * 32768 + atom * 32768 + sw128_offset(d & 127, chunk);
    *reinterpret_cast<uint4*>(xo + dof) = __ldg(reinterpret_cast<const uint4*>(src));
    *reinterpret_cast<uint4*>(xo + dof + 16384) = __ldg(reinterpret_cast<const uint4*>(src + 16384));
    if (lane < 16) {  // rotary row: lanes 0-7 the cos chunks, 8-15 the sin chunks
      const size_t rs = static_cast<size_t>(tile) * 32768 + (lane >> 3) * 16384 + sw128_offset(r, lane & 7);
      const size_t rd = static_cast<size_t>(d >> 7) * 32768 + (lane >> 3) * 16384 + sw128_offset(d & 127, lane & 7);
      *reinterpret_cast<uint4*>(roto + rd) = __ldg(reinterpret_cast<const uint4*>(rot + rs));
    } else if (lane == 16) {
      const int64_t orig = ind ? ind[i] : i;
      indo[d] = orig;
      pro[orig] = pr[orig] + 1;  // kept original indices are distinct: no race
    }
  }
  // rows behind the kept ones, up to the end of this (the image's last valid) tile: zeros
  if (last) sp_zero_rows(xo, roto, indo, kept, nvt * 128, n_cur, warp, lane);
}

}  // namespace lg

extern "C" size_t lg_score_prune_ws_bytes(int B, int tp0, int tp1) {
  if (B <= 0 || tp0 <= 0 || tp1 <= 0) return 0;
  return (static_cast<size_t>(B) * (tp0 + tp1) + 1) * sizeof(unsigned long long);
}

extern "C" int lg_score_prune_tiles(const void* x_tiles, const void* rot_tiles, void* x_out, void* rot_out,
                                    const int64_t* ind0, const int64_t* ind1, int64_t* ind0_out, int64_t* ind1_out,
                                    const int64_t* prune0, const int64_t* prune1, int64_t* prune0_out,
                                    int64_t* prune1_out, const int* len0, const int* len1, int* len0_out, int* len1_out,
                                    int* counts, const float* w_tok, const float* b_tok, const float* w_match,
                                    const float* b_match, float thr, float min_match, int do_prune, void* ws, int B,
                                    int n0, int n1, int tp0, int tp1, int n0_orig, int n1_orig, cudaStream_t stream) {
  if (!x_tiles || !counts || !w_match || !b_match || !ws) return LG_ERR_NULL;
  if ((w_tok == nullptr) != (b_tok == nullptr)) return LG_ERR_NULL;
  if (do_prune && (!rot_tiles || !x_out || !rot_out || !ind0_out || !ind1_out || !prune0 || !prune1 || !prune0_out ||
                   !prune1_out))
    return LG_ERR_NULL;
  if (B <= 0 || n0 <= 0 || n1 <= 0 || tp0 * 128 < n0 || tp1 * 128 < n1 || n0_orig < n0 || n1_orig < n1) return LG_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(ws) & 7) != 0) return LG_ERR_ALIGN;
  cudaError_t e = cudaMemsetAsync(ws, 0, lg_score_prune_ws_bytes(B, tp0, tp1), stream);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (do_prune) {  // the counters of the tokens that stay out of this layer are carried over unchanged
    e = cudaMemcpyAsync(prune0_out, prune0, static_cast<size_t>(B) * n0_orig * sizeof(int64_t), cudaMemcpyDeviceToDevice, stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(prune1_out, prune1, static_cast<size_t>(B) * n1_orig * sizeof(int64_t), cudaMemcpyDeviceToDevice, stream);
    if (e != cudaSuccess) return static_cast<int>(e);
  }
  lg::ScorePruneParams p{};
  p.x = static_cast<const uint8_t*>(x_tiles); p.rot = static_cast<const uint8_t*>(rot_tiles);
  p.x_out = static_cast<uint8_t*>(x_out); p.rot_out = static_cast<uint8_t*>(rot_out);
  p.ind[0] = ind0; p.ind[1] = ind1; p.ind_out[0] = ind0_out; p.ind_out[1] = ind1_out;
  p.prune[0] = prune0; p.prune[1] = prune1; p.prune_out[0] = prune0_out; p.prune_out[1] = prune1_out;
  p.len[0] = len0; p.len[1] = len1; p.len_out[0] = len0_out; p.len_out[1] = len1_out;
  p.counts = counts; p.ws = static_cast<unsigned long long*>(ws);
  p.w_tok = w_tok; p.b_tok = b_tok; p.w_match = w_match; p.b_match = b_match;
  p.thr = thr; p.min_match = min_match; p.do_prune = do_prune; p.B = B;
  p.n_cur[0] = n0; p.n_cur[1] = n1; p.tp[0] = tp0; p.tp[1] = tp1; p.n_orig[0] = n0_orig; p.n_orig[1] = n1_orig;
  lg::score_prune_tiles_kernel<<<B * (tp0 + tp1), lg::kSpThreads, 0, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}
