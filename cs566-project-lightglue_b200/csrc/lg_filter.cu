// lg_filter_matches : mutual-nearest-neighbour match filter on the log-assignment matrix.
//
// Replaces filter_matches (reference gluefactory/models/matchers/lightglue.py:321-337).
// Pass 1 reads the (M x N) inner block of the scores ONCE: a warp owns a 32-row strip and a
// subset of the 32-column chunks; every lane keeps a running (value, first index) per row in
// registers and a column best over the strip's 32 rows, then publishes both with 64-bit
// atomicMax on an order-preserving key  (orderable(value) << 32 | ~index)  so that ties
// resolve to the LOWEST index exactly like torch.max on CPU (SURVEY.md hard part 7).
// Pass 2 (tiny) decodes the keys, checks mutuality, applies expf + threshold, writes int64.
// HBM-bound: algorithmic bytes = (M+1)(N+1)*4 read + (M+N)*12 written per pair.
#include "lg_common.cuh"
#include "lg_keys.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kFilterWarps = 4;

// grid (ceil(M/32), col_groups, B), block 128: warp w takes column split blockIdx.y*4 + w
__global__ void __launch_bounds__(kFilterWarps * 32) filter_scan_kernel(
    const float* __restrict__ scores, int64_t pitch, int64_t img_stride, const int* __restrict__ m_len,
    const int* __restrict__ n_len, unsigned long long* __restrict__ rowkey, unsigned long long* __restrict__ colkey,
    int M, int N, int nsplit) {
  const int b = blockIdx.z, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int split = blockIdx.y * kFilterWarps + warp;
  const int m = m_len ? min(m_len[b], M) : M, n = n_len ? min(n_len[b], N) : N;
  const int r0 = blockIdx.x * 32;
  if (r0 >= m || split >= nsplit) return;
  const int nrows = min(32, m - r0);
  const float* base = scores + b * img_stride + static_cast<int64_t>(r0) * pitch;

  float rbest[32];
  uint32_t ridx[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) { rbest[i] = -INFINITY; ridx[i] = 0xFFFFFFFFu; }
  bool any = false;

  const int nchunks = (n + 31) / 32;
  for (int ch = split; ch < nchunks; ch += nsplit) {
    const int c = ch * 32 + lane;
    const bool cok = c < n;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i)
      v[i] = (cok && i < nrows) ? __ldg(base + static_cast<int64_t>(i) * pitch + c) : -INFINITY;
    float cb = -INFINITY;
    uint32_t ci = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      // strict > keeps the first (lowest) index on ties; rows ascend with i, columns with ch
      if (v[i] > cb || (i == 0)) { cb = v[i]; ci = i; }
      if (v[i] > rbest[i] || ridx[i] == 0xFFFFFFFFu) { rbest[i] = v[i]; ridx[i] = c; }
    }
    if (cok) atomicMax(colkey + static_cast<int64_t>(b) * N + c, make_key(cb, r0 + ci));
    any = true;
  }
  if (!any) return;
  // reduce the per-lane row candidates across the warp (value desc, index asc), publish
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    // lanes whose column was out of range hold (-inf, c>=n): give them the worst index
    unsigned long long k = make_key(rbest[i], ridx[i] < static_cast<uint32_t>(n) ? ridx[i] : 0xFFFFFFFEu);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o);
      k = other > k ? other : k;
    }
    if (lane == (i & 31) && i < nrows) atomicMax(rowkey + static_cast<int64_t>(b) * M + r0 + i, k);
  }
}

// one thread per point of either image
__global__ void filter_finalize_kernel(const unsigned long long* __restrict__ rowkey,
                                       const unsigned long long* __restrict__ colkey, const int* __restrict__ m_len,
                                       const int* __restrict__ n_len, float th, int64_t* __restrict__ m0,
                                       int64_t* __restrict__ m1, float* __restrict__ ms0, float* __restrict__ ms1,
                                       int M, int N) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = m_len ? min(m_len[b], M) : M, n = n_len ? min(n_len[b], N) : N;
  const unsigned long long* rk = rowkey + static_cast<int64_t>(b) * M;
  const unsigned long long* ck = colkey + static_cast<int64_t>(b) * N;
  if (i < M) {
    int64_t match = -1;
    float score = 0.f;
    if (i < m && n > 0) {
      float mx, dummy;
      uint32_t j, iback;
      decode_key(rk[i], mx, j);
      decode_key(ck[j], dummy, iback);
      if (iback == static_cast<uint32_t>(i)) {  // mutual0
        score = expf(mx);
        if (score > th) match = j;
      }
    }
    m0[static_cast<int64_t>(b) * M + i] = match;
    ms0[static_cast<int64_t>(b) * M + i] = score;
  }
  if (i < N) {
    int64_t match = -1;
    float score = 0.f;
    if (i < n && m > 0) {
      float dummy, mx;
      uint32_t r, jback;
      decode_key(ck[i], dummy, r);
      decode_key(rk[r], mx, jback);
      if (jback == static_cast<uint32_t>(i)) {  // mutual1 (implies mutual0 of row r)
        score = expf(mx);                       // == mscores0[r]
        if (score > th) match = r;
      }
    }
    m1[static_cast<int64_t>(b) * N + i] = match;
    ms1[static_cast<int64_t>(b) * N + i] = score;
  }
}

}  // namespace lg

extern "C" size_t lg_filter_ws_bytes(int B, int M, int N) {
  if (B <= 0 || M < 0 || N < 0) return 0;
  return static_cast<size_t>(B) * (static_cast<size_t>(M) + N) * sizeof(unsigned long long);
}

extern "C" int lg_filter_matches(const float* scores, int64_t pitch, const int* m_len, const int* n_len, float th,
                                 int64_t* m0, int64_t* m1, float* ms0, float* ms1, void* ws, int B, int M, int N,
                                 cudaStream_t stream) {
  if (!scores || !m0 || !m1 || !ms0 || !ms1 || !ws) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0 || pitch < N) return LG_ERR_SHAPE;
  auto* rowkey = static_cast<unsigned long long*>(ws);
  auto* colkey = rowkey + static_cast<size_t>(B) * M;
  cudaError_t e = cudaMemsetAsync(ws, 0, lg_filter_ws_bytes(B, M, N), stream);
  if (e != cudaSuccess) return static_cast<int>(e);
  const int strips = (M + 31) / 32, nchunks = (N + 31) / 32;
  // enough warps to cover 148 SMs x 16 without splitting a row across more warps than chunks
  int nsplit = (148 * 16 + B * strips - 1) / (B * strips);
  nsplit = max(1, min(nsplit, nchunks));
  const int col_groups = (nsplit + lg::kFilterWarps - 1) / lg::kFilterWarps;
  lg::filter_scan_kernel<<<dim3(strips, col_groups, B), lg::kFilterWarps * 32, 0, stream>>>(
      scores, pitch, static_cast<int64_t>(M + 1) * pitch, m_len, n_len, rowkey, colkey, M, N, nsplit);
  e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  const int pts = M > N ? M : N;
  lg::filter_finalize_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(rowkey, colkey, m_len, n_len, th, m0, m1,
                                                                            ms0, ms1, M, N);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_filter_from_keys(const void* keys_ws, const int* m_len, const int* n_len, float th, int64_t* m0,
                                   int64_t* m1, float* ms0, float* ms1, int B, int M, int N, cudaStream_t stream) {
  if (!keys_ws || !m0 || !m1 || !ms0 || !ms1) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0) return LG_ERR_SHAPE;
  const auto* rowkey = static_cast<const unsigned long long*>(keys_ws);
  const auto* colkey = rowkey + static_cast<size_t>(B) * M;
  const int pts = M > N ? M : N;
  lg::filter_finalize_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(rowkey, colkey, m_len, n_len, th, m0, m1,
                                                                            ms0, ms1, M, N);
  return static_cast<int>(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------------------------
// lg_filter_matches_static: lg_filter_matches for callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  The row / column key scratch lives in a file-scope device array instead of a
// caller-supplied workspace.  No counterpart in the reference.
//   * limit: B * (M + N) <= LG_FILTER_STATIC_MAX_KEYS (1 MiB of keys, held by every GPU that loads the library);
//   * the array is shared by every call on a device: it is cleared on the stream at the start of each call, and
//     calls that overlap on the GPU (different streams) race on it;
//   * the host-side name of the array is not a device address: the address is resolved once per device with
//     cudaGetSymbolAddress (that device current) and cached.
namespace lg {
static __device__ __align__(256) unsigned long long g_filter_keys[LG_FILTER_STATIC_MAX_KEYS];
}  // namespace lg

extern "C" int lg_filter_matches_static(const float* scores, int64_t pitch, const int* m_len, const int* n_len,
                                        float th, int64_t* m0, int64_t* m1, float* ms0, float* ms1, int B, int M,
                                        int N, cudaStream_t stream) {
  if (!scores || !m0 || !m1 || !ms0 || !ms1) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0 || pitch < N) return LG_ERR_SHAPE;
  const size_t keys = static_cast<size_t>(B) * (static_cast<size_t>(M) + N);
  if (keys > static_cast<size_t>(LG_FILTER_STATIC_MAX_KEYS)) return LG_ERR_TOO_LARGE;
  static unsigned long long* dev_ptr[lg::kMaxDevices] = {};  // per device; a benign race resolves the same address twice
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= lg::kMaxDevices) return LG_ERR_SHAPE;
  if (!dev_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, lg::g_filter_keys);  // the current device's instance of the array
    if (e != cudaSuccess) return static_cast<int>(e);
    if (!p) return LG_ERR_NULL;
    dev_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  unsigned long long* rowkey = dev_ptr[dev];
  unsigned long long* colkey = rowkey + static_cast<size_t>(B) * M;
  e = cudaMemsetAsync(rowkey, 0, keys * sizeof(unsigned long long), stream);  // the last call's keys are still there
  if (e != cudaSuccess) return static_cast<int>(e);
  const int strips = (M + 31) / 32, nchunks = (N + 31) / 32;
  int nsplit = (148 * 16 + B * strips - 1) / (B * strips);
  nsplit = max(1, min(nsplit, nchunks));
  const int col_groups = (nsplit + lg::kFilterWarps - 1) / lg::kFilterWarps;
  lg::filter_scan_kernel<<<dim3(strips, col_groups, B), lg::kFilterWarps * 32, 0, stream>>>(
      scores, pitch, static_cast<int64_t>(M + 1) * pitch, m_len, n_len, rowkey, colkey, M, N, nsplit);
  e = cudaGetLastError();
  if (e != cudaSuccess) return static_cast<int>(e);
  const int pts = M > N ? M : N;
  lg::filter_finalize_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(rowkey, colkey, m_len, n_len, th, m0, m1,
                                                                            ms0, ms1, M, N);
  return static_cast<int>(cudaGetLastError());
}
