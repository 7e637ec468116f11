// lg_fp32 : the fp32 path (b200.precision = "fp32") -- every product and sum in IEEE fp32 on the CUDA cores, for
// the north-star's 1e-4 bound on log_assignment against the reference's fp32 CPU run.
//
// The throughput path rounds attention operands to bf16 and splits the linears into 3 bf16 terms (2e-2 bound,
// measured 1.3e-3); its tensor-core arithmetic cannot reach 1e-4 (3-term split bf16 alone leaves 2.1e-4, DESIGN.md
// section 4).  This file is the accuracy mode: a strided-batched FMA GEMM with fused bias / residual epilogue
// (Linear layers lightglue.py:168-176,184,203-211,223-228,310,370; q.k^T and P.v of Attention.forward :139-156 and of
// CrossBlock.forward :235-244), a masked row softmax (:147-149,236-243), the rotary apply (:36-43,187-188) and
// LayerNorm + GELU (:173-174).  Scores are materialised per (pair, head) like the reference's own math branch.
#include <cstdlib>
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

struct GemmF32Params {
  const float* A;      // (M, K) row-major, leading dimension lda
  const float* B;      // b_kn == 0: (N, K) row-major (C = A . B^T); b_kn == 1: (K, N) row-major (C = A . B)
  float* C;            // (M, N), leading dimension ldc
  const float* bias;   // (N) or null
  const float* resid;  // (M, N) with leading dimension ldr, or null (may alias C)
  int M, N, K;
  int64_t lda, ldb, ldc, ldr;
  int b_kn;
  float alpha;         // C = alpha * A.op(B) + bias + resid
  int inner;           // batch index z -> (z / inner, z % inner); offsets = outer * s?o + inner * s?i (elements)
  int64_t sAo, sAi, sBo, sBi, sCo, sCi, sRo, sRi;
  const int* m_len;    // optional valid rows per OUTER batch index (rows >= m_len are skipped)
  const int* k_len;    // optional valid K per OUTER batch index (b_kn == 1: keys beyond the count are not read)
};

constexpr int kGT = 64;   // output tile
constexpr int kGK = 16;   // K step

// 256 threads, 4x4 outputs per thread (packed FFMA2), operands staged in shared memory as [k][row] (float4 reads).
__global__ void __launch_bounds__(256) gemm_f32_kernel(const GemmF32Params p) {
  __shared__ __align__(16) float sA[kGK][kGT + 4];
  __shared__ __align__(16) float sB[kGK][kGT + 4];
  const int z = blockIdx.z, zo = z / p.inner, zi = z % p.inner;
  const float* A = p.A + zo * p.sAo + zi * p.sAi;
  const float* B = p.B + zo * p.sBo + zi * p.sBi;
  float* C = p.C + zo * p.sCo + zi * p.sCi;
  const float* R = p.resid ? p.resid + zo * p.sRo + zi * p.sRi : nullptr;
  const int M = p.m_len ? min(p.m_len[zo], p.M) : p.M;
  const int K = p.k_len ? min(p.k_len[zo], p.K) : p.K;
  const int row0 = blockIdx.y * kGT, col0 = blockIdx.x * kGT;
  if (row0 >= M) return;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float2 acc[4][2];  // f32x2 pairs along N: packed FFMA2, IEEE per element
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = make_float2(0.f, 0.f);
  for (int k0 = 0; k0 < K; k0 += kGK) {
    {  // A tile: 64 rows x 16 k, thread -> (row = t / 4, k = (t % 4) * 4 .. +3)
      const int r = threadIdx.x >> 2, kq = (threadIdx.x & 3) * 4;
      const int gr = row0 + r;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (gr < M) {
        const float* src = A + static_cast<int64_t>(gr) * p.lda + k0 + kq;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (k0 + kq + i < K) v[i] = src[i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) sA[kq + i][r] = v[i];
    }
    if (!p.b_kn) {  // B as (N, K): same access shape as A
      const int r = threadIdx.x >> 2, kq = (threadIdx.x & 3) * 4;
      const int gc = col0 + r;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (gc < p.N) {
        const float* src = B + static_cast<int64_t>(gc) * p.ldb + k0 + kq;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (k0 + kq + i < K) v[i] = src[i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) sB[kq + i][r] = v[i];
    } else {        // B as (K, N): thread -> (k = t / 16, n = (t % 16) * 4 .. +3)
      const int kk = threadIdx.x >> 4, nq = (threadIdx.x & 15) * 4;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (k0 + kk < K) {
        const float* src = B + static_cast<int64_t>(k0 + kk) * p.ldb + col0 + nq;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (col0 + nq + i < p.N) v[i] = src[i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) sB[kk][nq + i] = v[i];
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kGK; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(&sA[kk][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&sB[kk][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float2 b0 = make_float2(b4.x, b4.y), b1 = make_float2(b4.z, b4.w);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 aa = make_float2(a[i], a[i]);
        acc[i][0] = __ffma2_rn(aa, b0, acc[i][0]);
        acc[i][1] = __ffma2_rn(aa, b1, acc[i][1]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gr = row0 + ty * 4 + i;
    if (gr >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gc = col0 + tx * 4 + j;
      if (gc >= p.N) continue;
      float v = p.alpha * ((j & 1) ? acc[i][j >> 1].y : acc[i][j >> 1].x);
      if (p.bias) v += p.bias[gc];
      if (R) v += R[static_cast<int64_t>(gr) * p.ldr + gc];
      C[static_cast<int64_t>(gr) * p.ldc + gc] = v;
    }
  }
}

// The large-problem variant: 128 x BN output tile (BN = 128, or 64 for the P.v products whose N is the head dim),
// 8 x (BN/16) outputs per thread as 4x4 blocks (float4 shared-memory reads: 4 LDS.128 per 64 FMAs), K step 16 with
// the next step's operands prefetched into registers under the FMAs and two shared-memory buffers (one barrier per
// step), 16-byte global loads / stores where the leading dimensions and base addresses allow (vec flags), scalar
// guarded accesses otherwise.  Every accumulator sums its products in ascending k with fmaf: bit-identical to the
// 64 x 64 kernel above.
constexpr int kBM = 128;
template <int BN>
__global__ void __launch_bounds__(256, 2) gemm_f32_big_kernel(const GemmF32Params p, const int vec_a, const int vec_b,
                                                              const int vec_c) {
  constexpr int TN = BN / 16;  // outputs per thread along N and B-operand floats per thread per K step
  __shared__ __align__(16) float sA[2][kGK][kBM + 4];
  __shared__ __align__(16) float sB[2][kGK][BN + 4];
  const int z = blockIdx.z, zo = z / p.inner, zi = z % p.inner;
  const float* A = p.A + zo * p.sAo + zi * p.sAi;
  const float* B = p.B + zo * p.sBo + zi * p.sBi;
  float* C = p.C + zo * p.sCo + zi * p.sCi;
  const float* R = p.resid ? p.resid + zo * p.sRo + zi * p.sRi : nullptr;
  const int M = p.m_len ? min(p.m_len[zo], p.M) : p.M;
  const int K = p.k_len ? min(p.k_len[zo], p.K) : p.K;
  const int row0 = blockIdx.y * kBM, col0 = blockIdx.x * BN;
  if (row0 >= M) return;
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;

  // loader coordinates
  const int a_row = t >> 1, a_kq = (t & 1) * 8;
  const int b_row = BN == 128 ? (t >> 1) : (t >> 2), b_kq = BN == 128 ? (t & 1) * 8 : (t & 3) * 4;  // (N, K) layout
  const int bk_k = t >> 4, bk_n = (t & 15) * TN;                                                    // (K, N) layout
  float ra[8], rb[TN];

  auto load_run = [&](const float* src, int n, int valid, bool vec, float* dst) {
    // n floats from src, the first `valid` of them real (the rest 0)
    if (vec && valid >= n) {
#pragma unroll
      for (int q = 0; q < n / 4; ++q) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(src) + q);
        dst[4 * q] = v.x; dst[4 * q + 1] = v.y; dst[4 * q + 2] = v.z; dst[4 * q + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < n; ++i) dst[i] = i < valid ? __ldg(src + i) : 0.f;
    }
  };
  auto load_tiles = [&](int k0) {
    {
      const int gr = row0 + a_row;
      const int valid = gr < M ? max(0, min(8, K - (k0 + a_kq))) : 0;
      load_run(A + static_cast<int64_t>(gr) * p.lda + k0 + a_kq, 8, valid, vec_a, ra);
    }
    if (!p.b_kn) {
      const int gc = col0 + b_row;
      const int valid = gc < p.N ? max(0, min(TN, K - (k0 + b_kq))) : 0;
      load_run(B + static_cast<int64_t>(gc) * p.ldb + k0 + b_kq, TN, valid, vec_b, rb);
    } else {
      const int valid = k0 + bk_k < K ? max(0, min(TN, p.N - (col0 + bk_n))) : 0;
      load_run(B + static_cast<int64_t>(k0 + bk_k) * p.ldb + col0 + bk_n, TN, valid, vec_b, rb);
    }
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 8; ++i) sA[buf][a_kq + i][a_row] = ra[i];
    if (!p.b_kn) {
#pragma unroll
      for (int i = 0; i < TN; ++i) sB[buf][b_kq + i][b_row] = rb[i];
    } else {
#pragma unroll
      for (int q = 0; q < TN / 4; ++q)
        *reinterpret_cast<float4*>(&sB[buf][bk_k][bk_n + 4 * q]) = make_float4(rb[4 * q], rb[4 * q + 1], rb[4 * q + 2], rb[4 * q + 3]);
    }
  };

  // accumulators as f32x2 pairs along N: one packed FFMA2 (fma.rn.f32x2, IEEE per element) per two outputs halves the
  // issue slots of the inner loop (the scalar-FFMA version was issue bound: FMA pipe 55 % active at 65 % issue)
  float2 acc[8][TN / 2];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN / 2; ++j) acc[i][j] = make_float2(0.f, 0.f);

  load_tiles(0);
  store_tiles(0);
  __syncthreads();
  int buf = 0;
  for (int k0 = 0; k0 < K; k0 += kGK, buf ^= 1) {
    const bool more = k0 + kGK < K;
    if (more) load_tiles(k0 + kGK);
#pragma unroll
    for (int kk = 0; kk < kGK; ++kk) {
      float a[8];
      float2 b[TN / 2];
      const float4 a0 = *reinterpret_cast<const float4*>(&sA[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&sA[buf][kk][64 + ty * 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      const float4 b0 = *reinterpret_cast<const float4*>(&sB[buf][kk][tx * 4]);
      b[0] = make_float2(b0.x, b0.y); b[1] = make_float2(b0.z, b0.w);
      if (BN == 128) {
        const float4 b1 = *reinterpret_cast<const float4*>(&sB[buf][kk][64 + tx * 4]);
        b[TN / 2 - 2] = make_float2(b1.x, b1.y); b[TN / 2 - 1] = make_float2(b1.z, b1.w);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 aa = make_float2(a[i], a[i]);
#pragma unroll
        for (int j = 0; j < TN / 2; ++j) acc[i][j] = __ffma2_rn(aa, b[j], acc[i][j]);
      }
    }
    if (more) store_tiles(buf ^ 1);
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int gr = row0 + (i >> 2) * 64 + ty * 4 + (i & 3);
    if (gr >= M) continue;
#pragma unroll
    for (int jq = 0; jq < TN / 4; ++jq) {
      const int gc = col0 + jq * 64 + tx * 4;
      if (gc >= p.N) continue;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = p.alpha * ((j & 1) ? acc[i][jq * 2 + (j >> 1)].y : acc[i][jq * 2 + (j >> 1)].x);
      float* dst = C + static_cast<int64_t>(gr) * p.ldc + gc;
      if (vec_c && gc + 3 < p.N) {
        if (p.bias) {
          const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias + gc));
          v[0] += bb.x; v[1] += bb.y; v[2] += bb.z; v[3] += bb.w;
        }
        if (R) {
          const float4 rr = *reinterpret_cast<const float4*>(R + static_cast<int64_t>(gr) * p.ldr + gc);
          v[0] += rr.x; v[1] += rr.y; v[2] += rr.z; v[3] += rr.w;
        }
        *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (gc + j >= p.N) continue;
          float w = v[j];
          if (p.bias) w += p.bias[gc + j];
          if (R) w += R[static_cast<int64_t>(gr) * p.ldr + gc + j];
          dst[j] = w;
        }
      }
    }
  }
}

// In-place softmax over the first n_len[b] columns of every row (the rest -> 0); rows >= m_len[b] -> 0
// (the reference's fully masked rows are nan_to_num'd to 0, lightglue.py:145,149).  s: (B, H, M, pitch).
// one warp per row, grid (ceil(M/8), H, B)
__global__ void __launch_bounds__(256) softmax_rows_kernel(float* __restrict__ s, const int* __restrict__ m_len,
                                                           const int* __restrict__ n_len, int H, int M, int N,
                                                           int64_t pitch) {
  const int b = blockIdx.z, h = blockIdx.y, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= M) return;
  const int mv = m_len ? min(m_len[b], M) : M, nv = n_len ? min(n_len[b], N) : N;
  float* row = s + ((static_cast<int64_t>(b) * H + h) * M + r) * pitch;
  if (r >= mv || nv == 0) {
    for (int c = lane; c < pitch; c += 32) row[c] = 0.f;
    return;
  }
  float mx = -INFINITY;
  for (int c = lane; c < nv; c += 32) mx = fmaxf(mx, row[c]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int c = lane; c < nv; c += 32) {
    const float e = expf(row[c] - mx);
    row[c] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
  for (int c = lane; c < pitch; c += 32) row[c] = c < nv ? row[c] * inv : 0.f;
}

// rotary on the q and k blocks of a (B, N, ld) buffer, in place: columns [q_off + h*64, +64) and [k_off + h*64, +64)
// t' = t * cos + rotate_half(t) * sin with interleaved pairs (x0, x1) -> (-x1, x0) (lightglue.py:36-43), cos / sin
// from the compact (B, N, 32) tables (frequency j serves lanes 2j, 2j+1).  thread per (row, head, j)
__global__ void __launch_bounds__(256) rotary_rows_kernel(float* __restrict__ x, const float* __restrict__ cos_t,
                                                          const float* __restrict__ sin_t, int64_t rows, int64_t ld,
                                                          int q_off, int k_off, int H) {
  const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int j = idx & 31;
  const int h = (idx >> 5) % H;
  const int64_t r = (idx >> 5) / H;
  if (r >= rows) return;
  const float c = cos_t[r * 32 + j], s = sin_t[r * 32 + j];
#pragma unroll
  for (int part = 0; part < 2; ++part) {
    float* t = x + r * ld + (part ? k_off : q_off) + h * 64 + 2 * j;
    const float x0 = t[0], x1 = t[1];
    const float a0 = x0 * c, a1 = x1 * c;      // the reference's mul / mul / add rounding sequence
    const float b0 = -x1 * s, b1 = x0 * s;
    t[0] = a0 + b0;
    t[1] = a1 + b1;
  }
}

// LayerNorm (biased variance, eps inside the sqrt) + exact-erf GELU over rows of `width` floats, in place.
// one warp per row
__global__ void __launch_bounds__(256) ln_gelu_rows_kernel(float* __restrict__ x, const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, float eps, int64_t rows,
                                                           int width) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  float* row = x + r * width;
  float sum = 0.f;
  for (int c = lane; c < width; c += 32) sum += row[c];
  const float mean = warp_sum(sum) / width;
  float var = 0.f;
  for (int c = lane; c < width; c += 32) {
    const float d = row[c] - mean;
    var += d * d;
  }
  const float rstd = rsqrtf(warp_sum(var) / width + eps);
  for (int c = lane; c < width; c += 32) {
    const float y = (row[c] - mean) * rstd * gamma[c] + beta[c];
    row[c] = 0.5f * y * (1.f + erff(y * 0.70710678118654752f));
  }
}

}  // namespace lg

extern "C" int lg_gemm_f32(const float* A, const float* B, float* C, const float* bias, const float* resid, int M,
                           int N, int K, int64_t lda, int64_t ldb, int64_t ldc, int64_t ldr, int b_kn, float alpha,
                           int batch, int inner, const int64_t* strides, const int* m_len, const int* k_len,
                           cudaStream_t stream) {
  if (!A || !B || !C || !strides) return LG_ERR_NULL;
  if (M <= 0 || N <= 0 || K <= 0 || batch <= 0 || inner <= 0 || batch % inner) return LG_ERR_SHAPE;
  lg::GemmF32Params p{A, B, C, bias, resid, M, N, K, lda, ldb, ldc, ldr, b_kn, alpha, inner,
                      strides[0], strides[1], strides[2], strides[3], strides[4], strides[5], strides[6], strides[7],
                      m_len, k_len};
  // 128 x BN tiles whenever there is more than one 64-row tile of rows (measured faster or equal on every shape of the
  // path from one pair up, scripts/exp_gemm_f32.py); the 64 x 64 kernel serves the small remainder
  const int bn = N <= 64 ? 64 : 128;
  static const int force = [] {  // LG_GEMM_F32_TILE=small|big: experiments only (scripts/exp_gemm_f32.py)
    const char* e = getenv("LG_GEMM_F32_TILE");
    return !e ? 0 : e[0] == 's' ? 1 : e[0] == 'b' ? 2 : 0;
  }();
  if (force == 2 || (force == 0 && M > lg::kGT)) {
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    auto m4 = [](int64_t v) { return (v & 3) == 0; };
    const int vec_a = al16(A) && m4(lda) && m4(strides[0]) && m4(strides[1]);
    const int vec_b = al16(B) && m4(ldb) && m4(strides[2]) && m4(strides[3]);
    const int vec_c = al16(C) && m4(ldc) && m4(strides[4]) && m4(strides[5]) && (!bias || al16(bias)) &&
                      (!resid || (al16(resid) && m4(ldr) && m4(strides[6]) && m4(strides[7])));
    dim3 grid((N + bn - 1) / bn, (M + lg::kBM - 1) / lg::kBM, batch);
    if (grid.y > 65535 || grid.z > 65535) return LG_ERR_SHAPE;
    if (bn == 64)
      lg::gemm_f32_big_kernel<64><<<grid, 256, 0, stream>>>(p, vec_a, vec_b, vec_c);
    else
      lg::gemm_f32_big_kernel<128><<<grid, 256, 0, stream>>>(p, vec_a, vec_b, vec_c);
    return static_cast<int>(cudaGetLastError());
  }
  dim3 grid((N + lg::kGT - 1) / lg::kGT, (M + lg::kGT - 1) / lg::kGT, batch);
  if (grid.y > 65535 || grid.z > 65535) return LG_ERR_SHAPE;
  lg::gemm_f32_kernel<<<grid, 256, 0, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_softmax_rows_f32(float* s, const int* m_len, const int* n_len, int B, int H, int M, int N,
                                   int64_t pitch, cudaStream_t stream) {
  if (!s) return LG_ERR_NULL;
  if (B <= 0 || H <= 0 || M <= 0 || N <= 0 || pitch < N || B > 65535 || H > 65535) return LG_ERR_SHAPE;
  lg::softmax_rows_kernel<<<dim3((M + 7) / 8, H, B), 256, 0, stream>>>(s, m_len, n_len, H, M, N, pitch);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_rotary_rows_f32(float* x, const float* cos_t, const float* sin_t, int64_t rows, int64_t ld, int q_off,
                                  int k_off, int H, cudaStream_t stream) {
  if (!x || !cos_t || !sin_t) return LG_ERR_NULL;
  if (rows <= 0 || H <= 0 || ld < 64 * H) return LG_ERR_SHAPE;
  const int64_t threads = rows * H * 32;
  lg::rotary_rows_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(x, cos_t, sin_t, rows, ld,
                                                                                        q_off, k_off, H);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_ln_gelu_rows_f32(float* x, const float* gamma, const float* beta, float eps, int64_t rows, int width,
                                   cudaStream_t stream) {
  if (!x || !gamma || !beta) return LG_ERR_NULL;
  if (rows <= 0 || width <= 0) return LG_ERR_SHAPE;
  lg::ln_gelu_rows_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, stream>>>(x, gamma, beta, eps, rows, width);
  return static_cast<int>(cudaGetLastError());
}
