// lg_attention : tcgen05 / TMEM flash attention for LightGlue's SelfBlock and CrossBlock.
//
// Replaces Attention.forward (reference gluefactory/models/matchers/lightglue.py:139-156),
// i.e. softmax(q k^T * dh^-0.5 [+ key-padding mask]) v, without materialising the (N x N)
// score matrix.  For CrossBlock (lightglue.py:229-244) it is called once per direction on
// the shared to_qk projection (qk0 against qk1/v1, then qk1 against qk0/v0).
//
// Shape of one CTA (384 threads, 1 CTA / SM; setmaxnreg moves registers to the softmax warpgroups):
//   * 256 query rows = two 128-row tiles, one per softmax warpgroup (warps 0-3, 4-7);
//     both tiles share every staged K / V^T tile.
//   * warp 8 (one lane): bulk (TMA engine) copies of the pre-swizzled bf16 operand tiles
//     into a 3-stage shared-memory ring, mbarrier complete_tx signalling.
//   * warp 9 (one lane): tcgen05.mma issue.  S_t = Q_t K^T (128x128x64) and
//     O_t += P_t V (128x64x128) with accumulators in TMEM: S0 S1 O0 O1 = 384 columns.
//   * softmax warpgroups: thread == row.  tcgen05.ld S from TMEM, fp32 online softmax in the
//     exp2 domain with lazy (thresholded) rescaling of O, P written as bf16 into a
//     SWIZZLE_128B shared tile that feeds the second MMA.
// The MMA order  QK0_{j+1}, PV0_j, QK1_{j+1}, PV1_j  lets each warpgroup's softmax overlap the
// other tile's tensor work.
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kStages = 4;
constexpr int kTileBytes = 16384;
constexpr int kAttnThreads = 384;  // 2 softmax warpgroups + 1 warpgroup holding the producer and MMA warps
constexpr uint32_t kColS = 0, kColO = 384;  // TMEM column map: 3 rotating S buffers at i*128, O_t at 384 + t*64
constexpr float kRescaleThreshold = 8.0f;   // log2 units; P stays <= 2^8

__device__ __forceinline__ float ex2_approx(float x) {
#ifdef LG_NOEXP
  return x;  // timing experiment only
#endif
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T : A = bf16 rows packed two per 32-bit column (lane = row), issued by ONE thread
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
template <int N>
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

#ifdef LG_ATTN_TRACE
__device__ unsigned long long g_trace[8192];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TRACE(slot, idx) do { if (blockIdx.x == 1 && blockIdx.y == 1 && blockIdx.z == 1 && (idx) < 256) g_trace[(slot) * 256 + (idx)] = gtime(); } while (0)
#else
#define TRACE(slot, idx) do {} while (0)
#endif

struct AttnParams {
  const uint8_t* q_tiles;
  const uint8_t* k_tiles;
  const uint8_t* v_tiles;
  float* out;          // (B, nq, H*64) fp32, or null when out_tiles is set
  uint8_t* out_tiles;  // split tiles [rt][H atoms][hi 16 KB | lo 16 KB]: context as the A operand of lg_linear
  int out_rt_base;     // row tile of (b=0, query tile 0) inside out_tiles; image b starts at + b*tq_pad
  const int* q_len;    // per-image valid query rows, or null
  const int* k_len;    // per-image valid keys, or null
  int B, H, nq, nk, tq_pad, tk_pad;
  float scale_log2;    // dh^-0.5 * log2(e)
};

struct __align__(8) AttnBarriers {
  uint64_t q_full;
  uint64_t kv_full[kStages];
  uint64_t kv_empty[kStages];
  uint64_t s_full[3];
  uint64_t p_full[2];
  uint64_t pv_done[2];
  uint64_t o_final;
  uint32_t tmem_base;
};

constexpr int kAttnSmem = 1024 /*align slack*/ + 2 * kTileBytes + kStages * 2 * kTileBytes +
                          static_cast<int>(sizeof(AttnBarriers));

__global__ void __launch_bounds__(kAttnThreads, 1) attention_kernel(const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = base;                               // 2 x 16 KB
  uint8_t* sKV = sQ + 2 * kTileBytes;               // kStages x (K 16 KB | V^T 16 KB)
  AttnBarriers* bar = reinterpret_cast<AttnBarriers*>(sKV + kStages * 2 * kTileBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, qt0 = blockIdx.x * 2;
  const int q_len = p.q_len ? min(p.q_len[b], p.nq) : p.nq;
  const int k_len = p.k_len ? min(p.k_len[b], p.nk) : p.nk;
  const int nkt = (k_len + kTile - 1) / kTile;
  const int row_stride = p.H * kHeadDim;

  if (qt0 * kTile >= q_len || nkt == 0) {
    // nothing to attend from / to: rows are zero (reference: nan_to_num of an all-masked softmax)
    if (p.out_tiles) {
      for (int i = threadIdx.x; i < 2 * 32768 / 16; i += kAttnThreads) {
        const int t = i / 2048, o = i % 2048;
        uint8_t* dst = p.out_tiles + ((static_cast<size_t>(p.out_rt_base + b * p.tq_pad + qt0 + t) * p.H + h) * 32768);
        reinterpret_cast<uint4*>(dst)[o] = make_uint4(0u, 0u, 0u, 0u);
      }
      return;
    }
    const int r0 = qt0 * kTile, r1 = min(p.nq, r0 + 2 * kTile);
    for (int i = threadIdx.x; i < (r1 - r0) * (kHeadDim / 4); i += kAttnThreads) {
      const int r = r0 + i / (kHeadDim / 4), c4 = i % (kHeadDim / 4);
      reinterpret_cast<float4*>(p.out + (static_cast<size_t>(b) * p.nq + r) * row_stride + h * kHeadDim)[c4] =
          make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }

  if (threadIdx.x == 0) {
    mbar_init(&bar->q_full, 1);
    for (int s = 0; s < kStages; ++s) { mbar_init(&bar->kv_full[s], 1); mbar_init(&bar->kv_empty[s], 1); }
    for (int t = 0; t < 3; ++t) mbar_init(&bar->s_full[t], 1);
    for (int t = 0; t < 2; ++t) { mbar_init(&bar->p_full[t], 128); mbar_init(&bar->pv_done[t], 1); }
    mbar_init(&bar->o_final, 1);
    fence_mbar_init();
  }
  if (warp == 9) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp >= 8) {
  reg_dec<88>();  // pool = 384 x 168: 128*(208+208+88) registers
  if (warp == 8) {
    // ------------------------------------------------------------------ producer
    if (lane == 0) {
      const size_t bh = static_cast<size_t>(b) * p.H + h;
      const uint8_t* gq = p.q_tiles + (bh * p.tq_pad + qt0) * kTileBytes;
      const uint8_t* gk = p.k_tiles + bh * p.tk_pad * kTileBytes;
      const uint8_t* gv = p.v_tiles + bh * p.tk_pad * kTileBytes;
      mbar_arrive_expect_tx(&bar->q_full, 2 * kTileBytes);
      bulk_g2s(sQ, gq, 2 * kTileBytes, &bar->q_full);  // tiles qt0, qt0+1 are adjacent (tq_pad is even)
      for (int j = 0; j < nkt; ++j) {
        const int s = j % kStages;
        if (j >= kStages) mbar_wait(&bar->kv_empty[s], ((j / kStages) - 1) & 1);
        mbar_arrive_expect_tx(&bar->kv_full[s], 2 * kTileBytes);
        bulk_g2s(sKV + s * 2 * kTileBytes, gk + static_cast<size_t>(j) * kTileBytes, kTileBytes, &bar->kv_full[s]);
        bulk_g2s(sKV + s * 2 * kTileBytes + kTileBytes, gv + static_cast<size_t>(j) * kTileBytes, kTileBytes,
                 &bar->kv_full[s]);
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------------ MMA issuer
    {
      const bool leader = elect_one();  // all lanes run the loop (uniform operands); one lane issues
      constexpr uint32_t idesc_qk = umma_idesc(kFmtBF16, 128, 128);
      constexpr uint32_t idesc_pv = umma_idesc(kFmtBF16, 128, 64);
      const uint32_t aQ = smem_u32(sQ), aKV = smem_u32(sKV);
      // step n = 2j + t (key block j, query tile t) owns S buffer n % 3: QK(n+3) is queued as soon as the
      // sweep of step n has drained that buffer, i.e. 1.5 key blocks ahead of the softmax that consumes it.
      auto issue_qk = [&](int n) {
        const int t = n & 1, s = (n >> 1) % kStages, sb = n % 3;
        const uint64_t da = umma_desc_kmajor_sw128(aQ + t * kTileBytes);
        const uint64_t db = umma_desc_kmajor_sw128(aKV + s * 2 * kTileBytes);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)  // K = 64 = 4 x 16, 32 B per step inside the 128 B swizzle row
          if (leader) umma_bf16_ss(tmem + kColS + sb * 128, da + 2 * kk, db + 2 * kk, idesc_qk, kk > 0);
        if (leader) umma_commit(&bar->s_full[sb]);
      };
      auto issue_pv = [&](int n) {
        // P(n) sits in the first 64 columns of S buffer n % 3 as packed bf16 (A operand from TMEM)
        const int t = n & 1, j = n >> 1, s = j % kStages;
        const uint32_t tP = tmem + kColS + (n % 3) * 128;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {  // K = 128 keys = 8 x 16: 8 TMEM columns / 32 B of a V^T row per step
          const uint32_t offB = kTileBytes + (kk >> 2) * (kTileBytes / 2) + (kk & 3) * 32;
          if (leader)
            umma_bf16_ts(tmem + kColO + t * 64, tP + kk * 8, umma_desc_kmajor_sw128(aKV + s * 2 * kTileBytes + offB),
                         idesc_pv, (j > 0 || kk > 0));
        }
      };
      const int steps = 2 * nkt;
      mbar_wait(&bar->q_full, 0);
      mbar_wait(&bar->kv_full[0], 0);
      tc_fence_after();
      issue_qk(0);
      issue_qk(1);
      if (steps > 2) { mbar_wait(&bar->kv_full[1], 0); tc_fence_after(); issue_qk(2); }
      for (int n = 0; n < steps; ++n) {
        const int t = n & 1, j = n >> 1, s = j % kStages;
        mbar_wait(&bar->p_full[t], j & 1);  // P_t(j) staged AND its S buffer drained
        tc_fence_after();
        if (leader) TRACE(0, n);
        issue_pv(n);  // QK(n+3) below reuses this S buffer: same-thread MMAs execute in order
        if (leader) umma_commit(&bar->pv_done[t]);
        if (leader && t == 1) umma_commit(&bar->kv_empty[s]);  // K_j and V_j are dead once PV1_j retires
        const int n3 = n + 3;
        if (n3 < steps) {
          if ((n3 & 1) == 0) {  // first use of key block n3/2
            const int j3 = n3 >> 1;
            mbar_wait(&bar->kv_full[j3 % kStages], (j3 / kStages) & 1);
            tc_fence_after();
          }
          issue_qk(n3);
        }
        if (leader) TRACE(1, n);
      }
      if (leader) umma_commit(&bar->o_final);
    }
  }
  } else {
    reg_inc<208>();
    // ------------------------------------------------------------------ softmax warpgroups
    const int t = warp >> 2;                       // query tile of this warpgroup
    const int r = (warp & 3) * 32 + lane;          // row == TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS0 = tmem + lane_addr + kColS;
    const uint32_t tO = tmem + lane_addr + kColO + t * 64;
    const float c = p.scale_log2;
    float m_used = -INFINITY, l = 0.f;
    uint32_t va[32], vb[32];
    uint32_t pk0[32], pk1[32];  // P of the block as packed bf16 pairs: 64 TMEM columns
    const float2 c2 = make_float2(c, c);

    for (int j = 0; j < nkt; ++j) {
      const int rem = k_len - j * kTile;           // valid keys in this block (>= 1)
      const bool partial = rem < kTile;
      const int n = 2 * j + t;
      const uint32_t tS = tS0 + (n % 3) * 128;
      if (lane == 0 && (warp & 3) == 0) TRACE(2 + t * 3, j);
      mbar_wait(&bar->s_full[n % 3], (n / 3) & 1);
      tc_fence_after();
      if (lane == 0 && (warp & 3) == 0) TRACE(3 + t * 3, j);
      if (j == 0) {
        // first block: the reference max is not known yet -> one max-only sweep
        float mx = -INFINITY;
#pragma unroll
        for (int cb = 0; cb < 4; ++cb) {
          tmem_ld32(tS + cb * 32, va);
          tmem_wait_ld();
          float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const bool k0 = !partial || cb * 32 + i < rem, k1 = !partial || cb * 32 + i + 1 < rem;
            const bool k2 = !partial || cb * 32 + i + 2 < rem, k3 = !partial || cb * 32 + i + 3 < rem;
            m0 = fmax3(m0, k0 ? __uint_as_float(va[i]) : -INFINITY, k1 ? __uint_as_float(va[i + 1]) : -INFINITY);
            m1 = fmax3(m1, k2 ? __uint_as_float(va[i + 2]) : -INFINITY, k3 ? __uint_as_float(va[i + 3]) : -INFINITY);
          }
          mx = fmax3(mx, m0, m1);
        }
        m_used = mx;
      }
      // Single sweep: P = exp2(S*c - m_used*c) with the (possibly stale) reference max of the previous
      // blocks while the block max is gathered on the side.  Only when some row of the warp outgrew its
      // reference by more than 2^kRescaleThreshold is O rescaled and the sweep repeated (rare after the
      // first blocks) -- P <= 2^8 otherwise, harmless for the fp32 row sums and the bf16 P operand.
      float tile_sum = 0.f;
#pragma unroll 1
      for (int attempt = 0; attempt < 2; ++attempt) {
        const float ms = m_used * c;
        const float2 nms2 = make_float2(-ms, -ms);
        float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
        float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
        auto sweep = [&](uint32_t(&v)[32], const int cb) {
          if (partial) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (cb * 32 + i >= rem) v[i] = 0xff800000u;  // -inf: exp2 -> 0, ignored by max
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float e[8];
#pragma unroll
            for (int i = 0; i < 8; i += 4) {
              const float x0 = __uint_as_float(v[8 * q + i]), x1 = __uint_as_float(v[8 * q + i + 1]);
              const float x2 = __uint_as_float(v[8 * q + i + 2]), x3 = __uint_as_float(v[8 * q + i + 3]);
              if (i == 0) { m0 = fmax3(m0, x0, x1); m1 = fmax3(m1, x2, x3); }
              else { m2 = fmax3(m2, x0, x1); m3 = fmax3(m3, x2, x3); }
              const float2 t0 = __ffma2_rn(make_float2(x0, x1), c2, nms2);
              const float2 t1 = __ffma2_rn(make_float2(x2, x3), c2, nms2);
              e[i] = ex2_approx(t0.x); e[i + 1] = ex2_approx(t0.y);
              e[i + 2] = ex2_approx(t1.x); e[i + 3] = ex2_approx(t1.y);
              s0 = __fadd2_rn(s0, make_float2(e[i], e[i + 1]));
              s1 = __fadd2_rn(s1, make_float2(e[i + 2], e[i + 3]));
            }
            uint4 o;
            o.x = pack_bf16x2(e[0], e[1]); o.y = pack_bf16x2(e[2], e[3]);
            o.z = pack_bf16x2(e[4], e[5]); o.w = pack_bf16x2(e[6], e[7]);
            uint32_t(&pk)[32] = cb < 2 ? pk0 : pk1;  // cb, q are compile-time after unrolling
            pk[(cb & 1) * 16 + q * 4] = o.x; pk[(cb & 1) * 16 + q * 4 + 1] = o.y;
            pk[(cb & 1) * 16 + q * 4 + 2] = o.z; pk[(cb & 1) * 16 + q * 4 + 3] = o.w;
          }
        };
        // TMEM loads run one 32-column chunk ahead of the arithmetic
        tmem_ld32(tS, va);
        tmem_wait_ld();
        tmem_ld32(tS + 32, vb);
        sweep(va, 0);
        tmem_wait_ld();
        tmem_ld32(tS + 64, va);
        sweep(vb, 1);
        tmem_wait_ld();
        tmem_ld32(tS + 96, vb);
        sweep(va, 2);
        tmem_wait_ld();
        sweep(vb, 3);
        tile_sum = (s0.x + s0.y) + (s1.x + s1.y);
        if (attempt == 1) break;
        const float mx = fmaxf(fmax3(m0, m1, m2), m3);
        const bool need = (mx - m_used) * c > kRescaleThreshold;
        if (!__any_sync(0xffffffffu, need)) break;
        float alpha = 1.f;
        if (need) { alpha = ex2_approx((m_used - mx) * c); m_used = mx; }
        l *= alpha;
        if (j > 0) {
          // O_t must be quiescent: PV_t(j-1) retired.  The barrier is at most one phase ahead of (j-1), so a
          // parity wait is valid although earlier phases were never waited for.
          mbar_wait(&bar->pv_done[t], (j - 1) & 1);
          tc_fence_after();
#pragma unroll
          for (int cb = 0; cb < 2; ++cb) {
            tmem_ld32(tO + cb * 32, va);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) va[i] = __float_as_uint(__uint_as_float(va[i]) * alpha);
            tmem_st32(tO + cb * 32, va);
          }
          tmem_wait_st();
        }
      }
      l += tile_sum;
      // P overwrites the first 64 columns of its own S buffer (every S value has been consumed); the PV MMA reads
      // it from TMEM as the A operand, then QK(n+3) recycles the buffer.
      tmem_st32(tS, pk0);
      tmem_st32(tS + 32, pk1);
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(&bar->p_full[t]);
      if (lane == 0 && (warp & 3) == 0) TRACE(4 + t * 3, j);
    }
    // epilogue: O / l -> fp32 rows of the (B, nq, H*64) context tensor
    mbar_wait(&bar->o_final, 0);
    tc_fence_after();
    const int row = (qt0 + t) * kTile + r;
    const float inv = (row < q_len) ? 1.f / l : 0.f;  // padded query rows -> 0
    if (p.out_tiles) {
      // context rows as split bf16 tiles: head h is K atom h of the (rows x 256) matrix
      uint8_t* dt = p.out_tiles + ((static_cast<size_t>(p.out_rt_base + b * p.tq_pad + qt0 + t) * p.H + h) * 32768);
#pragma unroll
      for (int cb = 0; cb < 2; ++cb) {
        tmem_ld32(tO + cb * 32, va);
        tmem_wait_ld();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t hh[4], ll[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float x0 = __uint_as_float(va[8 * q + 2 * i]) * inv, x1 = __uint_as_float(va[8 * q + 2 * i + 1]) * inv;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
            hh[i] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
            ll[i] = pack_bf16x2(x0 - __bfloat162float(h0), x1 - __bfloat162float(h1));
          }
          uint8_t* at = dt + sw128_offset(r, cb * 4 + q);
          *reinterpret_cast<uint4*>(at) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
          *reinterpret_cast<uint4*>(at + 16384) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
        }
      }
    } else {
    float4* dst = reinterpret_cast<float4*>(p.out + (static_cast<size_t>(b) * p.nq + row) * row_stride + h * kHeadDim);
#pragma unroll
    for (int cb = 0; cb < 2; ++cb) {
      tmem_ld32(tO + cb * 32, va);
      tmem_wait_ld();
      if (row < p.nq) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          dst[cb * 8 + i] = make_float4(__uint_as_float(va[4 * i]) * inv, __uint_as_float(va[4 * i + 1]) * inv,
                                        __uint_as_float(va[4 * i + 2]) * inv, __uint_as_float(va[4 * i + 3]) * inv);
      }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

}  // namespace lg

static int launch_attention(const void* q_tiles, const void* k_tiles, const void* v_tiles, float* out, void* out_tiles,
                            int out_rt_base, const int* q_len, const int* k_len, int B, int H, int nq, int nk,
                            int tq_pad, int tk_pad, float softmax_scale, cudaStream_t stream) {
  if (!q_tiles || !k_tiles || !v_tiles || (!out && !out_tiles)) return LG_ERR_NULL;
  if (B <= 0 || H <= 0 || nq <= 0 || nk <= 0 || out_rt_base < 0) return LG_ERR_SHAPE;
  if (tq_pad * lg::kTile < nq || tk_pad * lg::kTile < nk || (tq_pad & 1) || (tk_pad & 1)) return LG_ERR_SHAPE;
  static bool attr_done = false;  // idempotent attribute; a benign race at worst sets it twice
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(lg::attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         lg::kAttnSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  lg::AttnParams p{static_cast<const uint8_t*>(q_tiles), static_cast<const uint8_t*>(k_tiles),
                   static_cast<const uint8_t*>(v_tiles), out, static_cast<uint8_t*>(out_tiles), out_rt_base, q_len,
                   k_len, B, H, nq, nk, tq_pad, tk_pad, softmax_scale * 1.4426950408889634f};
  lg::attention_kernel<<<dim3(tq_pad / 2, H, B), lg::kAttnThreads, lg::kAttnSmem, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_attention(const void* q_tiles, const void* k_tiles, const void* v_tiles, float* out,
                            const int* q_len, const int* k_len, int B, int H, int nq, int nk, int tq_pad, int tk_pad,
                            float softmax_scale, cudaStream_t stream) {
  if (!out) return LG_ERR_NULL;
  return launch_attention(q_tiles, k_tiles, v_tiles, out, nullptr, 0, q_len, k_len, B, H, nq, nk, tq_pad, tk_pad,
                          softmax_scale, stream);
}

extern "C" int lg_attention_tiles(const void* q_tiles, const void* k_tiles, const void* v_tiles, void* out_tiles,
                                  int out_rt_base, const int* q_len, const int* k_len, int B, int H, int nq, int nk,
                                  int tq_pad, int tk_pad, float softmax_scale, cudaStream_t stream) {
  if (!out_tiles) return LG_ERR_NULL;
  return launch_attention(q_tiles, k_tiles, v_tiles, nullptr, out_tiles, out_rt_base, q_len, k_len, B, H, nq, nk,
                          tq_pad, tk_pad, softmax_scale, stream);
}

#ifdef LG_ATTN_TRACE
extern "C" int lg_debug_attn_trace(unsigned long long* host_out, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host_out, lg::g_trace, sizeof(unsigned long long) * n));
}
#endif
