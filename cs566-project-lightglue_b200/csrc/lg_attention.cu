// lg_attention : tcgen05 / TMEM flash attention for LightGlue's SelfBlock and CrossBlock.
//
// Replaces Attention.forward (reference gluefactory/models/matchers/lightglue.py:139-156),
// i.e. softmax(q k^T * dh^-0.5 [+ key-padding mask]) v, without materialising the (N x N)
// score matrix.  For CrossBlock (lightglue.py:229-244) it is called once per direction on
// the shared to_qk projection (qk0 against qk1/v1, then qk1 against qk0/v0).
//
// Persistent kernel: one CTA per SM (384 threads) walks the work items (image b, head h, block of
// 256 query rows); neighbouring CTAs hold neighbouring query blocks of the same (b, h), so K / V^T
// tiles are shared through L2.  Roles (setmaxnreg moves registers to the softmax warpgroups):
//   * warp 8 : producer.  Bulk (TMA engine) copies of the pre-swizzled bf16 operand tiles: Q double
//     buffered per item, K | V^T in a 4-stage ring that runs across item boundaries.
//   * warp 9 : tcgen05.mma issue (all lanes run the loop so operands stay in uniform registers, one
//     elected lane issues).  S = Q_t K_j^T (128x128x64, SS) into one of THREE rotating TMEM buffers,
//     so the similarity of step n+3 is queued while the softmax of step n runs -- also across the
//     boundary to the next item, whose first similarities are ready when an item's epilogue ends;
//     O_t += P V_j (128x64x128) with P read from TMEM (TS): it overwrites the first 64 columns of
//     its own S buffer as packed bf16.  TMEM: 3 x 128 (S) + 2 x 64 (O) = 512 columns.
//   * warps 0-3 / 4-7 : softmax warpgroups, thread == row of query tile t.  One sweep per key block:
//     P = exp2(S*c - m*c) against the running reference max m of the row (exact for the first block),
//     FFMA2 / MUFU.EX2 / FADD2, TMEM loads one 32-column chunk ahead.  Overflow safety without a max
//     pass: the block's row sum bounds its largest P; only if a row sum reaches 2^60 is the block max
//     computed, O and l rescaled and the sweep repeated (rare: a logit 2^60 above the row's reference).
//     Epilogue O / l -> split bf16 tiles (or fp32 rows) overlaps the next item's first MMAs.
#include <cstdlib>
#include "lg_common.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kStages = 4;
constexpr int kTileBytes = 16384;
constexpr int kAttnThreads = 384;  // 2 softmax warpgroups + 1 warpgroup holding the producer and MMA warps
constexpr uint32_t kColS = 0, kColO = 384;  // TMEM column map: 3 rotating S buffers at i*128, O_t at 384 + t*64
#ifndef LG_SUMLIMIT
#define LG_SUMLIMIT 1.152921504606846976e18f  // 2^60
#endif
#ifndef LG_SWEEP
#define LG_SWEEP 2  // 0: pair-by-pair source order, 1: staged over the 32-column chunk, 2: software-pipelined groups
#endif
#ifndef LG_PREFETCH
#define LG_PREFETCH 0  // start loading the next block's similarity before the P hand-off of this block
#endif
#ifndef LG_WARP_ARRIVE
#define LG_WARP_ARRIVE 0  // P ready: one arrival per warp (count 4) instead of a warpgroup barrier + one arrival
#endif
#ifndef LG_ONE_LANE_WAIT
#define LG_ONE_LANE_WAIT 0  // full-warp barrier waits: one lane polls
#endif
#if LG_ONE_LANE_WAIT & 1
#define LG_WAIT_M(bar, par) mbar_wait_warp(bar, par)  // MMA warps (unused: deadlocks, see DESIGN)
#else
#define LG_WAIT_M(bar, par) mbar_wait(bar, par)
#endif
#if LG_ONE_LANE_WAIT & 2
#define LG_WAIT(bar, par) mbar_wait_warp(bar, par)  // softmax warps
#else
#define LG_WAIT(bar, par) mbar_wait(bar, par)
#endif
#ifndef LG_POLY_PAIRS
#define LG_POLY_PAIRS 2  // pairs of every 4 (8 columns) whose exp2 runs as a polynomial on the FMA pipe: 1 = 25 %, 2 = 50 %
#endif
#ifndef LG_SPLIT_POLY
#define LG_SPLIT_POLY 1  // split-row softmax: pairs of every 4 whose exp2 runs as a polynomial (0 .. 2)
#endif
#ifndef LG_SPLIT_ALT
#define LG_SPLIT_ALT 0
#endif
#ifndef LG_STAGGER
#define LG_STAGGER 0
#endif
#ifndef LG_ALTERNATE
#define LG_ALTERNATE 0  // the two softmax warpgroups take turns: a block starts when the other tile's block is half done
#endif
#ifndef LG_POLYEXP
#define LG_POLYEXP 1  // 0: none, 1: 25 %, 2: 37.5 %, 3: 50 %, 4: 75 % of the exponentials on the FMA pipe
#endif
constexpr bool kPolyExp = LG_POLYEXP != 0;
constexpr float kSumLimit = LG_SUMLIMIT;  // a block row sum below this bounds every P of the block: bf16 / fp32 hold
                                          // 2^127, so P, the row sums and O stay finite; precision is scale free

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// exp2 of two logits on the FMA / ALU pipes (MUFU.EX2 is the scarce unit at dh = 64: 16 lanes/clk/SM against
// 8192 MMA flop/clk/SM).  x = s*c - m*c is mapped to y = sat(x/252 + 1/2), which also clamps x to [-126, 126]
// (exponent field stays in 1..253; 2^126 still trips the row-sum overflow check, masked -inf gives 2^-126 ~ 0);
// n = round(x) by the magic-number add, 2^frac by a degree-3 minimax polynomial (7.5e-5 relative, P is rounded
// to bf16 = 3.9e-3), exponent inserted with an integer shift-add.
__device__ __forceinline__ float2 exp2_poly2(float sa, float sb, float cs, float bs) {
  constexpr float kMagic126 = 12582912.f - 126.f;  // 1.5 * 2^23 - 126
  float ya, yb;
  asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(ya) : "f"(sa), "f"(cs), "f"(bs));
  asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(yb) : "f"(sb), "f"(cs), "f"(bs));
  const float2 y = make_float2(ya, yb), k252 = make_float2(252.f, 252.f), km = make_float2(kMagic126, kMagic126);
  const float2 rr = __ffma2_rn(y, k252, km);                              // magic + n
  const float2 kk = __fadd2_rn(km, make_float2(-rr.x, -rr.y));            // -126 - n (exact)
  const float2 fr = __ffma2_rn(y, k252, kk);                              // x - n in [-0.5, 0.5]
  float2 pp = __ffma2_rn(make_float2(0.0551716685f, 0.0551716685f), fr, make_float2(0.2426111251f, 0.2426111251f));
  pp = __ffma2_rn(pp, fr, make_float2(0.6932609677f, 0.6932609677f));
  pp = __ffma2_rn(pp, fr, make_float2(0.9999280572f, 0.9999280572f));
  return make_float2(__int_as_float((__float_as_int(rr.x) << 23) + __float_as_int(pp.x)),
                     __int_as_float((__float_as_int(rr.y) << 23) + __float_as_int(pp.y)));
}

// D[tmem] (+)= A[tmem] * B[smem]^T : A = bf16 rows packed two per 32-bit column (lane = row)
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <int N>
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

#ifdef LG_ATTN_DBG
// progress markers into MAPPED PINNED HOST memory (survive a launch failure): 16 ints per CTA
__device__ volatile int* g_dbg = nullptr;
#define DBG(slot, val) do { if (g_dbg) { g_dbg[blockIdx.x * 16 + (slot)] = (val); __threadfence_system(); } } while (0)
#else
#define DBG(slot, val) do {} while (0)
#endif

#ifdef LG_ATTN_TRACE
__device__ unsigned long long g_trace[8192];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;  // SM cycle counter: every traced thread lives in one CTA, reading it costs a few cycles
  asm volatile("mov.u64 %0, %clock64;" : "=l"(t));
  return t;
}
#define TRACE(slot, idx) do { if (blockIdx.x == 5 && (idx) < 256) g_trace[(slot) * 256 + (idx)] = gtime(); } while (0)
#else
#define TRACE(slot, idx) do {} while (0)
#endif

// One launch serves up to two "sides" (the two images of a SelfBlock, or the two directions of a CrossBlock): half the
// launches per layer, and twice the work items per launch when the batch is small.
struct AttnSide {
  const uint8_t* q_tiles;
  const uint8_t* k_tiles;
  const uint8_t* v_tiles;
  float* out;          // (B, nq, H*64) fp32, or null when out_tiles is set
  int out_rt_base;     // row tile of (b=0, query tile 0) inside out_tiles; image b starts at + b*tq_pad
  const int* q_len;    // per-image valid query rows, or null
  const int* k_len;    // per-image valid keys, or null
  int nq, nk, tq_pad, tk_pad;
  int items;           // B * H * tq_pad / 2
};
struct AttnParams {
  AttnSide side[2];
  uint8_t* out_tiles;  // split tiles [rt][H atoms][hi 16 KB | lo 16 KB]: context as the A operand of lg_linear
  int B, H;
  float scale_log2;    // dh^-0.5 * log2(e)
  int items;           // side[0].items + side[1].items
};

struct __align__(8) AttnBarriers {
  uint64_t q_full[2];
  uint64_t q_empty[2];
  uint64_t kv_full[kStages];
  uint64_t kv_empty[kStages];
  uint64_t s_full[3];
  uint64_t p_full[4];  // [tile t][key block parity]: see the note at the arrival
  uint64_t pv_done[2];  // [tile t]: the last PV of an item retired (one completion per item)
  uint64_t t0_free[2];  // [key block parity]: PV_0(j) retired, its S buffer may be overwritten
  uint64_t o_free[2];
  uint32_t tmem_base;
  uint32_t oflag[4];  // split-row softmax: [tile][key block parity] a row sum of the block reached the limit
};

constexpr int kOutStage = 8 * 4096;  // per softmax warp: 32 rows x 128 B of the context tile, for bulk stores
constexpr int kAttnSmem = 1024 /*align slack*/ + 2 * 2 * kTileBytes + kStages * 2 * kTileBytes + kOutStage +
                          static_cast<int>(sizeof(AttnBarriers));

// One work item, decoded identically by every role.
struct AttnItem {
  int s, b, h, qt0, q_len, k_len, nkt;  // s: side
  bool skip;  // no valid query row in the block, or no key: rows are written as zeros, no MMA work
};
__device__ __forceinline__ AttnItem attn_item(const AttnParams& p, int w) {
  AttnItem it;
  it.s = w >= p.side[0].items;
  const AttnSide& sd = p.side[it.s];
  w -= it.s ? p.side[0].items : 0;
  const int nqb = sd.tq_pad >> 1;
  const int bh = w / nqb;
  it.qt0 = (w - bh * nqb) * 2;
  it.b = bh / p.H;
  it.h = bh - it.b * p.H;
  it.q_len = sd.q_len ? min(sd.q_len[it.b], sd.nq) : sd.nq;
  it.k_len = sd.k_len ? min(sd.k_len[it.b], sd.nk) : sd.nk;
  it.nkt = (it.k_len + kTile - 1) / kTile;
  it.skip = it.qt0 * kTile >= it.q_len || it.nkt == 0;
  return it;
}

// kSplit: FOUR softmax warpgroups instead of two -- every query row is shared by two threads (64 of the 128 key
// columns of a block each), so each scheduler holds four softmax warps whose MUFU / FMA work interleaves (two warps per
// scheduler could not keep the MUFU unit busy: every warp sits in a dependent FFMA -> MUFU -> FADD chain).
template <bool kSplit>
__global__ void __launch_bounds__(kSplit ? 640 : kAttnThreads, 1) attention_kernel(const AttnParams p) {
  constexpr int kSoft = kSplit ? 16 : 8;  // softmax warps; then producer, PV issuer, similarity issuer
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // 1 KB aligned, still provably shared
  uint8_t* sQ = base;                               // 2 buffers x (2 x 16 KB)
  uint8_t* sKV = sQ + 4 * kTileBytes;               // kStages x (K 16 KB | V^T 16 KB)
  uint8_t* sOut = sKV + kStages * 2 * kTileBytes;   // 8 x 4 KB epilogue staging
  AttnBarriers* bar = reinterpret_cast<AttnBarriers*>(sOut + kOutStage);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_stride = p.H * kHeadDim;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar->q_full[i], 1); mbar_init(&bar->q_empty[i], 1);
      mbar_init(&bar->p_full[2 * i], kSplit ? 8 : (LG_WARP_ARRIVE ? 4 : 1));
      mbar_init(&bar->p_full[2 * i + 1], kSplit ? 8 : (LG_WARP_ARRIVE ? 4 : 1));
      mbar_init(&bar->pv_done[i], 1); mbar_init(&bar->t0_free[i], 1); mbar_init(&bar->o_free[i], kSplit ? 256 : 128);
      bar->oflag[2 * i] = 0u; bar->oflag[2 * i + 1] = 0u;
    }
    for (int s = 0; s < kStages; ++s) { mbar_init(&bar->kv_full[s], 1); mbar_init(&bar->kv_empty[s], 1); }
    for (int i = 0; i < 3; ++i) mbar_init(&bar->s_full[i], 1);
    fence_mbar_init();
  }
  if (warp == kSoft + 1) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  pdl_launch_dependents();
  pdl_wait();  // operand tiles / lengths written by the previous kernels are complete and visible from here on

  if (warp >= kSoft) {
    if (kSplit) reg_dec<64>();  // pool = 640 x 96: 512 * 104 + 128 * 64 registers
    else reg_dec<88>();         // pool = 384 x 168: 128 * (208 + 208 + 88) registers
    if (warp == kSoft) {
      // ---------------------------------------------------------------- producer
      if (lane == 0) {
        uint32_t kvi = 0;  // running K/V block counter (ring position), across items
        int cnt = 0;       // non-skipped items so far (Q buffer = cnt & 1)
        for (int w = blockIdx.x; w < p.items; w += gridDim.x) {
          const AttnItem it = attn_item(p, w);
          if (it.skip) continue;
          DBG(0, w);
          const AttnSide& sd = p.side[it.s];
          const size_t bh = static_cast<size_t>(it.b) * p.H + it.h;
          const uint8_t* gq = sd.q_tiles + (bh * sd.tq_pad + it.qt0) * kTileBytes;
          const uint8_t* gk = sd.k_tiles + bh * sd.tk_pad * kTileBytes;
          const uint8_t* gv = sd.v_tiles + bh * sd.tk_pad * kTileBytes;
          const int qs = cnt & 1;
          if (cnt >= 2) mbar_wait(&bar->q_empty[qs], ((cnt >> 1) - 1) & 1);
          mbar_arrive_expect_tx(&bar->q_full[qs], 2 * kTileBytes);
          bulk_g2s(sQ + qs * 2 * kTileBytes, gq, 2 * kTileBytes, &bar->q_full[qs]);  // tiles qt0, qt0+1 are adjacent
          for (int j = 0; j < it.nkt; ++j, ++kvi) {
            const int s = kvi % kStages;
            DBG(1, static_cast<int>(kvi));
            if (kvi >= kStages) mbar_wait(&bar->kv_empty[s], ((kvi / kStages) - 1) & 1);
            mbar_arrive_expect_tx(&bar->kv_full[s], 2 * kTileBytes);
            bulk_g2s(sKV + s * 2 * kTileBytes, gk + static_cast<size_t>(j) * kTileBytes, kTileBytes, &bar->kv_full[s]);
            bulk_g2s(sKV + s * 2 * kTileBytes + kTileBytes, gv + static_cast<size_t>(j) * kTileBytes, kTileBytes,
                     &bar->kv_full[s]);
          }
          ++cnt;
        }
      }
    } else if (warp == kSoft + 2) {
      // ---------------------------------------------------------------- similarity issuer: S(G) = Q_t K_j^T
      // Two issuing warps (this one and the PV warp) instead of one: a single issuer measured 1250 cycles per half
      // step against 512 cycles of tensor work -- every ready mbarrier wait costs ~90 cycles, every tcgen05.commit
      // ~150, and the issue of an MMA blocks while the tensor pipe's short queue is full, so one warp's waits and
      // bookkeeping run with an idle tensor pipe.  With two warps the queue holds the other warp's MMAs meanwhile.
      // Global step G = (key block, query tile) pairs in order, across items; S buffer = G % 3.  S(G) may be written
      // once PV(G-3) has READ P(G-3) from the same buffer: for an even G-3 (tile 0) the PV warp commits
      // t0_free[block & 1]; for an odd G-3 (tile 1) its commit of kv_empty[stage] says the same.
      const bool leader = elect_one();  // all lanes run the loop (uniform operands); one lane issues
      constexpr uint32_t idesc_qk = umma_idesc(kFmtBF16, 128, 128);
      const uint64_t dQ0 = umma_desc_kmajor_sw128(smem_u32(sQ));
      const uint64_t dK0 = umma_desc_kmajor_sw128(smem_u32(sKV));
      uint32_t kv0 = 0, sb = 0, G = 0;  // kv0: key blocks of the earlier items; sb = G % 3
      int cnt = 0;                      // non-skipped items so far (Q buffer = cnt & 1)
      for (int w = blockIdx.x; w < p.items; w += gridDim.x) {
        const AttnItem it = attn_item(p, w);
        if (it.skip) continue;
        mbar_wait(&bar->q_full[cnt & 1], (cnt >> 1) & 1);
        const int nsteps = 2 * it.nkt;
        for (int n = 0; n < nsteps; ++n, ++G) {
          const int t = n & 1;
          const uint32_t kvi = kv0 + (n >> 1), s = kvi % kStages;
          if (t == 0) mbar_wait(&bar->kv_full[s], (kvi / kStages) & 1);
          if (G >= 3) {
            if (t == 0) {  // G-3 = tile 1 of block kvi-2
              const uint32_t kp = kvi - 2;
              mbar_wait(&bar->kv_empty[kp % kStages], (kp / kStages) & 1);
            } else {       // G-3 = tile 0 of block kvi-1
              const uint32_t kp = kvi - 1;
              mbar_wait(&bar->t0_free[kp & 1], (kp >> 1) & 1);
            }
          }
          tc_fence_after();
          const uint64_t da = dQ0 + static_cast<uint64_t>(((cnt & 1) * 2 + t) * (kTileBytes >> 4));
          const uint64_t db = dK0 + static_cast<uint64_t>(s * (2 * kTileBytes >> 4));
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)  // K = 64 = 4 x 16, 32 B per step inside the 128 B swizzle row
              umma_bf16_ss(tmem + kColS + sb * 128, da + 2 * kk, db + 2 * kk, idesc_qk, kk > 0);
            umma_commit(&bar->s_full[sb]);
            if (n + 1 == nsteps) umma_commit(&bar->q_empty[cnt & 1]);  // last similarity of the item: Q buffer free
            TRACE(27, G);
          }
          sb = sb == 2 ? 0 : sb + 1;
        }
        kv0 += it.nkt;
        ++cnt;
      }
    } else if (warp == kSoft + 1) {
      // ---------------------------------------------------------------- PV issuer: O_t += P_t(j) V_j
      const bool leader = elect_one();
      constexpr uint32_t idesc_pv = umma_idesc(kFmtBF16, 128, 64);
      // Descriptors differ only in their 14-bit (address >> 4) field, which cannot carry out for addresses below
      // 256 KB: one descriptor per operand array, plain 64-bit adds per step.
      const uint64_t dV0 = umma_desc_kmajor_sw128(smem_u32(sKV) + kTileBytes);
      uint32_t p_kv = 0, p_sb = 0;
      int p_cnt = 0;
#ifdef LG_ATTN_TRACE
      uint32_t G_p = 0;
#endif
      for (int w = blockIdx.x; w < p.items; w += gridDim.x) {
        const AttnItem it = attn_item(p, w);
        if (it.skip) continue;
        const int nsteps = 2 * it.nkt;
        for (int n = 0; n < nsteps; ++n) {
          const int t = n & 1, j = n >> 1;
          const uint32_t kvi = p_kv + j, s = kvi % kStages;
          // P sits in the first 64 columns of S buffer G % 3 as packed bf16 (A operand from TMEM)
          const uint32_t tP = tmem + kColS + p_sb * 128;
          const uint64_t dv = dV0 + static_cast<uint64_t>(s * (2 * kTileBytes >> 4));
          if (leader) TRACE(24, G_p);
          mbar_wait(&bar->p_full[2 * t + (kvi & 1)], (kvi >> 1) & 1);  // P_t(j) stored in its S buffer
          if (leader) TRACE(6, G_p);
          if (j == 0 && p_cnt > 0) mbar_wait(&bar->o_free[t], (p_cnt - 1) & 1);  // previous item's O_t was read out
          tc_fence_after();
          if (leader) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk)  // K = 128 keys = 8 x 16: 8 TMEM columns / 32 B of a V^T row per step
              umma_bf16_ts(tmem + kColO + t * 64, tP + (kSplit ? (kk & 3) * 8 + (kk >> 2) * 64 : kk * 8),
                           dv + static_cast<uint64_t>((kk >> 2) * (kTileBytes / 2 >> 4) + (kk & 3) * 2), idesc_pv,
                           (j > 0 || kk > 0));
            TRACE(26, G_p);
            // ONE commit per step says "PV(G) has retired": the S buffer may be overwritten (similarity warp), and for
            // tile 1 also "K_j, V_j are dead" (producer; the rare rescale path of the softmax warps waits on it too)
            if (t == 1) umma_commit(&bar->kv_empty[s]);
            else umma_commit(&bar->t0_free[kvi & 1]);
            if (n + 2 >= nsteps) umma_commit(&bar->pv_done[t]);  // last PV_t of the item: the item epilogue waits for it
          }
          p_sb = p_sb == 2 ? 0 : p_sb + 1;
          if (leader) TRACE(7, G_p);
#ifdef LG_ATTN_TRACE
          ++G_p;
#endif
        }
        p_kv += it.nkt;
        ++p_cnt;
      }
    }
  } else if constexpr (kSplit) {
    reg_inc<104>();
    // ------------------------------------------------------------------ softmax, two threads per row
    // warpgroup = (tile t, column half h): thread (row r, half h) owns columns [64 h, 64 h + 64) of every S block of its
    // row and columns [32 h, 32 h + 32) of O.  Both halves of a row use ONE reference max (exchanged through the
    // pair's staging slice at the first block and in the rare rescale), keep their own partial row sums (added in the
    // epilogue), and meet at a 256-thread barrier per block where they agree on the (rare) rescale.  P of half h goes
    // to the first 32 columns of the half's OWN 64 S columns (keys 0-63 at columns 0-31, keys 64-127 at 64-95: the PV
    // warp addresses them accordingly), so a thread only ever overwrites S values it has read itself and the store
    // is issued before the barrier.
    const int wgid = warp >> 2, t = wgid >> 1, h = wgid & 1, wq = warp & 3;
    const int r = wq * 32 + lane;                  // row == TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t tS0 = tmem + lane_addr + kColS;
    const uint32_t tO = tmem + lane_addr + kColO + t * 64 + h * 32;
    const float c = p.scale_log2;
    const float2 c2 = make_float2(c, c);
    const uint32_t tile_bar = 9 + t;               // named barrier of the tile's 256 threads
    uint8_t* slice = sOut + (t * 4 + wq) * 4096;   // the pair's staging slice: 32 rows x 128 B of the context tile image
    float* xch = reinterpret_cast<float*>(slice);  // exchange words [half][lane] (only between bulk stores)
    const bool issuer = h == 0 && lane == 0;       // issues the pair's bulk stores
    uint32_t pk[32];                               // P of the block: this thread's 64 columns as packed bf16 pairs
    uint32_t v[32];
    uint32_t kv_base = 0, item_cnt = 0;
    auto tile_sync = [&]() { asm volatile("bar.sync %0, 256;" ::"r"(tile_bar) : "memory"); };

    for (int w = blockIdx.x; w < p.items; w += gridDim.x) {
      const AttnItem it = attn_item(p, w);
      const int row = (it.qt0 + t) * kTile + r;
      const AttnSide& sd = p.side[it.s];
      const int side_nq = sd.nq;
      uint8_t* dt = p.out_tiles
                        ? p.out_tiles + ((static_cast<size_t>(sd.out_rt_base + it.b * sd.tq_pad + it.qt0 + t) * p.H + it.h) * 32768)
                        : nullptr;
      float4* drow = sd.out ? reinterpret_cast<float4*>(sd.out + (static_cast<size_t>(it.b) * sd.nq + row) * row_stride +
                                                        it.h * kHeadDim + h * 32)
                            : nullptr;
      if (it.skip) {  // nothing to attend from / to: rows are zero (reference: nan_to_num of an all-masked softmax)
        if (dt) {
#pragma unroll
          for (int ch = 4 * h; ch < 4 * h + 4; ++ch) {
            *reinterpret_cast<uint4*>(dt + sw128_offset(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(dt + 16384 + sw128_offset(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
          }
        } else if (row < side_nq) {
#pragma unroll
          for (int i = 0; i < 8; ++i) drow[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        continue;
      }

      float m_used = -INFINITY, l = 0.f;
      for (int j = 0; j < it.nkt; ++j) {
        const int rem = it.k_len - j * kTile - 64 * h;  // valid keys among this thread's 64 columns (may be <= 0)
        const bool partial = rem < 64;
        const uint32_t G = 2 * (kv_base + j) + t;       // global step: S buffer G % 3
        const uint32_t tS = tS0 + (G % 3) * 128 + 64 * h;
        uint32_t* oflag = &bar->oflag[2 * t + ((kv_base + j) & 1)];
        if ((threadIdx.x & 127) == 0 && h == 0) TRACE(t * 3, kv_base + j);
#if LG_SPLIT_ALT
        // The two tiles take turns on the MUFU unit: tile 1 starts block n when tile 0 is half way through its block n,
        // tile 0 starts block n+1 when tile 1 is half way through block n.  Left alone, tile 1 trails tile 0 by the
        // latency of its similarity (~300 cycles): both sweep together at half the MUFU rate each and then sit in
        // their P hand-off (barrier, TMEM store, wait for the next similarity: ~770 cycles) with the unit idle.
        if (t == 1) asm volatile("bar.sync 11, 512;" ::: "memory");
        else if (kv_base + j > 0) asm volatile("bar.sync 12, 512;" ::: "memory");
        bool signalled = false;
#endif
        mbar_wait(&bar->s_full[G % 3], (G / 3) & 1);
        tc_fence_after();
        if ((threadIdx.x & 127) == 0 && h == 0) TRACE(t * 3 + 1, kv_base + j);
        const bool first = j == 0;
        float tile_sum = 0.f;

        // one pass over this thread's 64 columns: P = exp2(S c - m c), half of the exponentials as polynomials
        auto sweep = [&]() {
          const float ms = m_used * c;
          const float2 nms2 = make_float2(-ms, -ms);
          const float cs252 = c * (1.f / 252.f), bs252 = fmaf(-ms, 1.f / 252.f, 0.5f);
          float2 s0 = make_float2(0.f, 0.f), s1 = s0;
#pragma unroll
          for (int cb = 0; cb < 2; ++cb) {
            tmem_ld32(tS + cb * 32, v);
            tmem_wait_ld();
            if (partial) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (cb * 32 + i >= rem) v[i] = 0xff800000u;  // -inf: exp2 -> 0
            }
#if LG_SPLIT_POLY == 2
#pragma unroll
            for (int q = 0; q < 8; ++q) {  // 4 columns: one MUFU pair, one polynomial pair
              const float2 ta = __ffma2_rn(make_float2(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1])), c2, nms2);
              const float2 ea = make_float2(ex2_approx(ta.x), ex2_approx(ta.y));
              const float2 eb = exp2_poly2(__uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]), cs252, bs252);
              s0 = __fadd2_rn(s0, ea);
              s1 = __fadd2_rn(s1, eb);
              pk[cb * 16 + 2 * q] = pack_bf16x2(ea.x, ea.y);
              pk[cb * 16 + 2 * q + 1] = pack_bf16x2(eb.x, eb.y);
            }
#else
#pragma unroll
            for (int q = 0; q < 4; ++q) {  // 8 columns: LG_SPLIT_POLY of the 4 pairs as polynomials, the others on MUFU
#pragma unroll
              for (int pi = 0; pi < 4; ++pi) {
                const float xa = __uint_as_float(v[8 * q + 2 * pi]), xb = __uint_as_float(v[8 * q + 2 * pi + 1]);
                float2 e;
                if (pi < LG_SPLIT_POLY) {
                  e = exp2_poly2(xa, xb, cs252, bs252);
                } else {
                  const float2 tt = __ffma2_rn(make_float2(xa, xb), c2, nms2);
                  e = make_float2(ex2_approx(tt.x), ex2_approx(tt.y));
                }
                if (pi & 1) s1 = __fadd2_rn(s1, e); else s0 = __fadd2_rn(s0, e);
                pk[cb * 16 + 4 * q + pi] = pack_bf16x2(e.x, e.y);
              }
            }
#endif
#if LG_SPLIT_ALT
            if (cb == 0 && !signalled) {  // half way: once per block
              signalled = true;
              if (t == 0) asm volatile("bar.arrive 11, 512;" ::: "memory");
              else asm volatile("bar.arrive 12, 512;" ::: "memory");
            }
#endif
          }
          tile_sum = (s0.x + s0.y) + (s1.x + s1.y);
        };

        bool exact = first, stored = false;
        if (!first) {
          sweep();
          // every P <= its row sum: below the limit nothing can have overflowed the bf16 / fp32 ranges
          const bool warp_over = __any_sync(0xffffffffu, !(tile_sum < kSumLimit));
          if (!warp_over) {  // P over the first 32 of this thread's own S columns, before the barrier
            tmem_st32(tS, pk);
            stored = true;
          } else if (lane == 0) {
            *reinterpret_cast<volatile uint32_t*>(oflag) = 1u;
          }
          if (threadIdx.x == 0) TRACE(16, kv_base + j);
          tile_sync();  // the flag is visible: both halves of every row agree on the (rare) rescale below
          if (threadIdx.x == 0) TRACE(17, kv_base + j);
          exact = *reinterpret_cast<volatile uint32_t*>(oflag) != 0u;
        }
        if (exact) {
          // New reference for the rows that need one: first block of an item, or a row sum that reached the limit.  A
          // warp that has already stored P (none of ITS rows over the limit) no longer holds S: its rows contribute
          // nothing to the exchange -- every S value of such a half is below old reference + 60 / c, i.e. below the
          // exact max of the half that did reach the limit -- and rescale the P they have.
          const bool intact = !stored;
          float mx = -INFINITY;
          if (intact) {
#pragma unroll
            for (int cb = 0; cb < 2; ++cb) {
              tmem_ld32(tS + cb * 32, v);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                const bool k0 = !partial || cb * 32 + i < rem, k1 = !partial || cb * 32 + i + 1 < rem;
                mx = fmax3(mx, k0 ? __uint_as_float(v[i]) : -INFINITY, k1 ? __uint_as_float(v[i + 1]) : -INFINITY);
              }
            }
          }
          if (issuer) bulk_wait_read_all();  // the slice may still be read by the previous item's context stores
          tile_sync();
          const bool over = !(tile_sum < kSumLimit);  // (tile_sum is 0 in a first block)
          xch[(h * 32 + lane) * 2] = mx;
          xch[(h * 32 + lane) * 2 + 1] = over ? 1.f : 0.f;
          tile_sync();
          mx = fmaxf(mx, xch[((1 - h) * 32 + lane) * 2]);
          // a ROW takes the new reference if either of its halves reached the limit (then the block max exceeds the old
          // reference and alpha < 1); the other rows of the tile keep theirs -- a lower reference would scale O up
          const bool need = first || over || xch[((1 - h) * 32 + lane) * 2 + 1] != 0.f;
          float alpha = 1.f;
          if (need) { alpha = first ? 0.f : ex2_approx((m_used - mx) * c); m_used = mx; }
          l *= alpha;
          if (!first) {
            // O_t must be quiescent: PV_t(j-1) retired (kv_empty of block j-1 completes behind PV_1(j-1))
            mbar_wait(&bar->kv_empty[(kv_base + j - 1) % kStages], ((kv_base + j - 1) / kStages) & 1);
            tc_fence_after();
            tmem_ld32(tO, v);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st32(tO, v);
            tmem_wait_st();
          }
          if (intact) {
            sweep();  // from S, against the row's (possibly new) reference
          } else {    // S is gone: P = P_old * 2^((m_old - m_new) c), and so is this block's row sum
            tile_sum *= alpha;
#pragma unroll
            for (int i = 0; i < 32; ++i)
              pk[i] = pack_bf16x2(__uint_as_float(pk[i] << 16) * alpha, __uint_as_float(pk[i] & 0xFFFF0000u) * alpha);
          }
          tmem_st32(tS, pk);
          if (!first && issuer && wq == 0) *reinterpret_cast<volatile uint32_t*>(oflag) = 0u;  // (read before the last barrier)
        }
        l += tile_sum;
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[2 * t + ((kv_base + j) & 1)]);  // 8 warp arrivals per block
        if ((threadIdx.x & 127) == 0 && h == 0) TRACE(t * 3 + 2, kv_base + j);
      }

      // epilogue: O / l
      mbar_wait(&bar->pv_done[t], item_cnt & 1);
      ++item_cnt;
      tc_fence_after();
      tmem_ld32(tO, v);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&bar->o_free[t]);
      kv_base += it.nkt;
      // the row sum: this half's + the partner's
      if (issuer) bulk_wait_read_all();
      tile_sync();
      xch[h * 32 + lane] = l;
      tile_sync();
      l += xch[(1 - h) * 32 + lane];
      tile_sync();  // the staging below overwrites the exchange words
      const float inv = (row < it.q_len) ? 1.f / l : 0.f;  // padded query rows -> 0
      if (dt) {
        // this thread's 32 context columns = 16-byte chunks 4 h .. 4 h + 3 of its row in the hi and the lo tile image
        const uint32_t stg_row = smem_u32(slice) + lane * 128;
        uint32_t lo_w[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t hh[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float x0 = __uint_as_float(v[8 * q + 2 * i]) * inv, x1 = __uint_as_float(v[8 * q + 2 * i + 1]) * inv;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
            hh[i] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
            lo_w[q * 4 + i] = pack_bf16x2(x0 - __bfloat162float(h0), x1 - __bfloat162float(h1));
          }
          st_shared_v4(stg_row + (((4 * h + q) ^ (lane & 7)) << 4), make_uint4(hh[0], hh[1], hh[2], hh[3]));
        }
        fence_proxy_async_smem();
        tile_sync();
        uint8_t* gdst = dt + wq * 4096;
        if (issuer) { bulk_s2g(gdst, slice, 4096); bulk_commit(); bulk_wait_read_all(); }
        tile_sync();
#pragma unroll
        for (int q = 0; q < 4; ++q)
          st_shared_v4(stg_row + (((4 * h + q) ^ (lane & 7)) << 4),
                       make_uint4(lo_w[4 * q], lo_w[4 * q + 1], lo_w[4 * q + 2], lo_w[4 * q + 3]));
        fence_proxy_async_smem();
        tile_sync();
        if (issuer) { bulk_s2g(gdst + 16384, slice, 4096); bulk_commit(); }
      } else if (row < side_nq) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          drow[i] = make_float4(__uint_as_float(v[4 * i]) * inv, __uint_as_float(v[4 * i + 1]) * inv,
                                __uint_as_float(v[4 * i + 2]) * inv, __uint_as_float(v[4 * i + 3]) * inv);
      }
    }
#if LG_SPLIT_ALT
    if (t == 0 && kv_base > 0) asm volatile("bar.sync 12, 512;" ::: "memory");  // absorbs tile 1's last half-way signal
#endif
    if (issuer) bulk_wait_all();  // outstanding context stores
  } else {
    reg_inc<208>();
    // ------------------------------------------------------------------ softmax warpgroups
    const int t = warp >> 2;                       // query tile of this warpgroup
    const int r = (warp & 3) * 32 + lane;          // row == TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS0 = tmem + lane_addr + kColS;
    const uint32_t tO = tmem + lane_addr + kColO + t * 64;
    const float c = p.scale_log2;
    const float2 c2 = make_float2(c, c);
    uint32_t va[32], vb[32];
    uint32_t pk0[32], pk1[32];  // P of the block as packed bf16 pairs: 64 TMEM columns
    uint32_t kv_base = 0;       // running key-block counter: barrier phases continue across items
    uint32_t item_cnt = 0;      // non-skipped items so far (phase of pv_done)
#if LG_STAGGER > 0
    if (t == 1) {  // start tile 1 half a block late: the sweeps of the two warpgroups then alternate on the MUFU unit
      long long c0 = clock64();
      while (clock64() - c0 < LG_STAGGER) {}
    }
#endif

    for (int w = blockIdx.x; w < p.items; w += gridDim.x) {
      const AttnItem it = attn_item(p, w);
      const int row = (it.qt0 + t) * kTile + r;
      const AttnSide& sd = p.side[it.s];
      const int side_nq = sd.nq;
      uint8_t* dt = p.out_tiles
                        ? p.out_tiles + ((static_cast<size_t>(sd.out_rt_base + it.b * sd.tq_pad + it.qt0 + t) * p.H + it.h) * 32768)
                        : nullptr;
      float4* drow = sd.out ? reinterpret_cast<float4*>(sd.out + (static_cast<size_t>(it.b) * sd.nq + row) * row_stride +
                                                        it.h * kHeadDim)
                            : nullptr;
      if ((threadIdx.x & 127) == 0) { DBG(6 + t, w); DBG(10 + t, it.skip ? 9 : 0); }
      if (it.skip) {
        // nothing to attend from / to: rows are zero (reference: nan_to_num of an all-masked softmax)
#ifdef LG_DBG_NOSKIPZERO
        continue;
#endif
        if (dt) {
#pragma unroll
          for (int ch = 0; ch < 8; ++ch) {
            *reinterpret_cast<uint4*>(dt + sw128_offset(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(dt + 16384 + sw128_offset(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
          }
        } else if (row < side_nq) {
#pragma unroll
          for (int i = 0; i < 16; ++i) drow[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        continue;
      }

      float m_used = -INFINITY, l = 0.f;
      bool pf = false;  // the first 32 columns of this block's S are already on their way into va (see the block's end)
      for (int j = 0; j < it.nkt; ++j) {
        const int rem = it.k_len - j * kTile;        // valid keys in this block (>= 1)
        const bool partial = rem < kTile;
        const uint32_t G = 2 * (kv_base + j) + t;    // global step: S buffer G % 3
        const uint32_t tS = tS0 + (G % 3) * 128;
        if ((threadIdx.x & 127) == 0) TRACE(t * 3, kv_base + j);
        if ((threadIdx.x & 127) == 0) { DBG(8 + t, j); DBG(10 + t, 1); }
#if LG_ALTERNATE
        // Tile 1 starts block n when tile 0 is half way through its block n; tile 0 starts block n+1 when tile 1 is
        // half way through block n.  Left alone the two warpgroups fall into step (tile 1 trails by the latency of its
        // similarity), sweep at the same time at half the MUFU rate each and then sit in their P hand-off together
        // with the MUFU unit idle; alternating, one sweeps alone while the other stores P, waits and loads.
        if (t == 1) asm volatile("bar.sync 3, 256;" ::: "memory");
        else if (kv_base + j > 0) asm volatile("bar.sync 4, 256;" ::: "memory");
#endif
        if (!pf) {
          LG_WAIT(&bar->s_full[G % 3], (G / 3) & 1);
          tc_fence_after();
        }
        if ((threadIdx.x & 127) == 0) DBG(10 + t, 2);
        if ((threadIdx.x & 127) == 0) TRACE(t * 3 + 1, kv_base + j);

        float tile_sum = 0.f;
        const bool first = j == 0;
#pragma unroll 1
        for (int attempt = 0; attempt < 2; ++attempt) {
          if (first || attempt == 1) {
            // exact block max (first block of a row, or a row sum that reached the limit): new reference, rescale
            // (taking the max of the first 32 columns only for a first block measured slower: 0.3435 vs 0.3410 ms)
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
            const bool need = first || !(tile_sum < kSumLimit);
            float alpha = 1.f;
            if (need) { alpha = first ? 0.f : ex2_approx((m_used - mx) * c); m_used = mx; }
            l *= alpha;
            if (!first) {
              // O_t must be quiescent: PV_t(j-1) retired.  kv_empty of block j-1 completes when PV_1(j-1) has retired
              // (and with it everything issued before); its next completion (block j+3) needs this block's P first,
              // so the parity wait cannot alias.
              LG_WAIT(&bar->kv_empty[(kv_base + j - 1) % kStages], ((kv_base + j - 1) / kStages) & 1);
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
          const float ms = m_used * c;
          const float2 nms2 = make_float2(-ms, -ms);
          const float cs252 = c * (1.f / 252.f), bs252 = fmaf(-ms, 1.f / 252.f, 0.5f);
          float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
#if LG_SWEEP == 0
          auto sweep = [&](uint32_t(&v)[32], const int cb) {
            if (partial) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (cb * 32 + i >= rem) v[i] = 0xff800000u;  // -inf: exp2 -> 0
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float e[8];
#pragma unroll
              for (int pi = 0; pi < 4; ++pi) {  // pairs of the 8-column group
                const float xa = __uint_as_float(v[8 * q + 2 * pi]), xb = __uint_as_float(v[8 * q + 2 * pi + 1]);
                // 6 of every 16 pairs take the polynomial path: balances the MUFU pipe against the issue slots
                const bool poly = LG_POLYEXP == 1 ? (pi == 3) : LG_POLYEXP == 2 ? (pi == 3 || ((q & 1) && pi == 1))
                                  : LG_POLYEXP == 3 ? (pi & 1) : LG_POLYEXP == 4 ? (pi != 0) : false;
                float2 ee;
                if (poly) {
                  ee = exp2_poly2(xa, xb, cs252, bs252);
                } else {
                  const float2 tt = __ffma2_rn(make_float2(xa, xb), c2, nms2);
                  ee = make_float2(ex2_approx(tt.x), ex2_approx(tt.y));
                }
                e[2 * pi] = ee.x; e[2 * pi + 1] = ee.y;
                if (pi & 1) s1 = __fadd2_rn(s1, ee); else s0 = __fadd2_rn(s0, ee);
              }
              uint32_t(&pk)[32] = cb < 2 ? pk0 : pk1;  // cb, q are compile-time after unrolling
              pk[(cb & 1) * 16 + q * 4] = pack_bf16x2(e[0], e[1]);
              pk[(cb & 1) * 16 + q * 4 + 1] = pack_bf16x2(e[2], e[3]);
              pk[(cb & 1) * 16 + q * 4 + 2] = pack_bf16x2(e[4], e[5]);
              pk[(cb & 1) * 16 + q * 4 + 3] = pack_bf16x2(e[6], e[7]);
            }
          };
#else
          // Same arithmetic, written stage by stage over the whole 32-column chunk (all scale FMAs, the polynomial
          // pairs advanced in lock step, all MUFUs, then sums / packs) with four sum accumulators: every dependent
          // pair of instructions is a whole stage apart.
          float2 s2 = make_float2(0.f, 0.f), s3 = make_float2(0.f, 0.f);
          auto sweep = [&](uint32_t(&v)[32], const int cb) {
            if (partial) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (cb * 32 + i >= rem) v[i] = 0xff800000u;  // -inf: exp2 -> 0
            }
            constexpr float kMagic126 = 12582912.f - 126.f;
            const float2 k252 = make_float2(252.f, 252.f), km = make_float2(kMagic126, kMagic126);
            float2 ee[16], yy[4], rr[4], fr[4], pp[4];
            // stage A: arguments
#pragma unroll
            for (int pr = 0; pr < 16; ++pr) {
              const float xa = __uint_as_float(v[2 * pr]), xb = __uint_as_float(v[2 * pr + 1]);
              if ((pr & 3) == 3) {
                float ya, yb;
                asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(ya) : "f"(xa), "f"(cs252), "f"(bs252));
                asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(yb) : "f"(xb), "f"(cs252), "f"(bs252));
                yy[pr >> 2] = make_float2(ya, yb);
              } else {
                ee[pr] = __ffma2_rn(make_float2(xa, xb), c2, nms2);
              }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) rr[k] = __ffma2_rn(yy[k], k252, km);
#pragma unroll
            for (int k = 0; k < 4; ++k) fr[k] = __fadd2_rn(km, make_float2(-rr[k].x, -rr[k].y));
            // stage B: exponentials (MUFU), polynomial steps in between
#pragma unroll
            for (int pr = 0; pr < 4; ++pr)
              if ((pr & 3) != 3) ee[pr] = make_float2(ex2_approx(ee[pr].x), ex2_approx(ee[pr].y));
#pragma unroll
            for (int k = 0; k < 4; ++k) fr[k] = __ffma2_rn(yy[k], k252, fr[k]);
#pragma unroll
            for (int pr = 4; pr < 8; ++pr)
              if ((pr & 3) != 3) ee[pr] = make_float2(ex2_approx(ee[pr].x), ex2_approx(ee[pr].y));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              pp[k] = __ffma2_rn(make_float2(0.0551716685f, 0.0551716685f), fr[k], make_float2(0.2426111251f, 0.2426111251f));
#pragma unroll
            for (int pr = 8; pr < 12; ++pr)
              if ((pr & 3) != 3) ee[pr] = make_float2(ex2_approx(ee[pr].x), ex2_approx(ee[pr].y));
#pragma unroll
            for (int k = 0; k < 4; ++k) pp[k] = __ffma2_rn(pp[k], fr[k], make_float2(0.6932609677f, 0.6932609677f));
#pragma unroll
            for (int pr = 12; pr < 16; ++pr)
              if ((pr & 3) != 3) ee[pr] = make_float2(ex2_approx(ee[pr].x), ex2_approx(ee[pr].y));
#pragma unroll
            for (int k = 0; k < 4; ++k) pp[k] = __ffma2_rn(pp[k], fr[k], make_float2(0.9999280572f, 0.9999280572f));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ee[4 * k + 3] = make_float2(__int_as_float((__float_as_int(rr[k].x) << 23) + __float_as_int(pp[k].x)),
                                          __int_as_float((__float_as_int(rr[k].y) << 23) + __float_as_int(pp[k].y)));
            // stage C: row sums (four independent chains) and bf16 packing
#pragma unroll
            for (int pr = 0; pr < 16; pr += 4) {
              s0 = __fadd2_rn(s0, ee[pr]);
              s1 = __fadd2_rn(s1, ee[pr + 1]);
              s2 = __fadd2_rn(s2, ee[pr + 2]);
              s3 = __fadd2_rn(s3, ee[pr + 3]);
            }
            uint32_t(&pk)[32] = cb < 2 ? pk0 : pk1;  // cb is a compile-time constant after inlining
#pragma unroll
            for (int pr = 0; pr < 16; ++pr) pk[(cb & 1) * 16 + pr] = pack_bf16x2(ee[pr].x, ee[pr].y);
          };
#endif
#if LG_SWEEP == 2
          // Software-pipelined over the 16 groups of 8 columns (3 MUFU pairs + 1 polynomial pair each).  The MUFU
          // unit takes a warp instruction every 8 cycles; ptxas keeps source order, so the source interleaves every
          // MUFU of group g with the scale FMAs / polynomial chain of group g+1 and the sums / packs of group g-1.
          {
            constexpr float kMagic126 = 12582912.f - 126.f;
            const float2 k252 = make_float2(252.f, 252.f), km = make_float2(kMagic126, kMagic126);
            float2 T[16][3], E[16][4], Y[16], RR[16], FR[16], PP[16];
#if LG_POLY_PAIRS == 2
            float2 Y2[16], RR2[16], FR2[16], PP2[16];
#endif
            float2 s2 = make_float2(0.f, 0.f), s3 = make_float2(0.f, 0.f);
            auto mask_chunk = [&](uint32_t(&v)[32], const int cb) {
              if (partial) {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (cb * 32 + i >= rem) v[i] = 0xff800000u;  // -inf: exp2 -> 0
              }
            };
#define LG_V(g, i) __uint_as_float((((g) >> 2) & 1) ? vb[((g) & 3) * 8 + (i)] : va[((g) & 3) * 8 + (i)])
#define LG_ARGS1(g)                                                                   \
  T[g][0] = __ffma2_rn(make_float2(LG_V(g, 0), LG_V(g, 1)), c2, nms2);                \
  T[g][1] = __ffma2_rn(make_float2(LG_V(g, 2), LG_V(g, 3)), c2, nms2);
#if LG_POLY_PAIRS == 2
#define LG_ARGS2_T2(g)                                                                \
  {                                                                                   \
    float ya, yb;                                                                     \
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(ya) : "f"(LG_V(g, 4)), "f"(cs252), "f"(bs252)); \
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(yb) : "f"(LG_V(g, 5)), "f"(cs252), "f"(bs252)); \
    Y2[g] = make_float2(ya, yb);                                                      \
  }
#define LG_QA(g) RR2[g] = __ffma2_rn(Y2[g], k252, km);
#define LG_QB(g) FR2[g] = __fadd2_rn(km, make_float2(-RR2[g].x, -RR2[g].y));
#define LG_QC(g) FR2[g] = __ffma2_rn(Y2[g], k252, FR2[g]);
#define LG_QD(g) PP2[g] = __ffma2_rn(make_float2(0.0551716685f, 0.0551716685f), FR2[g], make_float2(0.2426111251f, 0.2426111251f));
#define LG_QE(g) PP2[g] = __ffma2_rn(PP2[g], FR2[g], make_float2(0.6932609677f, 0.6932609677f));
#define LG_QF(g) PP2[g] = __ffma2_rn(PP2[g], FR2[g], make_float2(0.9999280572f, 0.9999280572f));
#define LG_QG(g)                                                                                        \
  E[g][2] = make_float2(__int_as_float((__float_as_int(RR2[g].x) << 23) + __float_as_int(PP2[g].x)),    \
                        __int_as_float((__float_as_int(RR2[g].y) << 23) + __float_as_int(PP2[g].y)));
#define LG_MUFU2X(g)
#define LG_MUFU2Y(g)
#else
#define LG_ARGS2_T2(g) T[g][2] = __ffma2_rn(make_float2(LG_V(g, 4), LG_V(g, 5)), c2, nms2);
#define LG_QA(g)
#define LG_QB(g)
#define LG_QC(g)
#define LG_QD(g)
#define LG_QE(g)
#define LG_QF(g)
#define LG_QG(g)
#define LG_MUFU2X(g) E[g][2].x = ex2_approx(T[g][2].x);
#define LG_MUFU2Y(g) E[g][2].y = ex2_approx(T[g][2].y);
#endif
#define LG_ARGS2(g)                                                                   \
  LG_ARGS2_T2(g)                                                                      \
  {                                                                                   \
    float ya, yb;                                                                     \
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(ya) : "f"(LG_V(g, 6)), "f"(cs252), "f"(bs252)); \
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(yb) : "f"(LG_V(g, 7)), "f"(cs252), "f"(bs252)); \
    Y[g] = make_float2(ya, yb);                                                       \
  }
#define LG_PA(g) RR[g] = __ffma2_rn(Y[g], k252, km);
#define LG_PB(g) FR[g] = __fadd2_rn(km, make_float2(-RR[g].x, -RR[g].y));
#define LG_PC(g) FR[g] = __ffma2_rn(Y[g], k252, FR[g]);
#define LG_PD(g) PP[g] = __ffma2_rn(make_float2(0.0551716685f, 0.0551716685f), FR[g], make_float2(0.2426111251f, 0.2426111251f));
#define LG_PE(g) PP[g] = __ffma2_rn(PP[g], FR[g], make_float2(0.6932609677f, 0.6932609677f));
#define LG_PF(g) PP[g] = __ffma2_rn(PP[g], FR[g], make_float2(0.9999280572f, 0.9999280572f));
#define LG_PG(g)                                                                                        \
  E[g][3] = make_float2(__int_as_float((__float_as_int(RR[g].x) << 23) + __float_as_int(PP[g].x)),      \
                        __int_as_float((__float_as_int(RR[g].y) << 23) + __float_as_int(PP[g].y)));
#define LG_PK(g, k) (((g) < 8 ? pk0 : pk1)[((g) & 7) * 4 + (k)])
            if (!pf) tmem_ld32(tS, va);
            pf = false;
            tmem_wait_ld();
            if (threadIdx.x == 0) TRACE(16, kv_base + j);
            tmem_ld32(tS + 32, vb);
            mask_chunk(va, 0);
            LG_ARGS1(0) LG_ARGS2(0) LG_PA(0) LG_QA(0) LG_PB(0) LG_QB(0) LG_PC(0) LG_QC(0) LG_PD(0) LG_QD(0) LG_PE(0) LG_QE(0) LG_PF(0) LG_QF(0) LG_PG(0) LG_QG(0)
#pragma unroll
            for (int g = 0; g < 16; ++g) {
              if (g == 3) { if (threadIdx.x == 0) TRACE(17, kv_base + j); tmem_wait_ld(); mask_chunk(vb, 1); }
              if (g == 7) { if (threadIdx.x == 0) TRACE(18, kv_base + j); tmem_wait_ld(); mask_chunk(va, 2); }
#if LG_ALTERNATE
              if (g == 8 && attempt == 0) {  // half way: once per block (the rare second attempt does not signal again)
                if (t == 0) asm volatile("bar.arrive 3, 256;" ::: "memory");
                else asm volatile("bar.arrive 4, 256;" ::: "memory");
              }
#endif
              if (g == 11) { if (threadIdx.x == 0) TRACE(19, kv_base + j); tmem_wait_ld(); mask_chunk(vb, 3); }
              E[g][0].x = ex2_approx(T[g][0].x);
              if (g + 1 < 16) { LG_ARGS1(g + 1) }
              E[g][0].y = ex2_approx(T[g][0].y);
              if (g + 1 < 16) { LG_ARGS2(g + 1) }
              if (g == 3) tmem_ld32(tS + 64, va);   // va is dead: the arguments of group 3 were its last readers
              if (g == 7) tmem_ld32(tS + 96, vb);
              E[g][1].x = ex2_approx(T[g][1].x);
              if (g + 1 < 16) { LG_PA(g + 1) LG_QA(g + 1) }
              if (g > 0) { s0 = __fadd2_rn(s0, E[g - 1][0]); LG_PK(g - 1, 0) = pack_bf16x2(E[g - 1][0].x, E[g - 1][0].y); }
              E[g][1].y = ex2_approx(T[g][1].y);
              if (g + 1 < 16) { LG_PB(g + 1) LG_QB(g + 1) }
              if (g > 0) { s1 = __fadd2_rn(s1, E[g - 1][1]); LG_PK(g - 1, 1) = pack_bf16x2(E[g - 1][1].x, E[g - 1][1].y); }
              LG_MUFU2X(g)
              if (g + 1 < 16) { LG_PC(g + 1) LG_QC(g + 1) LG_PD(g + 1) LG_QD(g + 1) }
              if (g > 0) { s2 = __fadd2_rn(s2, E[g - 1][2]); LG_PK(g - 1, 2) = pack_bf16x2(E[g - 1][2].x, E[g - 1][2].y); }
              LG_MUFU2Y(g)
              if (g + 1 < 16) { LG_PE(g + 1) LG_QE(g + 1) LG_PF(g + 1) LG_QF(g + 1) LG_PG(g + 1) LG_QG(g + 1) }
              if (g > 0) { s3 = __fadd2_rn(s3, E[g - 1][3]); LG_PK(g - 1, 3) = pack_bf16x2(E[g - 1][3].x, E[g - 1][3].y); }
            }
            s0 = __fadd2_rn(s0, E[15][0]); LG_PK(15, 0) = pack_bf16x2(E[15][0].x, E[15][0].y);
            s1 = __fadd2_rn(s1, E[15][1]); LG_PK(15, 1) = pack_bf16x2(E[15][1].x, E[15][1].y);
            s2 = __fadd2_rn(s2, E[15][2]); LG_PK(15, 2) = pack_bf16x2(E[15][2].x, E[15][2].y);
            s3 = __fadd2_rn(s3, E[15][3]); LG_PK(15, 3) = pack_bf16x2(E[15][3].x, E[15][3].y);
            tile_sum = ((s0.x + s0.y) + (s1.x + s1.y)) + ((s2.x + s2.y) + (s3.x + s3.y));
            if (threadIdx.x == 0) TRACE(20, kv_base + j);
#undef LG_V
#undef LG_ARGS1
#undef LG_ARGS2
#undef LG_PA
#undef LG_PB
#undef LG_PC
#undef LG_PD
#undef LG_PE
#undef LG_PF
#undef LG_PG
#undef LG_PK
          }
#else
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
#if LG_SWEEP == 0
          tile_sum = (s0.x + s0.y) + (s1.x + s1.y);
#else
          tile_sum = ((s0.x + s0.y) + (s1.x + s1.y)) + ((s2.x + s2.y) + (s3.x + s3.y));
#endif
#endif
          if (first || attempt == 1) break;
          // every P <= row sum: below the limit nothing can have overflowed the bf16 / fp32 ranges
          if (!__any_sync(0xffffffffu, !(tile_sum < kSumLimit))) break;
        }
        l += tile_sum;
        // The next block's similarity was queued three steps ahead and is normally complete by now: start loading its
        // first 32 columns (va is dead) so that the wake-up, fence and TMEM-load latency of block j+1 hide behind the
        // P store / hand-off of block j.  Non-blocking: if the similarity has not landed, block j+1 waits as before.
        if (LG_PREFETCH && j + 1 < it.nkt) {
          const uint32_t Gn = G + 2;
          pf = __all_sync(0xffffffffu, mbar_test_wait(&bar->s_full[Gn % 3], (Gn / 3) & 1));
          if (pf) {
            tc_fence_after();
            tmem_ld32(tS0 + (Gn % 3) * 128, va);
          }
        }
        // P overwrites the first 64 columns of its own S buffer (every S value has been consumed); the PV MMA reads
        // it from TMEM as the A operand, then the similarity of step G+3 recycles the buffer.
        tmem_st32(tS, pk0);
        tmem_st32(tS + 32, pk1);
        if (threadIdx.x == 0) TRACE(21, kv_base + j);
        tmem_wait_st();
        tc_fence_before();
        if (threadIdx.x == 0) TRACE(22, kv_base + j);
        // One arrival per WARP (barrier count 4), no warpgroup barrier: the warps of a warpgroup drift apart by less
        // than two blocks -- the similarity of block j+2 of this tile is queued behind PV(j) of the OTHER tile, which
        // the MMA warp issues after PV_t(j), i.e. after all four warps have arrived for block j -- and blocks j, j+1
        // use different barriers (below), so the four arrivals that complete a phase always belong to one block.
        // Two barriers per tile, alternating with the key block: the three-deep similarity lookahead lets a warp
        // finish blocks j and j+1 while the MMA warp is held up (at an item boundary it decodes the next items, which
        // with per-image lengths costs dependent global loads, and waits for their Q tile).  On ONE barrier those
        // arrivals would flip the phase twice and the MMA warp's parity wait for block j would never return; block
        // j+2 (the same barrier again) cannot be reached before PV(j) has consumed P(j).
#if LG_WARP_ARRIVE
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar->p_full[2 * t + ((kv_base + j) & 1)]);
#else
        asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");
        if ((threadIdx.x & 127) == 0) mbar_arrive(&bar->p_full[2 * t + ((kv_base + j) & 1)]);
#endif
        if ((threadIdx.x & 127) == 0) TRACE(t * 3 + 2, kv_base + j);
      }

      // epilogue: O / l.  O_t is copied to registers and released first so the next item's PV can start.
      if ((threadIdx.x & 127) == 0) TRACE(8 + t * 4, kv_base);
      if ((threadIdx.x & 127) == 0) DBG(10 + t, 3);
      // pv_done[t] completes once per (non-skipped) item, behind the item's last PV_t; the next completion needs this
      // warpgroup's P of the next item, so the barrier is never more than one completion away from this wait.
      LG_WAIT(&bar->pv_done[t], item_cnt & 1);
      ++item_cnt;
      if ((threadIdx.x & 127) == 0) DBG(10 + t, 4);
      tc_fence_after();
      if ((threadIdx.x & 127) == 0) TRACE(9 + t * 4, kv_base);
      tmem_ld32(tO, va);
      tmem_ld32(tO + 32, vb);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(&bar->o_free[t]);
      if ((threadIdx.x & 127) == 0) TRACE(10 + t * 4, kv_base);
      kv_base += it.nkt;
      const float inv = (row < it.q_len) ? 1.f / l : 0.f;  // padded query rows -> 0
      if (dt) {
        // context rows as split bf16 tiles (head h = K atom h of the rows x 256 matrix).  A single bf16 is NOT
        // enough here: rounding the context costs 1e-2 on log_assignment (P's rounding averages out over the
        // keys, the context's does not).  The warp's 32 rows are a contiguous 4 KB slice of the 16 KB hi (and
        // lo) tile image: stage in shared memory, one bulk store each.
        uint8_t* stg = sOut + warp * 4096;
        const uint32_t stg_row = smem_u32(stg) + lane * 128;
        uint32_t lo_w[32];
        if (lane == 0) bulk_wait_read_all();  // previous item's stores have drained the staging slice
        __syncwarp();
        auto split_store = [&](const uint32_t(&v)[32], const int cb) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint32_t hh[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float x0 = __uint_as_float(v[8 * q + 2 * i]) * inv, x1 = __uint_as_float(v[8 * q + 2 * i + 1]) * inv;
              const __nv_bfloat16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
              hh[i] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
              lo_w[cb * 16 + q * 4 + i] = pack_bf16x2(x0 - __bfloat162float(h0), x1 - __bfloat162float(h1));
            }
            st_shared_v4(stg_row + (((cb * 4 + q) ^ (lane & 7)) << 4), make_uint4(hh[0], hh[1], hh[2], hh[3]));
          }
        };
        split_store(va, 0);
        split_store(vb, 1);
        fence_proxy_async_smem();
        __syncwarp();
        uint8_t* gdst = dt + (warp & 3) * 4096;
        if (lane == 0) { bulk_s2g(gdst, stg, 4096); bulk_commit(); bulk_wait_read_all(); }
        __syncwarp();
#pragma unroll
        for (int ch = 0; ch < 8; ++ch)
          st_shared_v4(stg_row + ((ch ^ (lane & 7)) << 4),
                       make_uint4(lo_w[4 * ch], lo_w[4 * ch + 1], lo_w[4 * ch + 2], lo_w[4 * ch + 3]));
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) { bulk_s2g(gdst + 16384, stg, 4096); bulk_commit(); }
        if ((threadIdx.x & 127) == 0) TRACE(11 + t * 4, kv_base - it.nkt);
      } else if (row < side_nq) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          drow[i] = make_float4(__uint_as_float(va[4 * i]) * inv, __uint_as_float(va[4 * i + 1]) * inv,
                                __uint_as_float(va[4 * i + 2]) * inv, __uint_as_float(va[4 * i + 3]) * inv);
          drow[8 + i] = make_float4(__uint_as_float(vb[4 * i]) * inv, __uint_as_float(vb[4 * i + 1]) * inv,
                                    __uint_as_float(vb[4 * i + 2]) * inv, __uint_as_float(vb[4 * i + 3]) * inv);
        }
      }
    }
#if LG_ALTERNATE
    if (t == 0 && kv_base > 0) asm volatile("bar.sync 4, 256;" ::: "memory");  // absorbs tile 1's last half-way signal
#endif
    if (lane == 0) bulk_wait_all();  // outstanding context stores
    if ((threadIdx.x & 127) == 0) DBG(10 + t, 7);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kSoft + 1) tmem_dealloc<512>(tmem);
}

}  // namespace lg

struct AttnSideArgs {
  const void *q_tiles, *k_tiles, *v_tiles;
  float* out;
  int out_rt_base;
  const int *q_len, *k_len;
  int nq, nk, tq_pad, tk_pad;
};

static int check_side(const AttnSideArgs& a, bool tiles) {
  if (!a.q_tiles || !a.k_tiles || !a.v_tiles || (!tiles && !a.out)) return LG_ERR_NULL;
  if (a.nq <= 0 || a.nk <= 0 || a.out_rt_base < 0) return LG_ERR_SHAPE;
  if (a.tq_pad * lg::kTile < a.nq || a.tk_pad * lg::kTile < a.nk || (a.tq_pad & 1) || (a.tk_pad & 1)) return LG_ERR_SHAPE;
  return 0;
}

static int launch_attention(const AttnSideArgs* sides, int n_sides, void* out_tiles, int B, int H, float softmax_scale,
                            cudaStream_t stream) {
  if (B <= 0 || H <= 0 || n_sides < 1 || n_sides > 2) return LG_ERR_SHAPE;
  static bool attr_done_dev[lg::kMaxDevices] = {};  // per device; a benign race sets it twice
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  const int sms = lg::device_sm_count();
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(lg::attention_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         lg::kAttnSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(lg::attention_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lg::kAttnSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  lg::AttnParams p{};
  p.out_tiles = static_cast<uint8_t*>(out_tiles);
  p.B = B; p.H = H; p.scale_log2 = softmax_scale * 1.4426950408889634f;
  long long items = 0;
  for (int i = 0; i < n_sides; ++i) {
    if (int e = check_side(sides[i], out_tiles != nullptr)) return e;
    const AttnSideArgs& a = sides[i];
    p.side[i] = lg::AttnSide{static_cast<const uint8_t*>(a.q_tiles), static_cast<const uint8_t*>(a.k_tiles),
                             static_cast<const uint8_t*>(a.v_tiles), a.out, a.out_rt_base, a.q_len, a.k_len, a.nq, a.nk,
                             a.tq_pad, a.tk_pad, B * H * (a.tq_pad / 2)};
    items += p.side[i].items;
  }
  if (items > 0x7fffffffLL) return LG_ERR_SHAPE;
  p.items = static_cast<int>(items);
  static const bool split = [] {  // LG_ATTN_SPLIT=0: the two-warpgroup softmax (one thread per row)
    const char* e = getenv("LG_ATTN_SPLIT");
    return !(e && e[0] == '0');
  }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(p.items < sms ? p.items : sms));
  cfg.blockDim = dim3(split ? 640 : lg::kAttnThreads);
  cfg.dynamicSmemBytes = lg::kAttnSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  cfg.attrs = attr;
  cfg.numAttrs = lg::add_pdl_attr(attr, 0);
  if (split) return static_cast<int>(cudaLaunchKernelEx(&cfg, lg::attention_kernel<true>, p));
  return static_cast<int>(cudaLaunchKernelEx(&cfg, lg::attention_kernel<false>, p));
}

extern "C" int lg_attention(const void* q_tiles, const void* k_tiles, const void* v_tiles, float* out,
                            const int* q_len, const int* k_len, int B, int H, int nq, int nk, int tq_pad, int tk_pad,
                            float softmax_scale, cudaStream_t stream) {
  if (!out) return LG_ERR_NULL;
  const AttnSideArgs a{q_tiles, k_tiles, v_tiles, out, 0, q_len, k_len, nq, nk, tq_pad, tk_pad};
  return launch_attention(&a, 1, nullptr, B, H, softmax_scale, stream);
}

extern "C" int lg_attention_tiles(const void* q_tiles, const void* k_tiles, const void* v_tiles, void* out_tiles,
                                  int out_rt_base, const int* q_len, const int* k_len, int B, int H, int nq, int nk,
                                  int tq_pad, int tk_pad, float softmax_scale, cudaStream_t stream) {
  if (!out_tiles) return LG_ERR_NULL;
  const AttnSideArgs a{q_tiles, k_tiles, v_tiles, nullptr, out_rt_base, q_len, k_len, nq, nk, tq_pad, tk_pad};
  return launch_attention(&a, 1, out_tiles, B, H, softmax_scale, stream);
}

// Both images of a SelfBlock (side 0: image 0 attends to itself, side 1: image 1) or both directions of a CrossBlock
// (side 0: queries of image 0 over the keys / values of image 1, side 1: the reverse) in ONE launch.
extern "C" int lg_attention_tiles2(const void* q0, const void* k0, const void* v0, int rt_base0, const int* q_len0,
                                   const int* k_len0, int nq0, int nk0, int tq_pad0, int tk_pad0, const void* q1,
                                   const void* k1, const void* v1, int rt_base1, const int* q_len1, const int* k_len1,
                                   int nq1, int nk1, int tq_pad1, int tk_pad1, void* out_tiles, int B, int H,
                                   float softmax_scale, cudaStream_t stream) {
  if (!out_tiles) return LG_ERR_NULL;
  const AttnSideArgs a[2] = {{q0, k0, v0, nullptr, rt_base0, q_len0, k_len0, nq0, nk0, tq_pad0, tk_pad0},
                             {q1, k1, v1, nullptr, rt_base1, q_len1, k_len1, nq1, nk1, tq_pad1, tk_pad1}};
  return launch_attention(a, 2, out_tiles, B, H, softmax_scale, stream);
}

#ifdef LG_ATTN_DBG
extern "C" int lg_debug_attn_set(int* host_mapped) {
  return static_cast<int>(cudaMemcpyToSymbol(lg::g_dbg, &host_mapped, sizeof(host_mapped)));
}
#endif

#ifdef LG_ATTN_TRACE
extern "C" int lg_debug_attn_trace(unsigned long long* host_out, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host_out, lg::g_trace, sizeof(unsigned long long) * n));
}
#endif
