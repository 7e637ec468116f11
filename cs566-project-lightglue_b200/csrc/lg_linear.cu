// lg_linear : the per-token linears of SelfBlock / CrossBlock / MatchAssignment on tcgen05.
//
// Replaces the nn.Linear calls of the reference (gluefactory/models/matchers/lightglue.py):
//   SelfBlock   Wqkv :184 (+ unflatten/rotary :185-188), out_proj :190, ffn.0 / ffn.3 :171-176,191
//   CrossBlock  to_qk / to_v :223-228, to_out :246, ffn.0 / ffn.3 :205-210,247-248
//   MatchAssignment.final_proj :310, LightGlue.input_proj :370-373,478-479
//
// Precision: bf16 tensor-core operands alone cost 1e-1 on log_assignment when used in the FFN
// (scripts/exp_linear_precision.py), so every linear runs as a 3-term split-bf16 product
//   x = xh + xl, w = wh + wl (bf16 each):  y = xh.wh + xl.wh + xh.wl     (fp32 accumulate in TMEM)
// which carries ~16 mantissa bits (2e-4 end to end) at 3 bf16 MMAs per product.
//
// Data layout ("split tiles"): an activation matrix [rows][K] lives in HBM as
//   [row tile of 128][K atom of 64][hi 16 KB | lo 16 KB]
// where each 16 KB block is the exact byte image of a SWIZZLE_128B K-major shared-memory tile,
// so one flat bulk (TMA engine) copy stages a 32 KB hi|lo atom.  Weights [Nout][K] are packed once
// as [N tile][K atom][hi | lo][Nt x 128 B] in the same pattern (lg_pack_weight).
//
// One persistent CTA per SM (or CTA pair per TPC), 320 threads:
//   warp 8 (one lane)  producer: bulk copies A atom (32 KB) + W atom into a 2..4-stage ring
//   warp 9 (one lane)  tcgen05.mma issue, 128 (pair: 256) x Nt x 16 per instruction, 12 per stage;
//                      two TMEM accumulators (2 x 256 columns) so tile i+1 accumulates while
//   warps 0-7          the epilogue of tile i runs: 2 threads per row (= TMEM lane), 32 of every 64 columns each
//                      (4 warps were not enough: at K = 256 the epilogue, not the MMA, set the tile period)
// Epilogues (template): split tiles (+bias, +residual), fp32 rows, and the attention operand
// producers (self: de-interleaved Wqkv + rotary; cross: to_qk | to_v) that write the bf16
// Q / K / V^T tiles lg_attention consumes, through shared-memory staging + bulk stores.
#include <cstdlib>

#include "lg_common.cuh"
#include <cuda_fp16.h>
#include "lg_tiles.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kLinThreads = 320;  // warps 0-7 epilogue, 8 producer, 9 MMA issue
constexpr int kLinMaxStages = 4;
constexpr int kLinSmemMax = 232448;  // 227 KB per CTA
constexpr int kStageOut = 32768;    // epilogue staging (one split atom, or two bf16 tiles)

enum LinearEpilogue { kEpiSplit = 0, kEpiF32 = 1, kEpiQkvSelf = 2, kEpiQkvCross = 3, kEpiF16 = 4 };

struct LinearParams {
  const uint8_t* a0;       // A source 0: [row_tiles][a0_atoms][32 KB]
  const uint8_t* a1;       // A source 1 (concatenated along K after a0), may be null
  int a0_atoms, a1_atoms;
  int a1_hi_only;          // a1 holds plain bf16 tiles [row_tiles][a1_atoms][16 KB] (no lo half): 2-term product
  int one_term;            // hi . hi only: the lo halves of A and W are not even staged (half the operand bytes, a third
                           // of the MMAs) -- for outputs that are rounded to bf16 attention operands anyway
  const uint8_t* w;        // [n_tiles][k_atoms][hi | lo][Nt*128]
  const float* bias;       // [n_tiles * Nt] (same row order as w)
  int row_tiles, n_tiles;
  // kEpiSplit
  uint8_t* out_tiles;            // [row_tiles][n_tiles*Nt/64][32 KB]
  const uint8_t* resid_tiles;    // same shape, or null
  // kEpiF32
  float* out_f32;
  int64_t ld;                    // row pitch in floats
  // kEpiQkv*: attention operand tiles per side (image 0 rows first, then image 1)
  uint8_t* q_tiles[2];
  uint8_t* k_tiles[2];
  uint8_t* v_tiles[2];
  const uint8_t* rot_tiles;      // [row_tiles][cos 16 KB | sin 16 KB], fp32, chunk-swizzled rows
  int B, H, tp0, tp1;            // row tile rt -> side 0: (b, t) = divmod(rt, tp0); side 1 after B*tp0
};

struct __align__(8) LinearBarriers {
  uint64_t full[kLinMaxStages];
  uint64_t peer_full[kLinMaxStages];  // even CTA of a pair: the odd CTA's stage has landed (relayed arrival)
  uint64_t empty[kLinMaxStages];
  uint64_t acc_full[2];
  uint64_t acc_empty[2];
  uint64_t aux_full;     // residual atom / rotary table landed
  uint64_t resid_full[3];  // pair mode: residual atom a landed in buffer a % 3
  uint32_t tmem_base;
  uint32_t relay_sink[kLinMaxStages];  // target words of the odd CTA's st.async relays
};

// kCtas == 2: a CTA pair (cluster of 2 on one TPC) computes two row tiles with one 256-row MMA; each CTA stages its
// own A atom but only HALF of the weight atom, which cuts the L2 -> shared-memory traffic per flop by a third and
// leaves room for a deeper ring.
template <int kNt, int kEpi, int kCtas>
constexpr int linear_stage_bytes() { return kSplitAtom + kNt * 256 / kCtas; }
// kEpiSplit with a CTA pair: the half-size weight stage leaves room for a double-buffered residual atom, fetched
// one atom ahead of the epilogue (single CTA: the residual lands in the output staging buffer on demand).
template <int kEpi, int kCtas>
constexpr bool resid_prefetch() { return kEpi == kEpiSplit && kCtas == 2; }
template <int kNt, int kEpi, int kCtas>
constexpr int linear_fixed_bytes() {
  return 1024 + kStageOut + (kEpi == kEpiQkvSelf ? 32768 : 0) + (resid_prefetch<kEpi, kCtas>() ? 2 * kSplitAtom : 0) +
         static_cast<int>(sizeof(LinearBarriers));
}
template <int kNt, int kEpi, int kCtas>
constexpr int linear_stages() {
  constexpr int n = (kLinSmemMax - linear_fixed_bytes<kNt, kEpi, kCtas>()) / linear_stage_bytes<kNt, kEpi, kCtas>();
  return n > kLinMaxStages ? kLinMaxStages : n;
}
template <int kNt, int kEpi, int kCtas>
constexpr int linear_smem_bytes() {
  return linear_fixed_bytes<kNt, kEpi, kCtas>() +
         linear_stages<kNt, kEpi, kCtas>() * linear_stage_bytes<kNt, kEpi, kCtas>();
}

// the accumulator buffer of this CTA has been read: tell the MMA issuer (which lives in the even CTA of a pair)
// (a remote arrive carries a cluster-scope release = MEMBAR.ALL.GPU: one lane per warp pays it)
__device__ __forceinline__ void acc_release(uint64_t* bar, uint32_t rank) {
  if (rank == 0) {
    mbar_arrive(bar);
  } else {
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive_remote(bar, 0);
  }
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// attention tile address of (row tile rt, head h) inside the per-side [B][H][tp][16 KB] buffers
__device__ __forceinline__ size_t attn_tile_offset(const LinearParams& p, int rt, int h, int& side) {
  int b, t, tp;
  if (rt < p.B * p.tp0) { side = 0; tp = p.tp0; b = rt / tp; t = rt - b * tp; }
  else { side = 1; tp = p.tp1; const int r2 = rt - p.B * p.tp0; b = r2 / tp; t = r2 - b * tp; }
  return ((static_cast<size_t>(b) * p.H + h) * tp + t) * static_cast<size_t>(kHalfAtom);
}

template <int kNt, int kEpi, int kCtas>
__global__ void __launch_bounds__(kLinThreads, 1) linear_kernel(const LinearParams p) {
  constexpr int kWAtom = kNt * 256;                       // hi | lo of one K atom of the weight tile
  constexpr int kWPart = kNt * 128 / kCtas;               // this CTA's rows of the hi (or lo) half
  constexpr int kStage = linear_stage_bytes<kNt, kEpi, kCtas>();
  constexpr int kLinStages = linear_stages<kNt, kEpi, kCtas>();
  static_assert(kLinStages >= 2, "ring needs two stages");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // 1 KB aligned, still provably shared
  uint8_t* ring = base;
  uint8_t* stg = ring + kLinStages * kStage;              // 32 KB epilogue staging (kEpiSplit, pair mode: 3 x 32 KB)
  uint8_t* rot = stg + kStageOut;                         // 32 KB rotary table (kEpiQkvSelf only)
  LinearBarriers* bar = reinterpret_cast<LinearBarriers*>(
      stg + kStageOut + (kEpi == kEpiQkvSelf ? 32768 : 0) + (resid_prefetch<kEpi, kCtas>() ? 2 * kSplitAtom : 0));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_atoms = p.a0_atoms + p.a1_atoms;
  // work item = (unit of kCtas adjacent row tiles, n tile); the pair walks the same item list
  const int items = (p.row_tiles / kCtas) * p.n_tiles;
  const uint32_t rank = kCtas == 2 ? cluster_ctarank() : 0;
  const int first = blockIdx.x / kCtas, stride = gridDim.x / kCtas;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kLinStages; ++s) {
      mbar_init(&bar->full[s], 1);
      mbar_init(&bar->peer_full[s], 1);
      mbar_init(&bar->empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) { mbar_init(&bar->acc_full[i], 1); mbar_init(&bar->acc_empty[i], kCtas == 2 ? 256 + 8 : 256); }
    mbar_init(&bar->aux_full, 1);
    for (int i = 0; i < 3; ++i) mbar_init(&bar->resid_full[i], 1);
    fence_mbar_init();
  }
  if (warp == 9) {
    if (kCtas == 2) tmem_alloc_pair<512>(&bar->tmem_base);
    else tmem_alloc<512>(&bar->tmem_base);
  }
  tc_fence_before();
  if (kCtas == 2) cluster_sync_all();  // peer barriers are initialised before any remote arrive / multicast commit
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  pdl_launch_dependents();
  pdl_wait();  // the activations of the previous kernel are complete and visible from here on

  if (warp == 8) {
    // ------------------------------------------------------------------ producer
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = first; item < items; item += stride) {
        const int unit = item / p.n_tiles, nt = item - unit * p.n_tiles;
        const int rt = unit * kCtas + rank;
        const uint8_t* gw = p.w + static_cast<size_t>(nt) * k_atoms * kWAtom + rank * kWPart;
        for (int ka = 0; ka < k_atoms; ++ka, ++it) {
          const int s = it % kLinStages;
          if (it >= kLinStages) mbar_wait(&bar->empty[s], ((it / kLinStages) - 1) & 1);
          const bool hi_only = ka >= p.a0_atoms && p.a1_hi_only;
          const uint8_t* ga = ka < p.a0_atoms
                                  ? p.a0 + (static_cast<size_t>(rt) * p.a0_atoms + ka) * kSplitAtom
                                  : p.a1 + (static_cast<size_t>(rt) * p.a1_atoms + (ka - p.a0_atoms)) *
                                               (hi_only ? kHalfAtom : kSplitAtom);
          const bool one = p.one_term != 0;
          const uint32_t a_bytes = (hi_only || one) ? kHalfAtom : kSplitAtom;
          mbar_arrive_expect_tx(&bar->full[s], a_bytes + (one ? 1 : 2) * kWPart);
          bulk_g2s(ring + s * kStage, ga, a_bytes, &bar->full[s]);
          if (kCtas == 1) {  // hi block first, lo block behind it
            bulk_g2s(ring + s * kStage + kSplitAtom, gw + static_cast<size_t>(ka) * kWAtom, one ? kWAtom / 2 : kWAtom,
                     &bar->full[s]);
          } else {  // this CTA's half of the output columns: rows [rank * Nt/2, +Nt/2) of the hi and of the lo block
            bulk_g2s(ring + s * kStage + kSplitAtom, gw + static_cast<size_t>(ka) * kWAtom, kWPart, &bar->full[s]);
            if (!one)
              bulk_g2s(ring + s * kStage + kSplitAtom + kWPart, gw + static_cast<size_t>(ka) * kWAtom + kNt * 128, kWPart,
                       &bar->full[s]);
          }
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------------ MMA issuer
    const bool leader = elect_one();  // all lanes run the loop (operands stay uniform); one lane issues
    if (kCtas == 2 && rank != 0) {
      // odd CTA of the pair: relay "my stage has landed" to the even CTA, which issues the MMAs for both
      uint32_t it = 0;
      for (int item = first; item < items; item += stride) {
        for (int ka = 0; ka < k_atoms; ++ka, ++it) {
          const int s = it % kLinStages;
          mbar_wait(&bar->full[s], (it / kLinStages) & 1);
          // st.async + complete_tx on the even CTA's barrier: no fence (a remote mbarrier.arrive carries a
          // cluster-scope release = MEMBAR.ALL.GPU); the operands themselves were written by the async proxy
          if (leader) st_async_remote(&bar->relay_sink[s], it, &bar->peer_full[s], 0);
          __syncwarp();
        }
      }
    } else {
      constexpr uint32_t idesc = umma_idesc(kFmtBF16, 128 * kCtas, kNt);
      const uint32_t ring_a = smem_u32(ring);
      uint32_t it = 0;
      int local = 0;
      for (int item = first; item < items; item += stride, ++local) {
        const int buf = local & 1;
        if (local >= 2) {
          mbar_wait(&bar->acc_empty[buf], ((local >> 1) - 1) & 1);
          tc_fence_after();
        }
        const uint32_t acc = tmem + buf * 256;
        for (int ka = 0; ka < k_atoms; ++ka, ++it) {
          const int s = it % kLinStages;
          mbar_wait(&bar->full[s], (it / kLinStages) & 1);
          if (kCtas == 2) {
            if (leader) mbar_arrive_expect_tx(&bar->peer_full[s], 4);  // armed per phase; the relay may land first
            mbar_wait(&bar->peer_full[s], (it / kLinStages) & 1);
          }
          tc_fence_after();
          const bool one = p.one_term != 0;
          const bool hi_only = ka >= p.a0_atoms && p.a1_hi_only;
          const uint32_t a_hi = ring_a + s * kStage, a_lo = a_hi + kHalfAtom;
          const uint32_t w_hi = a_hi + kSplitAtom, w_lo = w_hi + kWPart;
          const uint64_t dah = umma_desc_kmajor_sw128(a_hi), dal = umma_desc_kmajor_sw128(a_lo);
          const uint64_t dwh = umma_desc_kmajor_sw128(w_hi), dwl = umma_desc_kmajor_sw128(w_lo);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {  // 64 bf16 per atom = 4 x K16, 32 B per step
            if (kCtas == 2) {
              if (leader && !hi_only && !one) umma_bf16_ss_pair(acc, dal + 2 * kk, dwh + 2 * kk, idesc, (ka > 0 || kk > 0));
              if (leader && !one) umma_bf16_ss_pair(acc, dah + 2 * kk, dwl + 2 * kk, idesc, hi_only ? (ka > 0 || kk > 0) : 1);
              if (leader) umma_bf16_ss_pair(acc, dah + 2 * kk, dwh + 2 * kk, idesc, one ? (ka > 0 || kk > 0) : 1);
            } else {
              if (leader && !hi_only && !one) umma_bf16_ss(acc, dal + 2 * kk, dwh + 2 * kk, idesc, (ka > 0 || kk > 0));
              if (leader && !one) umma_bf16_ss(acc, dah + 2 * kk, dwl + 2 * kk, idesc, hi_only ? (ka > 0 || kk > 0) : 1);
              if (leader) umma_bf16_ss(acc, dah + 2 * kk, dwh + 2 * kk, idesc, one ? (ka > 0 || kk > 0) : 1);
            }
          }
          if (leader) { if (kCtas == 2) umma_commit_pair(&bar->empty[s]); else umma_commit(&bar->empty[s]); }
        }
        if (leader) { if (kCtas == 2) umma_commit_pair(&bar->acc_full[buf]); else umma_commit(&bar->acc_full[buf]); }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue: 2 threads per row; thread (r, g)
    // owns columns [32 g, 32 g + 32) of every 64-column group.  TMEM loads run one group ahead of the arithmetic
    // (packed f32x2 instructions throughout).
    const int q = warp & 3, g = warp >> 2;
    const int r = q * 32 + lane;
    const bool leader = threadIdx.x == 0;
    constexpr int kGroups = kNt / 64;
    uint32_t aux_phase = 0, sbi = 0;  // sbi: running half-buffer index of the bf16 tile staging
    uint32_t ra = 0;                  // running residual atom index (pair mode prefetch)
    uint32_t v[32], u[32];
    int local = 0;
    for (int item = first; item < items; item += stride, ++local) {
      const int unit = item / p.n_tiles, nt = item - unit * p.n_tiles;
      const int rt = unit * kCtas + rank;
      const int buf = local & 1;
      const uint32_t tacc = tmem + (static_cast<uint32_t>(q * 32) << 16) + buf * 256 + g * 32;
      const float* bias = p.bias + nt * kNt + g * 32;

      if (kEpi == kEpiQkvSelf) {
        // rotary table of this row tile (all reads of the previous table ended before the
        // previous item's last epi_bar_sync)
        if (leader) {
          mbar_arrive_expect_tx(&bar->aux_full, 32768);
          bulk_g2s(rot, p.rot_tiles + static_cast<size_t>(rt) * 32768, 32768, &bar->aux_full);
        }
      }
      mbar_wait(&bar->acc_full[buf], (local >> 1) & 1);
      tc_fence_after();
      tmem_ld32(tacc, v);
      // group j of the accumulator -> u (+ nothing yet); group j + 1 requested; after the last group the buffer is free
      auto next_group = [&](int j) {
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) u[i] = v[i];
        if (j + 1 < kGroups) {
          tmem_ld32(tacc + (j + 1) * 64, v);
        } else {
          tc_fence_before();
          acc_release(&bar->acc_empty[buf], rank);
        }
      };
      // x[0..3] = columns [8 c, 8 c + 8) of this thread's 32 of group j, bias added
      auto biased8 = [&](int j, int c, float2* x) {
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + j * 64 + c * 8));
        const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + j * 64 + c * 8) + 1);
        const uint32_t* uu = u + c * 8;
        x[0] = __fadd2_rn(make_float2(__uint_as_float(uu[0]), __uint_as_float(uu[1])), make_float2(b0.x, b0.y));
        x[1] = __fadd2_rn(make_float2(__uint_as_float(uu[2]), __uint_as_float(uu[3])), make_float2(b0.z, b0.w));
        x[2] = __fadd2_rn(make_float2(__uint_as_float(uu[4]), __uint_as_float(uu[5])), make_float2(b1.x, b1.y));
        x[3] = __fadd2_rn(make_float2(__uint_as_float(uu[6]), __uint_as_float(uu[7])), make_float2(b1.z, b1.w));
      };

      if (kEpi == kEpiF32) {
        float* dst = p.out_f32 + (static_cast<int64_t>(rt) * kTile + r) * p.ld + nt * kNt + g * 32;
#pragma unroll 1
        for (int j = 0; j < kGroups; ++j) {
          next_group(j);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            float2 x[4];
            biased8(j, c, x);
            reinterpret_cast<float4*>(dst + j * 64 + c * 8)[0] = make_float4(x[0].x, x[0].y, x[1].x, x[1].y);
            reinterpret_cast<float4*>(dst + j * 64 + c * 8)[1] = make_float4(x[2].x, x[2].y, x[3].x, x[3].y);
          }
        }
      } else if (kEpi == kEpiSplit) {
        const int out_atoms = p.n_tiles * kGroups;
        constexpr bool kPrefetch = resid_prefetch<kEpi, kCtas>();
        const bool has_resid = p.resid_tiles != nullptr;
#pragma unroll 1
        for (int j = 0; j < kGroups; ++j) {
          const size_t goff = (static_cast<size_t>(rt) * out_atoms + nt * kGroups + j) * kSplitAtom;
          uint8_t* ob = stg;  // buffer of this atom: residual in, result in place, bulk store out
          if (kPrefetch) {
            // three 32 KB buffers in rotation (atom ra uses ra % 3): the residual of atom ra + 1 is fetched while atom
            // ra is computed and atom ra - 1 is still being stored; nobody waits for the store just issued.
            if (leader) {
              bulk_wait_read<1>();  // stores up to atom ra - 2 have drained: buffer (ra + 1) % 3 is free
              if (has_resid) {
                if (ra == 0) {
                  mbar_arrive_expect_tx(&bar->resid_full[0], kSplitAtom);
                  bulk_g2s(stg, p.resid_tiles + goff, kSplitAtom, &bar->resid_full[0]);
                }
                size_t next = goff + kSplitAtom;  // next atom of this item, else atom 0 of the next item
                bool more = true;
                if (j == kGroups - 1) {
                  const int item2 = item + stride;
                  more = item2 < items;
                  const int unit2 = item2 / p.n_tiles, nt2 = item2 - unit2 * p.n_tiles;
                  next = (static_cast<size_t>(unit2 * kCtas + rank) * out_atoms + nt2 * kGroups) * kSplitAtom;
                }
                if (more) {
                  const uint32_t sl = (ra + 1) % 3;
                  mbar_arrive_expect_tx(&bar->resid_full[sl], kSplitAtom);
                  bulk_g2s(stg + sl * kSplitAtom, p.resid_tiles + next, kSplitAtom, &bar->resid_full[sl]);
                }
              }
            }
            if (has_resid) mbar_wait(&bar->resid_full[ra % 3], (ra / 3) & 1);
            else epi_bar_sync();
            ob = stg + (ra % 3) * kSplitAtom;
            ++ra;
          } else {
            if (leader) {
              bulk_wait_read<0>();  // the previous atom's store has drained the staging buffer
              if (has_resid) {
                mbar_arrive_expect_tx(&bar->aux_full, kSplitAtom);
                bulk_g2s(stg, p.resid_tiles + goff, kSplitAtom, &bar->aux_full);
              }
            }
            if (has_resid) { mbar_wait(&bar->aux_full, aux_phase); aux_phase ^= 1; }
            else epi_bar_sync();
          }
          next_group(j);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint8_t* at = ob + sw128_offset(r, g * 4 + c);
            float2 x[4];
            biased8(j, c, x);
            if (has_resid) {
              const uint4 rh = *reinterpret_cast<const uint4*>(at), rl = *reinterpret_cast<const uint4*>(at + kHalfAtom);
              const uint32_t hh[4] = {rh.x, rh.y, rh.z, rh.w}, ll[4] = {rl.x, rl.y, rl.z, rl.w};
#pragma unroll
              for (int i = 0; i < 4; ++i)
                x[i] = __fadd2_rn(__fadd2_rn(make_float2(bf16lo(hh[i]), bf16hi(hh[i])),
                                             make_float2(bf16lo(ll[i]), bf16hi(ll[i]))), x[i]);
            }
            uint4 hi, lo;
            split8(x, hi, lo);
            *reinterpret_cast<uint4*>(at) = hi;
            *reinterpret_cast<uint4*>(at + kHalfAtom) = lo;
          }
          fence_proxy_async_smem();
          epi_bar_sync();
          if (leader) { bulk_s2g(p.out_tiles + goff, ob, kSplitAtom); bulk_commit(); }
        }
      } else if (kEpi == kEpiF16) {
        // ---------------------------------------------------------------- plain fp16 tiles [rt][n_out/64][16 KB]
        // (final_proj for the single-product fp16 similarity of lg_assign_tiles; values clamped to the fp16 range)
        const int out_atoms = p.n_tiles * kGroups;
#pragma unroll 1
        for (int j = 0; j < kGroups; ++j) {
          uint8_t* sb = stg + (sbi++ & 1) * kHalfAtom;
          if (leader) bulk_wait_read<1>();  // the store issued two atoms ago has drained this half
          epi_bar_sync();
          next_group(j);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            float2 x[4];
            biased8(j, c, x);
            uint32_t h[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const __half2 v = __floats2half2_rn(fminf(fmaxf(x[i].x, -65504.f), 65504.f),
                                                  fminf(fmaxf(x[i].y, -65504.f), 65504.f));
              h[i] = *reinterpret_cast<const uint32_t*>(&v);
            }
            *reinterpret_cast<uint4*>(sb + sw128_offset(r, g * 4 + c)) = make_uint4(h[0], h[1], h[2], h[3]);
          }
          fence_proxy_async_smem();
          epi_bar_sync();
          if (leader) {
            bulk_s2g(p.out_tiles + (static_cast<size_t>(rt) * out_atoms + nt * kGroups + j) * kHalfAtom, sb, kHalfAtom);
            bulk_commit();
          }
        }
      } else {
        // ---------------------------------------------------------------- attention operand tiles
        // kEpiQkvSelf : n tile == head; columns [q 64 | k 64 | v 64]; rotary on q, k
        // kEpiQkvCross: n tile 0 = qk of the 4 heads, n tile 1 = v of the 4 heads
        if (kEpi == kEpiQkvSelf) { mbar_wait(&bar->aux_full, aux_phase); aux_phase ^= 1; }
#pragma unroll 1
        for (int part = 0; part < kGroups; ++part) {
          const int h = kEpi == kEpiQkvSelf ? nt : part;
          const int kind = kEpi == kEpiQkvSelf ? part : (nt == 0 ? 0 : 2);  // 0 q, 1 k, 2 v
          uint8_t* sb = stg + (sbi++ & 1) * kHalfAtom;
          if (leader) bulk_wait_read<1>();  // the store issued two parts ago has drained this half
          epi_bar_sync();
          next_group(part);
          float2 f[16];
#pragma unroll
          for (int c = 0; c < 4; ++c) biased8(part, c, f + 4 * c);
          if (kEpi == kEpiQkvSelf && kind < 2) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {  // 4 frequencies per 16-byte table chunk
              const uint32_t off = r * 128 + (((g * 4 + c) ^ (r & 7)) << 4);
              const float4 c4 = *reinterpret_cast<const float4*>(rot + off);
              const float4 s4 = *reinterpret_cast<const float4*>(rot + kHalfAtom + off);
              const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                // (t*cos) + (rot(t)*sin), reference rounding sequence (lightglue.py:36-43): two roundings of the
                // products, one of the sum (the packed multiplies / add are never contracted)
                const float2 x = f[4 * c + e];
                f[4 * c + e] = __fadd2_rn(__fmul2_rn(x, make_float2(cc[e], cc[e])),
                                          __fmul2_rn(make_float2(-x.y, x.x), make_float2(ss[e], ss[e])));
              }
            }
          }
          if (kind < 2) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint4 o;
              o.x = pack_bf16x2(f[4 * c].x, f[4 * c].y); o.y = pack_bf16x2(f[4 * c + 1].x, f[4 * c + 1].y);
              o.z = pack_bf16x2(f[4 * c + 2].x, f[4 * c + 2].y); o.w = pack_bf16x2(f[4 * c + 3].x, f[4 * c + 3].y);
              *reinterpret_cast<uint4*>(sb + sw128_offset(r, g * 4 + c)) = o;
            }
          } else {
            // V^T: element (dh d, point r) of 2 atoms x [64 dh][64 points]
            uint8_t* col = sb + (r >> 6) * 8192 + (r & 7) * 2;
            const uint32_t pc = (r & 63) >> 3;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int d = g * 32 + 2 * i;
              *reinterpret_cast<__nv_bfloat16*>(col + d * 128 + ((pc ^ (d & 7)) << 4)) = __float2bfloat16_rn(f[i].x);
              *reinterpret_cast<__nv_bfloat16*>(col + (d + 1) * 128 + ((pc ^ ((d + 1) & 7)) << 4)) =
                  __float2bfloat16_rn(f[i].y);
            }
          }
          fence_proxy_async_smem();
          epi_bar_sync();
          if (leader) {
            int side;
            const size_t off = attn_tile_offset(p, rt, h, side);
            uint8_t* dst = (kind == 0 ? p.q_tiles[side] : kind == 1 ? p.k_tiles[side] : p.v_tiles[side]) + off;
            bulk_s2g(dst, sb, kHalfAtom);
            bulk_commit();
          }
        }
      }
    }
    if (leader) bulk_wait_all();
  }
  tc_fence_before();
  if (kCtas == 2) {
    cluster_sync_all();  // neither CTA leaves (or frees TMEM) while the pair's MMAs / remote arrivals are in flight
    if (warp == 9) tmem_dealloc_pair<512>(tmem);
  } else {
    __syncthreads();
    if (warp == 9) tmem_dealloc<512>(tmem);
  }
}

}  // namespace lg

// ============================================================================ format kernels
namespace lg {

// W (n_out, k) f32 row-major -> [n_out/nt][k/64][hi | lo][nt x 128 B] swizzled bf16 tiles.
// one thread per (row, 8-element chunk)
__global__ void pack_weight_kernel(const float* __restrict__ w, uint8_t* __restrict__ out, int n_out, int k, int nt) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int chunks = k / 8;
  if (idx >= n_out * chunks) return;
  const int n = idx / chunks, ch = idx - n * chunks;
  const float4 a = __ldg(reinterpret_cast<const float4*>(w + static_cast<size_t>(n) * k + ch * 8));
  const float4 b = __ldg(reinterpret_cast<const float4*>(w + static_cast<size_t>(n) * k + ch * 8) + 1);
  const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  uint4 hi, lo;
  split8(x, hi, lo);
  const int tile = n / nt, r = n - tile * nt, ka = ch >> 3, c = ch & 7;
  uint8_t* at = out + (static_cast<size_t>(tile) * (k / 64) + ka) * (static_cast<size_t>(nt) * 256) + sw128_offset(r, c);
  *reinterpret_cast<uint4*>(at) = hi;
  *reinterpret_cast<uint4*>(at + static_cast<size_t>(nt) * 128) = lo;
}

// x (B, N, K) f32 -> split tiles [B*tp][K/64][32 KB]; rows >= N of each image are zero.
// grid (tp, B), block 1024: thread -> (row, chunk) tasks
__global__ void __launch_bounds__(1024) rows_to_tiles_kernel(const float* __restrict__ x, uint8_t* __restrict__ tiles,
                                                             int N, int K, int tp) {
  const int t = blockIdx.x, b = blockIdx.y, chunks = K / 8, atoms = K / 64;
  uint8_t* dst = tiles + (static_cast<size_t>(b) * tp + t) * atoms * kSplitAtom;
  for (int task = threadIdx.x; task < kTile * chunks; task += blockDim.x) {
    const int r = task / chunks, ch = task - r * chunks, n = t * kTile + r;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (n < N) {
      const float4* src = reinterpret_cast<const float4*>(x + (static_cast<size_t>(b) * N + n) * K + ch * 8);
      const float4 a = __ldg(src), c = __ldg(src + 1);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
    }
    uint4 hi, lo;
    split8(v, hi, lo);
    uint8_t* at = dst + (ch >> 3) * kSplitAtom + sw128_offset(r, ch & 7);
    *reinterpret_cast<uint4*>(at) = hi;
    *reinterpret_cast<uint4*>(at + kHalfAtom) = lo;
  }
}

// inverse: x (B, N, K) f32 = hi + lo
__global__ void __launch_bounds__(1024) tiles_to_rows_kernel(const uint8_t* __restrict__ tiles, float* __restrict__ x,
                                                             int N, int K, int tp) {
  const int t = blockIdx.x, b = blockIdx.y, chunks = K / 8, atoms = K / 64;
  const uint8_t* src = tiles + (static_cast<size_t>(b) * tp + t) * atoms * kSplitAtom;
  for (int task = threadIdx.x; task < kTile * chunks; task += blockDim.x) {
    const int r = task / chunks, ch = task - r * chunks, n = t * kTile + r;
    if (n >= N) continue;
    const uint8_t* at = src + (ch >> 3) * kSplitAtom + sw128_offset(r, ch & 7);
    const uint4 h = *reinterpret_cast<const uint4*>(at), l = *reinterpret_cast<const uint4*>(at + kHalfAtom);
    float4 o0, o1;
    o0.x = bf16lo(h.x) + bf16lo(l.x); o0.y = bf16hi(h.x) + bf16hi(l.x);
    o0.z = bf16lo(h.y) + bf16lo(l.y); o0.w = bf16hi(h.y) + bf16hi(l.y);
    o1.x = bf16lo(h.z) + bf16lo(l.z); o1.y = bf16hi(h.z) + bf16hi(l.z);
    o1.z = bf16lo(h.w) + bf16lo(l.w); o1.w = bf16hi(h.w) + bf16hi(l.w);
    float4* dst = reinterpret_cast<float4*>(x + (static_cast<size_t>(b) * N + n) * K + ch * 8);
    dst[0] = o0;
    dst[1] = o1;
  }
}

// compact cos / sin tables (B, N, 32) f32 -> [B*tp][cos 16 KB | sin 16 KB], 128-byte rows whose
// 16-byte chunks are XOR-swizzled with (row & 7) (conflict-free thread-per-row reads)
__global__ void __launch_bounds__(1024) rotary_tiles_kernel(const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                                            uint8_t* __restrict__ rot, int N, int tp) {
  const int t = blockIdx.x, b = blockIdx.y;
  const int r = threadIdx.x >> 3, c = threadIdx.x & 7, n = t * kTile + r;
  float4 cv = make_float4(1.f, 1.f, 1.f, 1.f), sv = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n < N) {
    cv = __ldg(reinterpret_cast<const float4*>(cos_t + (static_cast<size_t>(b) * N + n) * 32) + c);
    sv = __ldg(reinterpret_cast<const float4*>(sin_t + (static_cast<size_t>(b) * N + n) * 32) + c);
  }
  uint8_t* dst = rot + (static_cast<size_t>(b) * tp + t) * 32768 + r * 128 + ((c ^ (r & 7)) << 4);
  *reinterpret_cast<float4*>(dst) = cv;
  *reinterpret_cast<float4*>(dst + kHalfAtom) = sv;
}

// h (rows, 512) f32 -> GELU(LayerNorm(h)) as split tiles [rows/128][8][32 KB]
// (lightglue.py:173-174: LayerNorm(2d, eps, affine) then exact-erf GELU).  Warp per row.
__global__ void __launch_bounds__(256) ln_gelu_tiles_kernel(const float* __restrict__ h, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps,
                                                            uint8_t* __restrict__ tiles, int rows) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float4* src = reinterpret_cast<const float4*>(h + static_cast<size_t>(row) * 512);
  float x[16];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float4 a = __ldg(src + j * 32 + lane);
    x[4 * j] = a.x; x[4 * j + 1] = a.y; x[4 * j + 2] = a.z; x[4 * j + 3] = a.w;
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  const float mean = warp_sum(s) * (1.f / 512.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { const float d = x[i] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / 512.f) + eps);
  const int rt = row >> 7, r = row & 127;
  uint8_t* tile = tiles + static_cast<size_t>(rt) * 8 * kSplitAtom;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int col = j * 128 + lane * 4;
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + col)), bt = __ldg(reinterpret_cast<const float4*>(beta + col));
    const float gg[4] = {g.x, g.y, g.z, g.w}, bb[4] = {bt.x, bt.y, bt.z, bt.w};
    float y[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float z = (x[4 * j + e] - mean) * rstd * gg[e] + bb[e];
      y[e] = 0.5f * z * (1.f + erff(z * 0.70710678118654752f));
    }
    const __nv_bfloat16 h0 = __float2bfloat16_rn(y[0]), h1 = __float2bfloat16_rn(y[1]), h2 = __float2bfloat16_rn(y[2]),
                        h3 = __float2bfloat16_rn(y[3]);
    uint2 hi, lo;
    hi.x = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
    hi.y = static_cast<uint32_t>(__bfloat16_as_ushort(h2)) | (static_cast<uint32_t>(__bfloat16_as_ushort(h3)) << 16);
    lo.x = pack_bf16x2(y[0] - __bfloat162float(h0), y[1] - __bfloat162float(h1));
    lo.y = pack_bf16x2(y[2] - __bfloat162float(h2), y[3] - __bfloat162float(h3));
    const int ka = col >> 6, cc = col & 63;
    uint8_t* at = tile + ka * kSplitAtom + sw128_offset(r, cc >> 3) + (cc & 7) * 2;
    *reinterpret_cast<uint2*>(at) = hi;
    *reinterpret_cast<uint2*>(at + kHalfAtom) = lo;
  }
}

static int sm_count() { return device_sm_count(); }

template <int kNt, int kEpi, int kCtas>
static int launch_linear_n(const LinearParams& p, cudaStream_t stream) {
  constexpr int smem = linear_smem_bytes<kNt, kEpi, kCtas>();
  static_assert(smem <= kLinSmemMax, "shared memory budget");
  static bool attr_done_dev[lg::kMaxDevices] = {};  // per device; a benign race sets it twice
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  if (!attr_done) {
    cudaError_t e =
        cudaFuncSetAttribute(linear_kernel<kNt, kEpi, kCtas>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  const int items = (p.row_tiles / kCtas) * p.n_tiles;
  const int slots = sm_count() / kCtas;  // persistent: one CTA (or CTA pair) per SM (or TPC)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>((items < slots ? items : slots) * kCtas));
  cfg.blockDim = dim3(kLinThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = add_pdl_attr(attr, 1);
  return static_cast<int>(cudaLaunchKernelEx(&cfg, linear_kernel<kNt, kEpi, kCtas>, p));
}

// LG_QKV_TERMS=1 in the environment: Wqkv / to_qk|to_v as single hi.hi products (experiment)
static int qkv_one_term() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("LG_QKV_TERMS");
    v = (e && e[0] == '1') ? 1 : 0;
  }
  return v;
}

// LG_LINEAR_PAIR=0 in the environment forces the single-CTA kernels (A/B measurements)
static bool pair_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("LG_LINEAR_PAIR");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

template <int kNt, int kEpi>
static int launch_linear(const LinearParams& p, cudaStream_t stream) {
  if ((p.row_tiles & 1) == 0 && pair_enabled()) return launch_linear_n<kNt, kEpi, 2>(p, stream);
  return launch_linear_n<kNt, kEpi, 1>(p, stream);
}

}  // namespace lg

// ============================================================================ C ABI
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int lg_pack_weight(const float* w, void* out, int n_out, int k, int nt, cudaStream_t stream) {
  if (!w || !out) return LG_ERR_NULL;
  if (n_out <= 0 || k <= 0 || k % 64 || (nt != 192 && nt != 256) || n_out % nt) return LG_ERR_SHAPE;
  if (!aligned16(w) || !aligned16(out)) return LG_ERR_ALIGN;
  const int total = n_out * (k / 8);
  lg::pack_weight_kernel<<<(total + 255) / 256, 256, 0, stream>>>(w, static_cast<uint8_t*>(out), n_out, k, nt);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_rows_to_tiles(const float* x, void* tiles, int B, int N, int K, int tp, cudaStream_t stream) {
  if (!x || !tiles) return LG_ERR_NULL;
  if (B <= 0 || N <= 0 || K <= 0 || K % 64 || tp * lg::kTile < N) return LG_ERR_SHAPE;
  if (!aligned16(x) || !aligned16(tiles)) return LG_ERR_ALIGN;
  lg::rows_to_tiles_kernel<<<dim3(tp, B), 1024, 0, stream>>>(x, static_cast<uint8_t*>(tiles), N, K, tp);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_tiles_to_rows(const void* tiles, float* x, int B, int N, int K, int tp, cudaStream_t stream) {
  if (!x || !tiles) return LG_ERR_NULL;
  if (B <= 0 || N <= 0 || K <= 0 || K % 64 || tp * lg::kTile < N) return LG_ERR_SHAPE;
  if (!aligned16(x) || !aligned16(tiles)) return LG_ERR_ALIGN;
  lg::tiles_to_rows_kernel<<<dim3(tp, B), 1024, 0, stream>>>(static_cast<const uint8_t*>(tiles), x, N, K, tp);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_rotary_tiles(const float* cos_t, const float* sin_t, void* rot_tiles, int B, int N, int tp,
                               cudaStream_t stream) {
  if (!cos_t || !sin_t || !rot_tiles) return LG_ERR_NULL;
  if (B <= 0 || N <= 0 || tp * lg::kTile < N) return LG_ERR_SHAPE;
  lg::rotary_tiles_kernel<<<dim3(tp, B), 1024, 0, stream>>>(cos_t, sin_t, static_cast<uint8_t*>(rot_tiles), N, tp);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_ln_gelu_tiles(const float* h, const float* gamma, const float* beta, float eps, void* tiles, int rows,
                                cudaStream_t stream) {
  if (!h || !gamma || !beta || !tiles) return LG_ERR_NULL;
  if (rows <= 0 || rows % lg::kTile) return LG_ERR_SHAPE;
  lg::ln_gelu_tiles_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(h, gamma, beta, eps, static_cast<uint8_t*>(tiles), rows);
  return static_cast<int>(cudaGetLastError());
}

static int check_linear(const void* a0, int a0_atoms, const void* a1, int a1_atoms, const void* w, const float* bias,
                        int row_tiles, int n_out, int nt) {
  if (!a0 || !w || !bias) return LG_ERR_NULL;
  if ((a1 == nullptr) != (a1_atoms == 0)) return LG_ERR_NULL;
  if (a0_atoms <= 0 || a1_atoms < 0 || row_tiles <= 0 || n_out <= 0 || n_out % nt) return LG_ERR_SHAPE;
  if (!aligned16(a0) || !aligned16(a1) || !aligned16(w) || !aligned16(bias)) return LG_ERR_ALIGN;
  return 0;
}

extern "C" int lg_linear_split(const void* a0, int a0_atoms, const void* a1, int a1_atoms, int a1_hi_only, const void* w,
                               const float* bias, const void* resid_tiles, void* out_tiles, int row_tiles, int n_out,
                               cudaStream_t stream) {
  if (int e = check_linear(a0, a0_atoms, a1, a1_atoms, w, bias, row_tiles, n_out, 256)) return e;
  if (!out_tiles) return LG_ERR_NULL;
  lg::LinearParams p{};
  p.a0 = static_cast<const uint8_t*>(a0); p.a1 = static_cast<const uint8_t*>(a1);
  p.a0_atoms = a0_atoms; p.a1_atoms = a1_atoms; p.a1_hi_only = a1_hi_only;
  p.w = static_cast<const uint8_t*>(w); p.bias = bias; p.row_tiles = row_tiles; p.n_tiles = n_out / 256;
  p.out_tiles = static_cast<uint8_t*>(out_tiles); p.resid_tiles = static_cast<const uint8_t*>(resid_tiles);
  return lg::launch_linear<256, lg::kEpiSplit>(p, stream);
}

extern "C" int lg_linear_f16(const void* a0, int a0_atoms, const void* a1, int a1_atoms, int a1_hi_only, const void* w,
                             const float* bias, void* out_tiles, int row_tiles, int n_out, cudaStream_t stream) {
  if (int e = check_linear(a0, a0_atoms, a1, a1_atoms, w, bias, row_tiles, n_out, 256)) return e;
  if (!out_tiles) return LG_ERR_NULL;
  lg::LinearParams p{};
  p.a0 = static_cast<const uint8_t*>(a0); p.a1 = static_cast<const uint8_t*>(a1);
  p.a0_atoms = a0_atoms; p.a1_atoms = a1_atoms; p.a1_hi_only = a1_hi_only;
  p.w = static_cast<const uint8_t*>(w); p.bias = bias; p.row_tiles = row_tiles; p.n_tiles = n_out / 256;
  p.out_tiles = static_cast<uint8_t*>(out_tiles);
  return lg::launch_linear<256, lg::kEpiF16>(p, stream);
}

extern "C" int lg_linear_f32(const void* a0, int a0_atoms, const void* a1, int a1_atoms, int a1_hi_only, const void* w,
                             const float* bias, float* out, int64_t ld, int row_tiles, int n_out, cudaStream_t stream) {
  if (int e = check_linear(a0, a0_atoms, a1, a1_atoms, w, bias, row_tiles, n_out, 256)) return e;
  if (!out) return LG_ERR_NULL;
  if (ld < n_out || ld % 4 || !aligned16(out)) return LG_ERR_ALIGN;
  lg::LinearParams p{};
  p.a0 = static_cast<const uint8_t*>(a0); p.a1 = static_cast<const uint8_t*>(a1);
  p.a0_atoms = a0_atoms; p.a1_atoms = a1_atoms; p.a1_hi_only = a1_hi_only;
  p.w = static_cast<const uint8_t*>(w); p.bias = bias; p.row_tiles = row_tiles; p.n_tiles = n_out / 256;
  p.out_f32 = out; p.ld = ld;
  return lg::launch_linear<256, lg::kEpiF32>(p, stream);
}

extern "C" int lg_linear_qkv_self(const void* a, const void* w, const float* bias, const void* rot_tiles, void* q0,
                                  void* k0, void* v0, void* q1, void* k1, void* v1, int B, int H, int tp0, int tp1,
                                  cudaStream_t stream) {
  if (!a || !w || !bias || !rot_tiles || !q0 || !k0 || !v0) return LG_ERR_NULL;
  if (tp1 > 0 && (!q1 || !k1 || !v1)) return LG_ERR_NULL;
  if (B <= 0 || H != 4 || tp0 <= 0 || tp1 < 0) return LG_ERR_SHAPE;
  lg::LinearParams p{};
  p.a0 = static_cast<const uint8_t*>(a); p.a0_atoms = 4;
  p.w = static_cast<const uint8_t*>(w); p.bias = bias; p.row_tiles = B * (tp0 + tp1); p.n_tiles = H;
  p.q_tiles[0] = static_cast<uint8_t*>(q0); p.k_tiles[0] = static_cast<uint8_t*>(k0); p.v_tiles[0] = static_cast<uint8_t*>(v0);
  p.q_tiles[1] = static_cast<uint8_t*>(q1); p.k_tiles[1] = static_cast<uint8_t*>(k1); p.v_tiles[1] = static_cast<uint8_t*>(v1);
  p.rot_tiles = static_cast<const uint8_t*>(rot_tiles);
  p.B = B; p.H = H; p.tp0 = tp0; p.tp1 = tp1;
  p.one_term = lg::qkv_one_term();
  return lg::launch_linear<192, lg::kEpiQkvSelf>(p, stream);
}

extern "C" int lg_linear_qkv_cross(const void* a, const void* w, const float* bias, void* qk0, void* v0, void* qk1,
                                   void* v1, int B, int H, int tp0, int tp1, cudaStream_t stream) {
  if (!a || !w || !bias || !qk0 || !v0) return LG_ERR_NULL;
  if (tp1 > 0 && (!qk1 || !v1)) return LG_ERR_NULL;
  if (B <= 0 || H != 4 || tp0 <= 0 || tp1 < 0) return LG_ERR_SHAPE;
  lg::LinearParams p{};
  p.a0 = static_cast<const uint8_t*>(a); p.a0_atoms = 4;
  p.w = static_cast<const uint8_t*>(w); p.bias = bias; p.row_tiles = B * (tp0 + tp1); p.n_tiles = 2;
  p.q_tiles[0] = static_cast<uint8_t*>(qk0); p.v_tiles[0] = static_cast<uint8_t*>(v0);
  p.q_tiles[1] = static_cast<uint8_t*>(qk1); p.v_tiles[1] = static_cast<uint8_t*>(v1);
  p.B = B; p.H = H; p.tp0 = tp0; p.tp1 = tp1;
  p.one_term = lg::qkv_one_term();
  return lg::launch_linear<256, lg::kEpiQkvCross>(p, stream);
}
