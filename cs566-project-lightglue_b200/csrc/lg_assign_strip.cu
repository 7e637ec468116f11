// lg_assign_strip : MatchAssignment similarity + double log-softmax straight from split-tile operands.
//
// Replaces MatchAssignment.forward from the einsum on (reference
// gluefactory/models/matchers/lightglue.py:311-315) and sigmoid_log_double_softmax (:284-296), like
// lg_assign.cu, but consumes the split-bf16 row tiles that the final_proj linear writes (no fp32 rows, no pack
// pass) and is organised around the L2->SM bytes the TF32x3 tile kernel was bound by (512 KB per 128x128 tile):
//   * work item = (pair, row tile of image 0, chunk of column tiles): the A operand (128 rows x 256, hi|lo bf16,
//     128 KB) stays RESIDENT in shared memory for the whole chunk, only the B tiles (128 KB each) stream through
//     a ring of 16 KB half-atoms -> 128..144 KB per tile instead of 512 KB;
//   * 3-term split-bf16 products on tcgen05 kind::f16 (lo.hi, hi.hi, hi.lo per K16 step) into ONE fp32 accumulator:
//     the split error (~1e-5 on |sim| ~ 10) dominates the accumulator's truncation over the 48-step chain;
//   * four TMEM accumulators, each owned by one of four epilogue warpgroups (16 warps: the epilogue is a long
//     dependent instruction stream, it needs the warps to fill the issue slots): the tensor core runs up to three
//     tiles ahead of the reductions / stores; a thread owns a ROW (its TMEM lane), pulls 16 columns at a time, and
//     hands the 32x16 block to its own warp through a private shared-memory transpose so that the column statistics
//     (pass 1, per-warp partials straight to global: no block barrier) and the global stores (pass 2: 64-byte row
//     segments) run with lane == column; the chunk loop is NOT unrolled (an unrolled 128-column epilogue is ~7000
//     instructions and thrashes the instruction cache);
//   * the next item's A atoms stream in behind the last tile of the current item (per-atom full / empty barriers);
//   * pass 1 = row / column (max, sum exp) partials, reduce (lg_assign.cu), pass 2 = recompute + write, and --
//     with keys -- the row / column argmax of filter_matches (lightglue.py:321-337) as packed 64-bit keys.
#include "lg_assign.cuh"
#include <cstdlib>
#include "lg_common.cuh"
#include "lg_keys.cuh"
#include "lg_tiles.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kStripAtoms = 4;                       // K = 256 = 4 atoms of 64 bf16
constexpr int kStripATile = kStripAtoms * kSplitAtom;  // 128 KB: one row tile, hi | lo per atom
constexpr int kStripStages = 3;                      // ring of 16 KB half-atoms of B
constexpr int kWGs = 4;                              // epilogue warpgroups == TMEM accumulators (128 columns each)
constexpr int kEpiWarps = 4 * kWGs;
constexpr int kStripThreads = 32 * (kEpiWarps + 2);  // + producer warp + MMA warp
constexpr int kChunk = 16;                           // columns an epilogue thread pulls out of TMEM per step
constexpr int kChunkPitch = kChunk + 1;
constexpr int kWarpStage = 32 * kChunkPitch * 4;     // private 32 x 16 fp32 transpose buffer of a warp (pitch 17)
constexpr uint32_t kNone = 0xFFFFFFFFu;

struct StripParams {
  const uint8_t* a_tiles;   // split tiles of final_proj(desc0): image b starts at row tile b * tpa
  const uint8_t* b_tiles;   // split tiles of final_proj(desc1): image b starts at row tile b * tpb
  int tpa, tpb;
  const int* m_len;
  const int* n_len;
  float2* row_part;         // pass 1: [B][ct][M] (max, sum)
  float2* col_part;         // pass 1: [B][rt][4 warps][N]
  const float* row_lse;     // pass 2: [B][M]
  const float* col_lse;     // pass 2: [B][N]
  const float* row_ls;      // pass 2: logsigmoid(z0) [B][M]
  const float* col_ls;      // pass 2: logsigmoid(z1) [B][N]
  float* scores;            // pass 2: (B, M+1, N+1)
  unsigned long long* rowkey;  // pass 2, optional
  unsigned long long* colkey;
  float scale;              // similarity = scale * <a, b>  (prescale^2 of the reference's d^-1/4 on both sides)
  int B, M, N, rt, ct, cw, chunks;  // tile counts; cw = column tiles per item, chunks = ceil(ct / cw)
};

struct __align__(8) StripBarriers {
  uint64_t a_full[kStripAtoms], a_empty[kStripAtoms];  // per K atom: the next item's A streams in behind the last tile
  uint64_t full[kStripStages], empty[kStripStages];
  uint64_t acc_full[kWGs], acc_empty[kWGs];
  uint32_t tmem_base;
};

constexpr int kStripScratch = kWGs * kTile * 12;  // pass 2: column terms [4][128] f32 + column keys [4][128] u64
constexpr int kStripSmem = 1024 + kStripATile + kStripStages * kHalfAtom + kEpiWarps * kWarpStage + kStripScratch +
                           static_cast<int>(sizeof(StripBarriers));

struct StripItem {
  int b, rtile, c0, c1, row0, vrows, n, nreal;  // column tiles [c0, c1); the first nreal of them hold valid columns
};

__device__ __forceinline__ StripItem strip_item(const StripParams& p, int it) {
  StripItem t;
  const int chunk = it % p.chunks, br = it / p.chunks;
  t.rtile = br % p.rt;
  t.b = br / p.rt;
  const int m = p.m_len ? min(p.m_len[t.b], p.M) : p.M;
  t.n = p.n_len ? min(p.n_len[t.b], p.N) : p.N;
  t.row0 = t.rtile * kTile;
  t.vrows = max(0, min(kTile, m - t.row0));
  t.c0 = chunk * p.cw;
  t.c1 = min(p.ct, t.c0 + p.cw);
  const int valid_ct = (t.n + kTile - 1) / kTile;
  t.nreal = t.vrows > 0 ? max(0, min(t.c1, valid_ct) - t.c0) : 0;
  return t;
}

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr float kLog2e = 1.4426950408889634f;

template <int kPass>
__global__ void __launch_bounds__(kStripThreads, 1) assign_strip_kernel(const StripParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // 1 KB aligned, still provably shared
  uint8_t* a_res = base;
  uint8_t* ring = base + kStripATile;
  float* wstage = reinterpret_cast<float*>(ring + kStripStages * kHalfAtom);
  uint8_t* scratch = reinterpret_cast<uint8_t*>(wstage) + kEpiWarps * kWarpStage;
  StripBarriers* bar = reinterpret_cast<StripBarriers*>(scratch + kStripScratch);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = p.B * p.rt * p.chunks;
  const int64_t pitch = p.N + 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStripAtoms; ++s) { mbar_init(&bar->a_full[s], 1); mbar_init(&bar->a_empty[s], 1); }
    for (int s = 0; s < kStripStages; ++s) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
    for (int s = 0; s < kWGs; ++s) { mbar_init(&bar->acc_full[s], 1); mbar_init(&bar->acc_empty[s], kTile); }
    fence_mbar_init();
  }
  if (warp == kEpiWarps + 1) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp == kEpiWarps) {
    // ------------------------------------------------------------------ producer: A atoms once per item, B half-atoms
    if (lane == 0) {
      uint32_t g = 0, loaded = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const StripItem t = strip_item(p, it);
        if (t.nreal == 0) continue;
        const uint8_t* ga = p.a_tiles + (static_cast<size_t>(t.b) * p.tpa + t.rtile) * kStripATile;
        for (int j = 0; j < t.nreal; ++j) {
          const uint8_t* gb = p.b_tiles + (static_cast<size_t>(t.b) * p.tpb + t.c0 + j) * kStripATile;
          for (int ka = 0; ka < kStripAtoms; ++ka) {
            if (j == 0) {  // atom ka of the new row tile, as soon as the previous item's last tile is past it
              if (loaded > 0) mbar_wait(&bar->a_empty[ka], (loaded - 1) & 1);
              mbar_arrive_expect_tx(&bar->a_full[ka], kSplitAtom);
              bulk_g2s(a_res + ka * kSplitAtom, ga + ka * kSplitAtom, kSplitAtom, &bar->a_full[ka]);
            }
            for (int h = 0; h < 2; ++h, ++g) {  // hi, lo of atom ka are contiguous in the tile
              const int s = g % kStripStages;
              if (g >= kStripStages) mbar_wait(&bar->empty[s], ((g / kStripStages) - 1) & 1);
              mbar_arrive_expect_tx(&bar->full[s], kHalfAtom);
              bulk_g2s(ring + s * kHalfAtom, gb + (2 * ka + h) * kHalfAtom, kHalfAtom, &bar->full[s]);
            }
          }
        }
        ++loaded;
      }
    }
  } else if (warp == kEpiWarps + 1) {
    // ------------------------------------------------------------------ tensor core issue (warp-uniform, one elected lane)
    constexpr uint32_t idesc = umma_idesc(kFmtBF16, 128, 128);
    const uint32_t a_addr = smem_u32(a_res), ring_addr = smem_u32(ring);
    const bool leader = elect_one();
    uint32_t g = 0, tl = 0, loaded = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const StripItem t = strip_item(p, it);
      if (t.nreal == 0) continue;
      for (int j = 0; j < t.nreal; ++j, ++tl) {
        const uint32_t set = tl % kWGs;  // accumulator == epilogue warpgroup
        if (tl >= kWGs) {  // the epilogue has pulled the tile that used this accumulator out of TMEM
          mbar_wait(&bar->acc_empty[set], ((tl / kWGs) - 1) & 1);
          tc_fence_after();
        }
        const uint32_t acc = tmem + set * kTile;
        for (int ka = 0; ka < kStripAtoms; ++ka) {
          if (j == 0) {
            mbar_wait(&bar->a_full[ka], loaded & 1);
            tc_fence_after();
          }
          const uint64_t dah = umma_desc_kmajor_sw128(a_addr + ka * kSplitAtom);
          const uint64_t dal = umma_desc_kmajor_sw128(a_addr + ka * kSplitAtom + kHalfAtom);
          {  // b_hi: lo.hi and hi.hi
            const int s = g % kStripStages;
            mbar_wait(&bar->full[s], (g / kStripStages) & 1);
            tc_fence_after();
            const uint64_t dbh = umma_desc_kmajor_sw128(ring_addr + s * kHalfAtom);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                umma_bf16_ss(acc, dal + 2 * kk, dbh + 2 * kk, idesc, (ka > 0 || kk > 0));
                umma_bf16_ss(acc, dah + 2 * kk, dbh + 2 * kk, idesc, 1);
              }
              umma_commit(&bar->empty[s]);
            }
            __syncwarp();
            ++g;
          }
          {  // b_lo: hi.lo
            const int s = g % kStripStages;
            mbar_wait(&bar->full[s], (g / kStripStages) & 1);
            tc_fence_after();
            const uint64_t dbl = umma_desc_kmajor_sw128(ring_addr + s * kHalfAtom);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(acc, dah + 2 * kk, dbl + 2 * kk, idesc, 1);
              umma_commit(&bar->empty[s]);
              if (j == t.nreal - 1) umma_commit(&bar->a_empty[ka]);  // this item is done with atom ka of A
            }
            __syncwarp();
            ++g;
          }
        }
        if (leader) umma_commit(&bar->acc_full[set]);
        __syncwarp();
      }
      ++loaded;
    }
  } else {
    // ------------------------------------------------------------------ epilogue: 4 warpgroups, thread == row (TMEM lane)
    const int wg = warp >> 2, wq = warp & 3;
    const int r = wq * 32 + lane;
    const int half = lane >> 4, cl = lane & 15;
    const uint32_t bar_id = 1 + wg;
    float* stg = wstage + warp * (kWarpStage / 4);
    const uint32_t t_acc = tmem + (static_cast<uint32_t>(wq * 32) << 16) + wg * kTile;  // warpgroup == accumulator
    float* s_colc = reinterpret_cast<float*>(scratch) + wg * kTile;                                       // pass 2 [128]
    unsigned long long* s_colkey = reinterpret_cast<unsigned long long*>(scratch + kWGs * kTile * 4) + wg * kTile;
    const bool keys = kPass == 2 && p.rowkey != nullptr;
    const float s2 = 2.f * p.scale;
    uint32_t tl = 0, pad = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const StripItem t = strip_item(p, it);
      const int b = t.b, row0 = t.row0, vrows = t.vrows;
      const bool row_ok = r < vrows;
      float rowc = -INFINITY;  // log-sigmoid term - logsumexp of this row, once per item
      if (kPass == 2 && t.nreal > 0 && row_ok)
        rowc = p.row_ls[static_cast<size_t>(b) * p.M + row0 + r] - p.row_lse[static_cast<size_t>(b) * p.M + row0 + r];
      for (int ctile = t.c0; ctile < t.c1; ++ctile) {
        const int col0 = ctile * kTile;
        float2* cpart = p.col_part + (static_cast<size_t>(b) * p.rt + t.rtile) * 4 * p.N + col0;  // [4 warps][N]
        if (ctile - t.c0 >= t.nreal) {
          // tile entirely in the padding of a variable-count batch: no MMA, the warpgroups take turns filling it
          if ((pad++ % kWGs) != static_cast<uint32_t>(wg)) continue;
          if (kPass == 1) {
            if (row0 + r < p.M)
              p.row_part[(static_cast<size_t>(b) * p.ct + ctile) * p.M + row0 + r] = make_float2(-INFINITY, 0.f);
            if (col0 + r < p.N)
              for (int q = 0; q < 4; ++q) cpart[static_cast<size_t>(q) * p.N + r] = make_float2(-INFINITY, 0.f);
          } else {
            float* out = p.scores + static_cast<int64_t>(b) * (p.M + 1) * pitch;
            for (int i = wq; i < kTile; i += 4)
              for (int q = 0; q < 4; ++q) {
                const int rr = row0 + i, c = col0 + q * 32 + lane;
                if (rr < p.M && c < p.N) out[rr * pitch + c] = -INFINITY;
              }
          }
          continue;
        }
        const uint32_t my = tl++;
        if ((my % kWGs) != static_cast<uint32_t>(wg)) continue;
        const int vcols = min(kTile, t.n - col0);
        const int nld = (vcols + kChunk - 1) / kChunk;  // chunks that hold similarities
        if (kPass == 2) {  // column terms of this tile, staged under the MMA
          s_colc[r] = r < vcols ? p.col_ls[static_cast<size_t>(b) * p.N + col0 + r] -
                                      p.col_lse[static_cast<size_t>(b) * p.N + col0 + r]
                                : -INFINITY;
          s_colkey[r] = 0ull;
        }
        mbar_wait(&bar->acc_full[wg], (my / kWGs) & 1);
        tc_fence_after();
#ifdef LG_STRIP_NOEPI  // experiment: MMA + operand ring alone
        tc_fence_before();
        mbar_arrive(&bar->acc_empty[wg]);
        continue;
#endif

        if (kPass == 1) {
          float rmax = -INFINITY, rsum = 0.f;
#pragma unroll 1
          for (int ch = 0; ch < nld; ++ch) {
            uint32_t v[kChunk];
            tmem_ld16(t_acc + ch * kChunk, v);
            tmem_wait_ld();
            if (ch == nld - 1) {  // the rest of the accumulator is padding: TMEM goes back to the tensor core
              tc_fence_before();
              mbar_arrive(&bar->acc_empty[wg]);
            }
            float s[kChunk], cmax = -INFINITY;
#pragma unroll
            for (int i = 0; i < kChunk; ++i) {
              s[i] = ch * kChunk + i < vcols ? __uint_as_float(v[i]) * p.scale : -INFINITY;
              cmax = fmaxf(cmax, s[i]);
            }
            if (cmax > rmax) { rsum *= ex2_fast((rmax - cmax) * kLog2e); rmax = cmax; }
            if (rmax > -INFINITY) {
              const float rm2 = rmax * kLog2e;
#pragma unroll
              for (int i = 0; i < kChunk; ++i) rsum += ex2_fast(fmaf(s[i], kLog2e, -rm2));
            }
            // column (max, sum exp) over this warp's 32 rows: private transpose, lane == (row parity, column)
#pragma unroll
            for (int i = 0; i < kChunk; ++i) stg[lane * kChunkPitch + i] = row_ok ? s[i] : -INFINITY;
            __syncwarp();
            float w[kChunk], m = -INFINITY;
#pragma unroll
            for (int j = 0; j < kChunk; ++j) {
              w[j] = stg[(2 * j + half) * kChunkPitch + cl];
              m = fmaxf(m, w[j]);
            }
            float sum = 0.f;
            if (m > -INFINITY) {
              const float m2 = m * kLog2e;
#pragma unroll
              for (int j = 0; j < kChunk; ++j) sum += ex2_fast(fmaf(w[j], kLog2e, -m2));
            }
            const float mo = __shfl_xor_sync(0xffffffffu, m, 16), so = __shfl_xor_sync(0xffffffffu, sum, 16);
            const float mm = fmaxf(m, mo);
            float tot = 0.f;
            if (sum > 0.f) tot += sum * ex2_fast((m - mm) * kLog2e);
            if (so > 0.f) tot += so * ex2_fast((mo - mm) * kLog2e);
            // per-warp partial straight to global (the reduce kernel merges 4 per row tile): no block barrier in pass 1
            if (half == 0 && col0 + ch * kChunk + cl < p.N)
              cpart[static_cast<size_t>(wq) * p.N + ch * kChunk + cl] = make_float2(mm, tot);
            __syncwarp();
          }
          for (int ch = nld; ch < kTile / kChunk; ++ch)  // columns of this tile beyond the pair's valid count
            if (half == 0 && col0 + ch * kChunk + cl < p.N)
              cpart[static_cast<size_t>(wq) * p.N + ch * kChunk + cl] = make_float2(-INFINITY, 0.f);
          if (row0 + r < p.M)
            p.row_part[(static_cast<size_t>(b) * p.ct + ctile) * p.M + row0 + r] =
                row_ok ? make_float2(rmax, rsum) : make_float2(-INFINITY, 0.f);
        } else {
          float* out = p.scores + static_cast<int64_t>(b) * (p.M + 1) * pitch;
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // column terms / key slots are in place
          const int nch = (min(kTile, p.N - col0) + kChunk - 1) / kChunk;  // chunks that are stored
          // every element of this warp's 32 x 128 block valid and inside the matrix: no masks, no store predicates
          const bool full = vcols == kTile && wq * 32 + 32 <= vrows;
          float rv = -INFINITY;
          uint32_t rc = kNone;
#pragma unroll 1
          for (int ch = 0; ch < nch; ++ch) {
            uint32_t v[kChunk];
            if (ch < nld) {
              tmem_ld16(t_acc + ch * kChunk, v);
              tmem_wait_ld();
              if (ch == nld - 1) {
                tc_fence_before();
                mbar_arrive(&bar->acc_empty[wg]);
              }
            }
            // thread == row: final values of 16 columns, best column of the row so far
            float cc[kChunk];
#pragma unroll
            for (int q = 0; q < kChunk / 4; ++q) {
              const float4 c4 = *reinterpret_cast<const float4*>(s_colc + ch * kChunk + 4 * q);
              cc[4 * q] = c4.x; cc[4 * q + 1] = c4.y; cc[4 * q + 2] = c4.z; cc[4 * q + 3] = c4.w;
            }
#pragma unroll
            for (int i = 0; i < kChunk; ++i) {
              const int c = ch * kChunk + i;
              // (s - lse_r) + (s - lse_c) + (ls_r + ls_c) as one fma: 2 s + ((ls_r - lse_r) + (ls_c - lse_c))
              float val = fmaf(__uint_as_float(v[i]), s2, rowc + cc[i]);
              if (!full) val = (row_ok && c < vcols) ? val : -INFINITY;  // padding may hold anything (NaN + -inf)
              if (keys) {  // strict > keeps the lowest column on ties
                const bool up = val > rv;
                rv = up ? val : rv;
                rc = up ? static_cast<uint32_t>(col0 + c) : rc;
              }
              stg[lane * kChunkPitch + i] = val;
            }
            __syncwarp();
            // lane == (row parity, column): 64-byte row segments, best row of the column among this warp's rows
            float cbv = -INFINITY;
            uint32_t cbi = kNone;
            const int c = ch * kChunk + cl;
            float* orow = out + static_cast<int64_t>(row0 + wq * 32 + half) * pitch + col0 + c;
            if (full) {
#pragma unroll
              for (int j = 0; j < kChunk; ++j) {
                const float x = stg[(2 * j + half) * kChunkPitch + cl];
#ifndef LG_STRIP_NOSTORE
                orow[2 * j * pitch] = x;
#endif
                if (keys) {
                  const bool up = x > cbv;
                  cbv = up ? x : cbv;
                  cbi = up ? static_cast<uint32_t>(row0 + wq * 32 + 2 * j + half) : cbi;
                }
              }
            } else {
              const bool col_st = col0 + c < p.N, col_ok = c < vcols;
#pragma unroll
              for (int j = 0; j < kChunk; ++j) {
                const float x = stg[(2 * j + half) * kChunkPitch + cl];
                const int rr = wq * 32 + 2 * j + half;
                if (col_st && row0 + rr < p.M) orow[2 * j * pitch] = x;
                if (keys && col_ok && rr < vrows && (x > cbv || cbi == kNone)) { cbv = x; cbi = row0 + rr; }
              }
            }
            if (keys) {
              if (full && cbi == kNone) cbi = row0 + wq * 32 + half;  // a column of -inf: its first row, like torch.max
              if (cbi != kNone) atomicMax(&s_colkey[c], make_key(cbv, cbi));
            }
            __syncwarp();
          }
          if (keys) {
            if (row_ok) {
              if (rc == kNone) rc = col0;  // a row of -inf: its first column
              atomicMax(p.rowkey + static_cast<size_t>(b) * p.M + row0 + r, make_key(rv, rc));
            }
            asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            if (r < vcols) atomicMax(p.colkey + static_cast<size_t>(b) * p.N + col0 + r, s_colkey[r]);
          }
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // terms / keys may be overwritten for the next tile
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps + 1) tmem_dealloc<512>(tmem);
}
}  // namespace lg

static inline size_t strip_align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

extern "C" size_t lg_assign_tiles_ws_bytes(int B, int M, int N) {
  if (B <= 0 || M <= 0 || N <= 0) return 0;
  const size_t rt = (M + 127) / 128, ct64 = (N + 63) / 64;
  size_t bytes = 0;
  bytes += strip_align256(B * ct64 * static_cast<size_t>(M) * sizeof(float2));  // row partials (64-col tiles of the TMEM variant)
  bytes += strip_align256(B * rt * 4 * static_cast<size_t>(N) * sizeof(float2));  // col partials, one per warp
  bytes += 2 * strip_align256(static_cast<size_t>(B) * M * 4) + 2 * strip_align256(static_cast<size_t>(B) * N * 4);
  return bytes;
}

extern "C" int lg_assign_tiles(const void* a_tiles, const void* b_tiles, int tpa, int tpb, const float* z0,
                               const float* z1, const int* m_len, const int* n_len, float scale, float* scores,
                               void* ws, void* keys_ws, int B, int M, int N, cudaStream_t stream) {
  if (!a_tiles || !b_tiles || !z0 || !z1 || !scores || !ws) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0) return LG_ERR_SHAPE;
  const int rt = (M + 127) / 128, ct = (N + 127) / 128;
  if (tpa < rt || tpb < ct) return LG_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(a_tiles) | reinterpret_cast<uintptr_t>(b_tiles)) & 15) return LG_ERR_ALIGN;
  // LG_ASSIGN_IMPL: "tmem" = both passes of the TMEM-resident variant, "strip" = both passes of the strip kernels;
  // default = strip pass 1 (128x128 SS tiles: the faster statistics pass) + TMEM pass 2 (the faster write pass)
  static const int use_tmem = [] {
    const char* e = getenv("LG_ASSIGN_IMPL");
    return !e ? 0 : e[0] == 't' ? 1 : e[0] == 's' ? 2 : 0;
  }();
  if (use_tmem == 1)
    return lg_assign_tiles_tmem(a_tiles, b_tiles, tpa, tpb, z0, z1, m_len, n_len, scale, scores, ws, keys_ws, B, M, N,
                                stream);
  uint8_t* w = static_cast<uint8_t*>(ws);
  float2* row_part = reinterpret_cast<float2*>(w); w += strip_align256(static_cast<size_t>(B) * ct * M * sizeof(float2));
  float2* col_part = reinterpret_cast<float2*>(w); w += strip_align256(static_cast<size_t>(B) * rt * 4 * N * sizeof(float2));
  float* row_lse = reinterpret_cast<float*>(w); w += strip_align256(static_cast<size_t>(B) * M * 4);
  float* row_ls = reinterpret_cast<float*>(w); w += strip_align256(static_cast<size_t>(B) * M * 4);
  float* col_lse = reinterpret_cast<float*>(w); w += strip_align256(static_cast<size_t>(B) * N * 4);
  float* col_ls = reinterpret_cast<float*>(w);

  unsigned long long* rowkey = nullptr;
  unsigned long long* colkey = nullptr;
  if (keys_ws) {
    cudaError_t e = cudaMemsetAsync(keys_ws, 0, lg_filter_ws_bytes(B, M, N), stream);
    if (e != cudaSuccess) return static_cast<int>(e);
    rowkey = static_cast<unsigned long long*>(keys_ws);
    colkey = rowkey + static_cast<size_t>(B) * M;
  }
  static bool attr_done_dev[lg::kMaxDevices] = {};  // per device; a benign race sets it twice
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(lg::assign_strip_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         lg::kStripSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(lg::assign_strip_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, lg::kStripSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  // column tiles per item: as many as keep >= 4 items per SM in flight (A is reloaded once per item)
  int cw = 8;
  while (cw > 1 && static_cast<long long>(B) * rt * ((ct + cw - 1) / cw) < 4LL * sms) cw >>= 1;
  const int chunks = (ct + cw - 1) / cw;
  const long long items = static_cast<long long>(B) * rt * chunks;
  if (items > 0x7fffffffLL) return LG_ERR_SHAPE;
  const int grid = items < sms ? static_cast<int>(items) : sms;  // persistent: one CTA per SM
  lg::StripParams p{static_cast<const uint8_t*>(a_tiles), static_cast<const uint8_t*>(b_tiles), tpa, tpb, m_len, n_len,
                    row_part, col_part, row_lse, col_lse, row_ls, col_ls, scores, rowkey, colkey, scale,
                    B, M, N, rt, ct, cw, chunks};
  lg::assign_strip_kernel<1><<<grid, lg::kStripThreads, lg::kStripSmem, stream>>>(p);
  lg::launch_assign_reduce(row_part, col_part, z0, z1, m_len, n_len, row_lse, col_lse, row_ls, col_ls, scores, B, M, N,
                           4 * rt, ct, stream);  // 4 column partials per row tile (one per epilogue warp)
  if (use_tmem == 0)
    return lg::launch_assign_tmem_pass2(a_tiles, b_tiles, tpa, tpb, m_len, n_len, row_lse, col_lse, row_ls, col_ls, scores,
                                        rowkey, colkey, scale, B, M, N, stream);
  lg::assign_strip_kernel<2><<<grid, lg::kStripThreads, lg::kStripSmem, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}
