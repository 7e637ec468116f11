// lg_assign_tmem : the strip assignment kernels (lg_assign_strip.cu) with the A operand RESIDENT IN TENSOR MEMORY.
//
// Same math, same interface (lg_assign_tiles), same two passes as lg_assign_strip.cu -- MatchAssignment.forward from the
// similarity on (reference gluefactory/models/matchers/lightglue.py:311-315) + sigmoid_log_double_softmax (:284-296) +
// the argmax keys of filter_matches (:321-337) -- but the strip kernel was bound by the latency of its 3-stage B ring
// (48 KB in flight per SM: the resident 128 KB A tile left no shared memory for more).  Here
//   * the row tile of image 0 (128 x 256, hi and lo bf16) lives in TMEM columns [0, 256): a loader warpgroup copies
//     each 32 KB atom from a 2-slot shared-memory staging buffer with tcgen05.st, per-atom full / empty barriers let
//     the next item's atoms stream in behind the last tile of the current item; the MMAs take A from TMEM
//     (tcgen05.mma kind::f16, A in tensor memory) and stream only B;
//   * tiles are 128 x 64: four 64-column fp32 accumulators in TMEM columns [256, 512), one per epilogue warpgroup;
//   * shared memory goes to a 14-stage ring of 8 KB B half-atoms (112 KB in flight).
// 22 warps: 0-15 epilogue (4 warpgroups), 16-19 A loader, 20 producer (one lane), 21 MMA issue (one elected lane).
#include "lg_assign.cuh"
#include "lg_common.cuh"
#include "lg_keys.cuh"
#include "lg_tiles.cuh"
#include "lg_b200.h"

namespace lg {
namespace tm {

constexpr int kAtoms = 4;                    // K = 256 = 4 atoms of 64 bf16
constexpr int kTileN = 64;                   // columns of a similarity tile
constexpr int kBStage = 8192;                // 64 rows x 128 B: half (hi or lo) of a K atom of a column tile
constexpr int kRing = 14;
constexpr int kWGs = 4, kEpiWarps = 16;
constexpr int kLoaderWarp0 = 16, kProducerWarp = 20, kMmaWarp = 21;
constexpr int kThreads = 32 * 22;
constexpr int kChunk = 16, kChunkPitch = kChunk + 1;
constexpr int kWarpStage = 32 * kChunkPitch * 4;
constexpr int kScratch = kWGs * kTileN * 12;  // pass 2: column terms [4][64] f32 + column keys [4][64] u64
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kColAhi = 0, kColAlo = 128, kColAcc = 256;
constexpr float kLog2e = 1.4426950408889634f;

struct Params {
  const uint8_t* a_tiles;
  const uint8_t* b_tiles;
  int tpa, tpb;
  const int* m_len;
  const int* n_len;
  float2* row_part;         // pass 1: [B][ct64][M] (max, sum)
  float2* col_part;         // pass 1: [B][rt][4 warps][N]
  const float* row_lse;
  const float* col_lse;
  const float* row_ls;
  const float* col_ls;
  float* scores;            // (B, M+1, N+1)
  unsigned long long* rowkey;
  unsigned long long* colkey;
  float scale;
  int B, M, N, rt, ct64, cw, chunks;  // ct64 = 64-column tiles per row; cw = tiles per item
};

struct __align__(8) Barriers {
  uint64_t as_full[2], as_empty[2];        // A staging slots: TMA landed / loader done reading
  uint64_t a_ready[kAtoms], a_empty[kAtoms];  // A atom in TMEM / the previous item's MMAs are past it
  uint64_t full[kRing], empty[kRing];
  uint64_t acc_full[kWGs], acc_empty[kWGs];
  uint32_t tmem_base;
};

constexpr int kSmem = 1024 + 2 * kSplitAtom + kRing * kBStage + kEpiWarps * kWarpStage + kScratch +
                      static_cast<int>(sizeof(Barriers));
static_assert(kSmem <= 232448, "shared memory budget");

struct Item {
  int b, rtile, c0, c1, row0, vrows, n, nreal;
};
__device__ __forceinline__ Item item_of(const Params& p, int it) {
  Item t;
  const int chunk = it % p.chunks, br = it / p.chunks;
  t.rtile = br % p.rt;
  t.b = br / p.rt;
  const int m = p.m_len ? min(p.m_len[t.b], p.M) : p.M;
  t.n = p.n_len ? min(p.n_len[t.b], p.N) : p.N;
  t.row0 = t.rtile * kTile;
  t.vrows = max(0, min(kTile, m - t.row0));
  t.c0 = chunk * p.cw;
  t.c1 = min(p.ct64, t.c0 + p.cw);
  const int valid_ct = (t.n + kTileN - 1) / kTileN;
  t.nreal = t.vrows > 0 ? max(0, min(t.c1, valid_ct) - t.c0) : 0;
  return t;
}

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// D[tmem] (+)= A[tmem] * B[smem]^T : A = bf16 rows packed two per 32-bit column (lane = row)
__device__ __forceinline__ void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

template <int kPass>
__global__ void __launch_bounds__(kThreads, 1) assign_tmem_kernel(const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // 1 KB aligned, still provably shared
  uint8_t* a_stage = base;                              // 2 x 32 KB
  uint8_t* ring = base + 2 * kSplitAtom;                // kRing x 8 KB
  float* wstage = reinterpret_cast<float*>(ring + kRing * kBStage);
  uint8_t* scratch = reinterpret_cast<uint8_t*>(wstage) + kEpiWarps * kWarpStage;
  Barriers* bar = reinterpret_cast<Barriers*>(scratch + kScratch);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = p.B * p.rt * p.chunks;
  const int64_t pitch = p.N + 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&bar->as_full[s], 1); mbar_init(&bar->as_empty[s], kTile); }
    for (int s = 0; s < kAtoms; ++s) { mbar_init(&bar->a_ready[s], kTile); mbar_init(&bar->a_empty[s], 1); }
    for (int s = 0; s < kRing; ++s) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
    for (int s = 0; s < kWGs; ++s) { mbar_init(&bar->acc_full[s], 1); mbar_init(&bar->acc_empty[s], kTile); }
    fence_mbar_init();
  }
  if (warp == kMmaWarp) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp == kProducerWarp) {
    // ------------------------------------------------------------------ producer: A atoms into the staging slots, B stages
    if (lane == 0) {
      uint32_t g = 0, na = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const Item t = item_of(p, it);
        if (t.nreal == 0) continue;
        const uint8_t* ga = p.a_tiles + (static_cast<size_t>(t.b) * p.tpa + t.rtile) * (kAtoms * kSplitAtom);
        for (int j = 0; j < t.nreal; ++j) {
          const int c = t.c0 + j;  // 64-column tile c = half (c & 1) of row tile c / 2 of image 1
          const uint8_t* gb = p.b_tiles + (static_cast<size_t>(t.b) * p.tpb + (c >> 1)) * (kAtoms * kSplitAtom) + (c & 1) * kBStage;
          for (int ka = 0; ka < kAtoms; ++ka) {
            if (j == 0) {
              const uint32_t slot = na & 1;
              if (na >= 2) mbar_wait(&bar->as_empty[slot], ((na >> 1) - 1) & 1);
              mbar_arrive_expect_tx(&bar->as_full[slot], kSplitAtom);
              bulk_g2s(a_stage + slot * kSplitAtom, ga + ka * kSplitAtom, kSplitAtom, &bar->as_full[slot]);
              ++na;
            }
            for (int h = 0; h < 2; ++h, ++g) {  // hi, lo: 64 rows x 128 B out of each 16 KB half-atom
              const int s = g % kRing;
              if (g >= kRing) mbar_wait(&bar->empty[s], ((g / kRing) - 1) & 1);
              mbar_arrive_expect_tx(&bar->full[s], kBStage);
              bulk_g2s(ring + s * kBStage, gb + ka * kSplitAtom + h * kHalfAtom, kBStage, &bar->full[s]);
            }
          }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ------------------------------------------------------------------ tensor core issue: A from TMEM, B from the ring
    constexpr uint32_t idesc = umma_idesc(kFmtBF16, 128, kTileN);
    const uint32_t ring_addr = smem_u32(ring);
    const bool leader = elect_one();
    uint32_t g = 0, tl = 0, done = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const Item t = item_of(p, it);
      if (t.nreal == 0) continue;
      for (int j = 0; j < t.nreal; ++j, ++tl) {
        const uint32_t set = tl % kWGs;
        if (tl >= kWGs) {
          mbar_wait(&bar->acc_empty[set], ((tl / kWGs) - 1) & 1);
          tc_fence_after();
        }
        const uint32_t acc = tmem + kColAcc + set * kTileN;
        for (int ka = 0; ka < kAtoms; ++ka) {
          if (j == 0) {
            mbar_wait(&bar->a_ready[ka], done & 1);
            tc_fence_after();
          }
          const uint32_t a_hi = tmem + kColAhi + ka * 32, a_lo = tmem + kColAlo + ka * 32;
          {  // b_hi: lo.hi and hi.hi
            const int s = g % kRing;
            mbar_wait(&bar->full[s], (g / kRing) & 1);
            tc_fence_after();
            const uint64_t dbh = umma_desc_kmajor_sw128(ring_addr + s * kBStage);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                umma_ts(acc, a_lo + kk * 8, dbh + 2 * kk, idesc, (ka > 0 || kk > 0));
                umma_ts(acc, a_hi + kk * 8, dbh + 2 * kk, idesc, 1);
              }
              umma_commit(&bar->empty[s]);
            }
            __syncwarp();
            ++g;
          }
          {  // b_lo: hi.lo
            const int s = g % kRing;
            mbar_wait(&bar->full[s], (g / kRing) & 1);
            tc_fence_after();
            const uint64_t dbl = umma_desc_kmajor_sw128(ring_addr + s * kBStage);
            if (leader) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) umma_ts(acc, a_hi + kk * 8, dbl + 2 * kk, idesc, 1);
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
      ++done;
    }
  } else if (warp >= kLoaderWarp0) {
    // ------------------------------------------------------------------ A loader: staging slot -> TMEM, thread == row
    const int r = (warp - kLoaderWarp0) * 32 + lane;
    const uint32_t t_lane = tmem + (static_cast<uint32_t>((warp - kLoaderWarp0) * 32) << 16);
    uint32_t na = 0, done = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const Item t = item_of(p, it);
      if (t.nreal == 0) continue;
      for (int ka = 0; ka < kAtoms; ++ka, ++na) {
        const uint32_t slot = na & 1;
        mbar_wait(&bar->as_full[slot], (na >> 1) & 1);
        if (done > 0) {  // the previous item's last tile is past atom ka
          mbar_wait(&bar->a_empty[ka], (done - 1) & 1);
          tc_fence_after();
        }
        const uint8_t* src = a_stage + slot * kSplitAtom;
        uint32_t w[32];
#pragma unroll
        for (int half = 0; half < 2; ++half) {  // hi block, lo block of the atom: 128 B of this row each
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint4 v = *reinterpret_cast<const uint4*>(src + half * kHalfAtom + sw128_offset(r, j));
            w[4 * j] = v.x; w[4 * j + 1] = v.y; w[4 * j + 2] = v.z; w[4 * j + 3] = v.w;
          }
          tmem_st32(t_lane + (half ? kColAlo : kColAhi) + ka * 32, w);
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&bar->a_ready[ka]);
        mbar_arrive(&bar->as_empty[slot]);
      }
      ++done;
    }
  } else {
    // ------------------------------------------------------------------ epilogue: 4 warpgroups, thread == row (TMEM lane)
    const int wg = warp >> 2, wq = warp & 3;
    const int r = wq * 32 + lane;
    const int half = lane >> 4, cl = lane & 15;
    const uint32_t bar_id = 1 + wg;
    float* stg = wstage + warp * (kWarpStage / 4);
    const uint32_t t_acc = tmem + (static_cast<uint32_t>(wq * 32) << 16) + kColAcc + wg * kTileN;
    float* s_colc = reinterpret_cast<float*>(scratch) + wg * kTileN;                                       // pass 2 [64]
    unsigned long long* s_colkey = reinterpret_cast<unsigned long long*>(scratch + kWGs * kTileN * 4) + wg * kTileN;
    const bool keys = kPass == 2 && p.rowkey != nullptr;
    const float s2 = 2.f * p.scale;
    uint32_t tl = 0, pad = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const Item t = item_of(p, it);
      const int b = t.b, row0 = t.row0, vrows = t.vrows;
      const bool row_ok = r < vrows;
      float rowc = -INFINITY;  // log-sigmoid term - logsumexp of this row, once per item
      if (kPass == 2 && t.nreal > 0 && row_ok)
        rowc = p.row_ls[static_cast<size_t>(b) * p.M + row0 + r] - p.row_lse[static_cast<size_t>(b) * p.M + row0 + r];
      for (int ctile = t.c0; ctile < t.c1; ++ctile) {
        const int col0 = ctile * kTileN;
        float2* cpart = p.col_part + (static_cast<size_t>(b) * p.rt + t.rtile) * 4 * p.N + col0;  // [4 warps][N]
        if (ctile - t.c0 >= t.nreal) {
          // tile entirely in the padding of a variable-count batch: no MMA, the warpgroups take turns filling it
          if ((pad++ % kWGs) != static_cast<uint32_t>(wg)) continue;
          if (kPass == 1) {
            if (row0 + r < p.M)
              p.row_part[(static_cast<size_t>(b) * p.ct64 + ctile) * p.M + row0 + r] = make_float2(-INFINITY, 0.f);
            if (r < kTileN && col0 + r < p.N)
              for (int q = 0; q < 4; ++q) cpart[static_cast<size_t>(q) * p.N + r] = make_float2(-INFINITY, 0.f);
          } else {
            float* out = p.scores + static_cast<int64_t>(b) * (p.M + 1) * pitch;
            for (int i = wq; i < kTile; i += 4)
              for (int q = 0; q < kTileN / 32; ++q) {
                const int rr = row0 + i, c = col0 + q * 32 + lane;
                if (rr < p.M && c < p.N) out[rr * pitch + c] = -INFINITY;
              }
          }
          continue;
        }
        const uint32_t my = tl++;
        if ((my % kWGs) != static_cast<uint32_t>(wg)) continue;
        const int vcols = min(kTileN, t.n - col0);
        const int nld = (vcols + kChunk - 1) / kChunk;  // chunks that hold similarities
        if (kPass == 2 && r < kTileN) {  // column terms of this tile, staged under the MMA
          s_colc[r] = r < vcols ? p.col_ls[static_cast<size_t>(b) * p.N + col0 + r] -
                                      p.col_lse[static_cast<size_t>(b) * p.N + col0 + r]
                                : -INFINITY;
          s_colkey[r] = 0ull;
        }
        mbar_wait(&bar->acc_full[wg], (my / kWGs) & 1);
        tc_fence_after();

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
            if (cmax > rmax) { rsum *= ex2f((rmax - cmax) * kLog2e); rmax = cmax; }
            if (rmax > -INFINITY) {
              const float rm2 = rmax * kLog2e;
#pragma unroll
              for (int i = 0; i < kChunk; ++i) rsum += ex2f(fmaf(s[i], kLog2e, -rm2));
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
              for (int j = 0; j < kChunk; ++j) sum += ex2f(fmaf(w[j], kLog2e, -m2));
            }
            const float mo = __shfl_xor_sync(0xffffffffu, m, 16), so = __shfl_xor_sync(0xffffffffu, sum, 16);
            const float mm = fmaxf(m, mo);
            float tot = 0.f;
            if (sum > 0.f) tot += sum * ex2f((m - mm) * kLog2e);
            if (so > 0.f) tot += so * ex2f((mo - mm) * kLog2e);
            if (half == 0 && col0 + ch * kChunk + cl < p.N)
              cpart[static_cast<size_t>(wq) * p.N + ch * kChunk + cl] = make_float2(mm, tot);
            __syncwarp();
          }
          for (int ch = nld; ch < kTileN / kChunk; ++ch)  // columns of this tile beyond the pair's valid count
            if (half == 0 && col0 + ch * kChunk + cl < p.N)
              cpart[static_cast<size_t>(wq) * p.N + ch * kChunk + cl] = make_float2(-INFINITY, 0.f);
          if (row0 + r < p.M)
            p.row_part[(static_cast<size_t>(b) * p.ct64 + ctile) * p.M + row0 + r] =
                row_ok ? make_float2(rmax, rsum) : make_float2(-INFINITY, 0.f);
        } else {
          float* out = p.scores + static_cast<int64_t>(b) * (p.M + 1) * pitch;
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // column terms / key slots are in place
          const int nch = (min(kTileN, p.N - col0) + kChunk - 1) / kChunk;  // chunks that are stored
          const bool full = vcols == kTileN && wq * 32 + 32 <= vrows;
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
            float cc[kChunk];
#pragma unroll
            for (int q = 0; q < kChunk / 4; ++q) {
              const float4 c4 = *reinterpret_cast<const float4*>(s_colc + ch * kChunk + 4 * q);
              cc[4 * q] = c4.x; cc[4 * q + 1] = c4.y; cc[4 * q + 2] = c4.z; cc[4 * q + 3] = c4.w;
            }
#pragma unroll
            for (int i = 0; i < kChunk; ++i) {
              const int c = ch * kChunk + i;
              float val = fmaf(__uint_as_float(v[i]), s2, rowc + cc[i]);
              if (!full) val = (row_ok && c < vcols) ? val : -INFINITY;
              if (keys) {
                const bool up = val > rv;
                rv = up ? val : rv;
                rc = up ? static_cast<uint32_t>(col0 + c) : rc;
              }
              stg[lane * kChunkPitch + i] = val;
            }
            __syncwarp();
            float cbv = -INFINITY;
            uint32_t cbi = kNone;
            const int c = ch * kChunk + cl;
            float* orow = out + static_cast<int64_t>(row0 + wq * 32 + half) * pitch + col0 + c;
            if (full) {
#pragma unroll
              for (int j = 0; j < kChunk; ++j) {
                const float x = stg[(2 * j + half) * kChunkPitch + cl];
                orow[2 * j * pitch] = x;
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
              if (full && cbi == kNone) cbi = row0 + wq * 32 + half;
              if (cbi != kNone) atomicMax(&s_colkey[c], make_key(cbv, cbi));
            }
            __syncwarp();
          }
          if (keys) {
            if (row_ok) {
              if (rc == kNone) rc = col0;
              atomicMax(p.rowkey + static_cast<size_t>(b) * p.M + row0 + r, make_key(rv, rc));
            }
            asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            if (r < vcols) atomicMax(p.colkey + static_cast<size_t>(b) * p.N + col0 + r, s_colkey[r]);
          }
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc<512>(tmem);
}

}  // namespace tm
}  // namespace lg

static inline size_t tm_align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// Same contract as lg_assign_tiles (include/lg_b200.h); ws must hold lg_assign_tiles_ws_bytes(B, M, N) bytes.
extern "C" int lg_assign_tiles_tmem(const void* a_tiles, const void* b_tiles, int tpa, int tpb, const float* z0,
                                    const float* z1, const int* m_len, const int* n_len, float scale, float* scores,
                                    void* ws, void* keys_ws, int B, int M, int N, cudaStream_t stream) {
  if (!a_tiles || !b_tiles || !z0 || !z1 || !scores || !ws) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0) return LG_ERR_SHAPE;
  const int rt = (M + 127) / 128, ct = (N + 127) / 128, ct64 = (N + 63) / 64;
  if (tpa < rt || tpb < ct) return LG_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(a_tiles) | reinterpret_cast<uintptr_t>(b_tiles)) & 15) return LG_ERR_ALIGN;
  uint8_t* w = static_cast<uint8_t*>(ws);
  float2* row_part = reinterpret_cast<float2*>(w); w += tm_align256(static_cast<size_t>(B) * ct64 * M * sizeof(float2));
  float2* col_part = reinterpret_cast<float2*>(w); w += tm_align256(static_cast<size_t>(B) * rt * 4 * N * sizeof(float2));
  float* row_lse = reinterpret_cast<float*>(w); w += tm_align256(static_cast<size_t>(B) * M * 4);
  float* row_ls = reinterpret_cast<float*>(w); w += tm_align256(static_cast<size_t>(B) * M * 4);
  float* col_lse = reinterpret_cast<float*>(w); w += tm_align256(static_cast<size_t>(B) * N * 4);
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
    cudaError_t e = cudaFuncSetAttribute(lg::tm::assign_tmem_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         lg::tm::kSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(lg::tm::assign_tmem_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, lg::tm::kSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  int cw = 16;  // 64-column tiles per item: as many as keep >= 4 items per SM (A is converted once per item)
  while (cw > 1 && static_cast<long long>(B) * rt * ((ct64 + cw - 1) / cw) < 4LL * sms) cw >>= 1;
  const int chunks = (ct64 + cw - 1) / cw;
  const long long items = static_cast<long long>(B) * rt * chunks;
  if (items > 0x7fffffffLL) return LG_ERR_SHAPE;
  const int grid = items < sms ? static_cast<int>(items) : sms;
  lg::tm::Params p{static_cast<const uint8_t*>(a_tiles), static_cast<const uint8_t*>(b_tiles), tpa, tpb, m_len, n_len,
                   row_part, col_part, row_lse, col_lse, row_ls, col_ls, scores, rowkey, colkey, scale,
                   B, M, N, rt, ct64, cw, chunks};
  lg::tm::assign_tmem_kernel<1><<<grid, lg::tm::kThreads, lg::tm::kSmem, stream>>>(p);
  lg::launch_assign_reduce(row_part, col_part, z0, z1, m_len, n_len, row_lse, col_lse, row_ls, col_ls, scores, B, M, N,
                           4 * rt, ct64, stream);
  lg::tm::assign_tmem_kernel<2><<<grid, lg::tm::kThreads, lg::tm::kSmem, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}

namespace lg {
int launch_assign_tmem_pass2(const void* a_tiles, const void* b_tiles, int tpa, int tpb, const int* m_len,
                             const int* n_len, const float* row_lse, const float* col_lse, const float* row_ls,
                             const float* col_ls, float* scores, unsigned long long* rowkey,
                             unsigned long long* colkey, float scale, int B, int M, int N, cudaStream_t stream) {
  const int rt = (M + 127) / 128, ct64 = (N + 63) / 64;
  static bool attr_done_dev[lg::kMaxDevices] = {};  // per device; a benign race sets it twice
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(tm::assign_tmem_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, tm::kSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  int cw = 16;
  while (cw > 1 && static_cast<long long>(B) * rt * ((ct64 + cw - 1) / cw) < 4LL * sms) cw >>= 1;
  const int chunks = (ct64 + cw - 1) / cw;
  const long long items = static_cast<long long>(B) * rt * chunks;
  if (items > 0x7fffffffLL) return LG_ERR_SHAPE;
  const int grid = items < sms ? static_cast<int>(items) : sms;
  tm::Params p{static_cast<const uint8_t*>(a_tiles), static_cast<const uint8_t*>(b_tiles), tpa, tpb, m_len, n_len,
               nullptr, nullptr, const_cast<float*>(row_lse), const_cast<float*>(col_lse), const_cast<float*>(row_ls),
               const_cast<float*>(col_ls), scores, rowkey, colkey, scale, B, M, N, rt, ct64, cw, chunks};
  tm::assign_tmem_kernel<2><<<grid, tm::kThreads, tm::kSmem, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}
}  // namespace lg
