// lg_assign_f16 : MatchAssignment similarity + double log-softmax from fp16 operand tiles (the matcher's path).
//
// Replaces MatchAssignment.forward from the einsum on (reference gluefactory/models/matchers/lightglue.py:311-315),
// sigmoid_log_double_softmax (:284-296) and -- with keys -- the row / column argmax of filter_matches (:321-337).
// Operands are the fp16 tiles lg_linear_f16 writes for final_proj: [row tile][4 K atoms][16 KB], each atom the byte
// image of a SWIZZLE_128B K-major tile, so ONE tcgen05.mma kind::f16 product per K16 step gives the similarity
// (11-bit operands: ~1e-3 on log_assignment inside the bf16 path's 2e-2 budget; the 3-term split-bf16 and TF32x3
// kernels this file replaces on the matcher's path spent 3x the tensor work and 2x the operand bytes on digits the
// bf16 attention had already lost).  lg_assign.cu keeps TF32x3 for b200.precision = fp32.
//
// Three launches + a small merge:
//   stats<rows>   persistent tcgen05 kernel, thread == row of the similarity: online (max, sum exp2) over the columns
//   stats<cols>   THE SAME kernel with the operands swapped (rows of S^T): no transposition, no column atomics
//   merge         partials -> logsumexp, log-sigmoid terms, dustbin row / column / corner
//   write         recompute S, value = 2 s scale + (ls_r - lse_r) + (ls_c - lse_c) as one FMA, staged per warp and
//                 written as 16-byte vectors to rows whose pitch is a multiple of 32 floats (every 64-byte row segment
//                 is two whole sectors), filter keys gathered on the way
// Work item = (pair, TWO row tiles, chunk of column tiles): both row tiles (128 KB) stay resident in shared memory and
// every streamed 16 KB column atom feeds two MMAs, i.e. 32 KB of L2 -> shared-memory traffic per 128 x 128 tile
// (the L2 -> SM path delivers ~33 B/clk/SM: at 64 KB per tile the tensor pipe would wait for operands).
// Roles: warps 0-15 epilogue (4 warpgroups = 4 TMEM accumulators = {row tile 0, 1} x {even, odd step}), warp 16
// producer (bulk copies), warp 17 MMA issue.
#include <cuda_fp16.h>
#include "lg_common.cuh"
#include "lg_keys.cuh"
#include "lg_b200.h"

namespace lg {
namespace af {

constexpr int kAtom = 16384;                 // [128 rows][64 halves]
constexpr int kKAtoms = 4;                   // K = 256
constexpr int kTileBytes = kKAtoms * kAtom;  // 64 KB: one 128 x 256 operand tile
// Stats pass: TWO row tiles per item stay resident (each streamed column atom feeds two MMAs: 32 KB of L2 -> SM
// traffic per 128 x 128 tile) behind a 5-atom ring.  Write pass: bound by the HBM write of 64 KB per tile, not by
// operand delivery: ONE row tile per item, a 6-atom ring, and four accumulators = four warpgroups in flight (each
// takes every fourth tile), so the tensor core runs three tiles ahead of the stores.
constexpr int kRT1 = 2, kRT2 = 1;
constexpr int kStages1 = 5, kStages2 = 6;
constexpr int kWGs = 4;
constexpr int kEpiWarps = 4 * kWGs;
constexpr int kThreads = 32 * (kEpiWarps + 2);
constexpr int kChunk = 16;                   // columns per store chunk (64-byte row segments)
constexpr int kStgPitch = 20;                // floats: rows 16-byte aligned, 8 lanes x 16 B cover 32 banks
constexpr int kWarpStage = 32 * kStgPitch * 4;
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kFmtF16 = 0;
constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;

struct Params {
  const uint8_t* a_tiles;  // row operand: image b starts at row tile b * tpa
  const uint8_t* b_tiles;  // column operand
  int tpa, tpb;
  const int* m_len;        // valid rows / columns per pair (null = all)
  const int* n_len;
  float2* part;            // stats: [B][2 * chunks][M] (max in log2 units, sum)
  const float* rowc;       // write: [B][M] log-sigmoid - logsumexp of the row
  const float* colc;       // write: [B][N]
  float* scores;           // write: (B, M+1, pitch)
  int64_t pitch;
  unsigned long long* rowkey;
  unsigned long long* colkey;
  float scale;             // similarity = scale * <a, b>
  int B, M, N, rt, ct, rt2, cw, chunks;  // rt2: groups of kRT row tiles
};

struct __align__(8) Barriers {
  uint64_t a_full[kKAtoms], a_empty[kKAtoms];  // per K atom (both row tiles): the next item's rows stream in behind the last step
  uint64_t full[8], empty[8];
  uint64_t acc_full[kWGs];   // stats pass: [step parity] (both accumulators of the step); write pass: [accumulator]
  uint64_t acc_empty[kWGs];  // per accumulator
  uint32_t tmem_base;
};

// write pass scratch: column terms [4 warpgroups][128] f32 + column keys [16 warps][128] u64 (one slot per warp: plain
// stores, merged by the warpgroup at the end of the tile -- a 64-bit shared-memory atomicMax is a CAS loop)
constexpr int kScratch = kWGs * kTile * 4 + kEpiWarps * kTile * 8;
constexpr int kSmem1 = 1024 + kRT1 * kTileBytes + kStages1 * kAtom + static_cast<int>(sizeof(Barriers));
constexpr int kSmem2 = 1024 + kRT2 * kTileBytes + kStages2 * kAtom + kEpiWarps * kWarpStage + kScratch +
                       static_cast<int>(sizeof(Barriers));
static_assert(kSmem1 <= 232448 && kSmem2 <= 232448, "shared memory budget");

struct Item {
  int b, rp, c0, c1, n, nrt, nreal;  // row tiles RT*rp ..; nrt of them hold valid rows; the first nreal column tiles hold valid columns
  int vrows[2];
};

template <int RT>
__device__ __forceinline__ Item decode(const Params& p, int it) {
  Item t;
  const int chunk = it % p.chunks, br = it / p.chunks;
  t.rp = br % p.rt2;
  t.b = br / p.rt2;
  const int m = p.m_len ? min(p.m_len[t.b], p.M) : p.M;
  t.n = p.n_len ? min(p.n_len[t.b], p.N) : p.N;
  t.vrows[0] = max(0, min(kTile, m - RT * t.rp * kTile));
  t.vrows[1] = RT == 2 ? max(0, min(kTile, m - (RT * t.rp + 1) * kTile)) : 0;
  t.nrt = (t.vrows[0] > 0) + (t.vrows[1] > 0);
  t.c0 = chunk * p.cw;
  t.c1 = min(p.ct, t.c0 + p.cw);
  const int valid_ct = (t.n + kTile - 1) / kTile;
  t.nreal = t.nrt > 0 ? max(0, min(t.c1, valid_ct) - t.c0) : 0;
  return t;
}

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ void st_shared_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// kPass 1: row statistics.  kPass 2: values + keys.
template <int kPass>
__global__ void __launch_bounds__(kThreads, 1) assign_f16_kernel(const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* a_res = base;                                  // [RT row tiles][4 atoms][16 KB]
  constexpr int RT = kPass == 1 ? kRT1 : kRT2;
  constexpr int kStages = kPass == 1 ? kStages1 : kStages2;
  uint8_t* ring = a_res + RT * kTileBytes;
  uint8_t* wstage = ring + kStages * kAtom;               // write pass only
  uint8_t* scratch = wstage + kEpiWarps * kWarpStage;
  Barriers* bar = reinterpret_cast<Barriers*>(kPass == 1 ? wstage : scratch + kScratch);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = p.B * p.rt2 * p.chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kKAtoms; ++s) { mbar_init(&bar->a_full[s], 1); mbar_init(&bar->a_empty[s], 1); }
    for (int s = 0; s < kStages; ++s) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
    for (int s = 0; s < kWGs; ++s) mbar_init(&bar->acc_full[s], 1);
    for (int s = 0; s < kWGs; ++s) mbar_init(&bar->acc_empty[s], kTile);
    fence_mbar_init();
  }
  if (warp == kEpiWarps + 1) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp == kEpiWarps) {
    // ------------------------------------------------------------------ producer
    if (lane == 0) {
      uint32_t g = 0, loaded = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const Item t = decode<RT>(p, it);
        if (t.nreal == 0) continue;
        const uint8_t* ga = p.a_tiles + (static_cast<size_t>(t.b) * p.tpa + RT * t.rp) * kTileBytes;
        for (int j = 0; j < t.nreal; ++j) {
          const uint8_t* gb = p.b_tiles + (static_cast<size_t>(t.b) * p.tpb + t.c0 + j) * kTileBytes;
          for (int ka = 0; ka < kKAtoms; ++ka) {
            if (j == 0) {  // atom ka of the new row tiles, as soon as the previous item's last step is past it
              if (loaded > 0) mbar_wait(&bar->a_empty[ka], (loaded - 1) & 1);
              mbar_arrive_expect_tx(&bar->a_full[ka], t.nrt * kAtom);
              for (int r = 0; r < t.nrt; ++r)
                bulk_g2s(a_res + (r * kKAtoms + ka) * kAtom, ga + (static_cast<size_t>(r) * kKAtoms + ka) * kAtom, kAtom,
                         &bar->a_full[ka]);
            }
            const int s = g % kStages;
            if (g >= kStages) mbar_wait(&bar->empty[s], ((g / kStages) - 1) & 1);
            mbar_arrive_expect_tx(&bar->full[s], kAtom);
            bulk_g2s(ring + s * kAtom, gb + ka * kAtom, kAtom, &bar->full[s]);
            ++g;
          }
        }
        ++loaded;
      }
    }
  } else if (warp == kEpiWarps + 1) {
    // ------------------------------------------------------------------ tensor core issue
    constexpr uint32_t idesc = umma_idesc(kFmtF16, 128, 128);
    const uint32_t a_addr = smem_u32(a_res), ring_addr = smem_u32(ring);
    const bool leader = elect_one();
    uint32_t g = 0, step = 0, loaded = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const Item t = decode<RT>(p, it);
      if (t.nreal == 0) continue;
      for (int j = 0; j < t.nreal; ++j, ++step) {
        // stats pass: accumulators 2*par + r of step parity par; write pass: accumulator step % 4
        const uint32_t par = step & 1, acc0 = kPass == 1 ? 2 * par : (step & 3);
        if (kPass == 1) {
          // BOTH warpgroups of the parity have released the step before the last one -- also the row-tile-1 warpgroup of
          // an item with one valid row tile, which holds no data but waits and arrives all the same: a warpgroup that
          // skipped completions of acc_full[par] could no longer wait on it by parity
          if (step >= 2) {
            mbar_wait(&bar->acc_empty[2 * par], ((step >> 1) - 1) & 1);
            mbar_wait(&bar->acc_empty[2 * par + 1], ((step >> 1) - 1) & 1);
          }
        } else if (step >= 4) {
          mbar_wait(&bar->acc_empty[acc0], ((step >> 2) - 1) & 1);
        }
        tc_fence_after();
        for (int ka = 0; ka < kKAtoms; ++ka) {
          if (j == 0) mbar_wait(&bar->a_full[ka], loaded & 1);
          const int s = g % kStages;
          mbar_wait(&bar->full[s], (g / kStages) & 1);
          tc_fence_after();
          const uint64_t db = umma_desc_kmajor_sw128(ring_addr + s * kAtom);
          if (leader) {
            for (int r = 0; r < t.nrt; ++r) {
              const uint64_t da = umma_desc_kmajor_sw128(a_addr + (r * kKAtoms + ka) * kAtom);
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)
                umma_bf16_ss(tmem + (acc0 + r) * kTile, da + 2 * kk, db + 2 * kk, idesc, (ka > 0 || kk > 0));
            }
            umma_commit(&bar->empty[s]);
            if (j == t.nreal - 1) umma_commit(&bar->a_empty[ka]);  // this item is done with atom ka of its rows
          }
          __syncwarp();
          ++g;
        }
        if (leader) umma_commit(&bar->acc_full[kPass == 1 ? par : acc0]);
        __syncwarp();
      }
      ++loaded;
    }
  } else {
    // ------------------------------------------------------------------ epilogue: thread == row (TMEM lane)
    const int wg = warp >> 2, wq = warp & 3;
    const int r = wg & 1;                 // stats pass: row tile of the item ...
    const uint32_t par = wg >> 1;         // ... and steps of this parity; write pass: warpgroup == accumulator == step % 4
    const int row = wq * 32 + lane;
    const uint32_t t_acc = tmem + (static_cast<uint32_t>(wq * 32) << 16) + wg * kTile;
    uint32_t gstep = 0;

    if (kPass == 1) {
      const float c = p.scale * kLog2e;
      const float2 c2 = make_float2(c, c);
      const int np = 2 * p.chunks;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const Item t = decode<RT>(p, it);
        const int chunk = it % p.chunks;
        const int grow = (2 * t.rp + r) * kTile + row;
        const bool row_ok = row < t.vrows[r];
        float m_run = -INFINITY, l_run = 0.f;
        {
          for (int j = 0; j < t.nreal; ++j) {
            const uint32_t s = gstep + j;
            if ((s & 1) != par) continue;
            const int vcols = min(kTile, t.n - (t.c0 + j) * kTile);
            mbar_wait(&bar->acc_full[par], (s >> 1) & 1);
            tc_fence_after();
            if (r >= t.nrt) {  // no valid row in this row tile: keep the hand-shake going
              tc_fence_before();
              mbar_arrive(&bar->acc_empty[wg]);
              continue;
            }
            auto process = [&](const uint32_t(&v)[32], int valid) {  // valid: columns of this group that count (<= 32)
              float mx = -INFINITY;
              if (valid >= 32) {
#pragma unroll
                for (int i = 0; i < 32; i += 2) mx = max3(mx, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
              } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) mx = fmaxf(mx, i < valid ? __uint_as_float(v[i]) : -INFINITY);
              }
              const float mn = fmaxf(m_run, mx * c);
              l_run *= ex2f(m_run - mn);  // -inf - finite = -inf -> 0
              m_run = mn;
              const float2 nm = make_float2(-mn, -mn);
              float2 s0 = make_float2(0.f, 0.f), s1 = s0, s2 = s0, s3 = s0;
#pragma unroll
              for (int pr = 0; pr < 16; ++pr) {
                const float2 x = __ffma2_rn(make_float2(__uint_as_float(v[2 * pr]), __uint_as_float(v[2 * pr + 1])), c2, nm);
                float2 e = make_float2(ex2f(x.x), ex2f(x.y));
                if (valid < 32) { e.x = 2 * pr < valid ? e.x : 0.f; e.y = 2 * pr + 1 < valid ? e.y : 0.f; }
                if ((pr & 3) == 0) s0 = __fadd2_rn(s0, e);
                else if ((pr & 3) == 1) s1 = __fadd2_rn(s1, e);
                else if ((pr & 3) == 2) s2 = __fadd2_rn(s2, e);
                else s3 = __fadd2_rn(s3, e);
              }
              l_run += ((s0.x + s0.y) + (s1.x + s1.y)) + ((s2.x + s2.y) + (s3.x + s3.y));
            };
            uint32_t va[32], vb[32];
            if (vcols == kTile) {  // loads one 32-column group ahead of the arithmetic
              tmem_ld32(t_acc, va);
              tmem_wait_ld();
              tmem_ld32(t_acc + 32, vb);
              process(va, 32);
              tmem_wait_ld();
              tmem_ld32(t_acc + 64, va);
              process(vb, 32);
              tmem_wait_ld();
              tmem_ld32(t_acc + 96, vb);
              process(va, 32);
              tmem_wait_ld();
              tc_fence_before();
              mbar_arrive(&bar->acc_empty[wg]);  // the accumulator is in registers: back to the tensor core
              process(vb, 32);
            } else {
              const int ng = (vcols + 31) / 32;
#pragma unroll 1
              for (int g4 = 0; g4 < ng; ++g4) {
                tmem_ld32(t_acc + g4 * 32, va);
                tmem_wait_ld();
                process(va, min(32, vcols - g4 * 32));
              }
              tc_fence_before();
              mbar_arrive(&bar->acc_empty[wg]);
            }
          }
        }
        gstep += t.nreal;
        if (grow < p.M && 2 * t.rp + r < p.rt)
          p.part[(static_cast<size_t>(t.b) * np + 2 * chunk + par) * p.M + grow] =
              row_ok ? make_float2(m_run, l_run) : make_float2(-INFINITY, 0.f);
      }
    } else {
      const int half = lane >> 4, cl = lane & 15;
      const uint32_t bar_id = 1 + wg;
      const uint32_t stg = smem_u32(wstage) + warp * kWarpStage;
      float* s_colc = reinterpret_cast<float*>(scratch) + wg * kTile;
      // column keys: one slot per (warp, column); the warpgroup's four warps cover rows 0-31 .. 96-127 of the tile
      unsigned long long* s_wkey = reinterpret_cast<unsigned long long*>(scratch + kWGs * kTile * 4) + wg * 4 * kTile;
      const bool keys = p.rowkey != nullptr;
      const float s2 = 2.f * p.scale;
      const int64_t pitch = p.pitch;
      uint32_t pad = 0;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const Item t = decode<RT>(p, it);
        const int row0 = t.rp * kTile;
        const int grow = row0 + row;
        const int vrows = t.vrows[0];
        const bool row_ok = row < vrows;
        const bool live = t.nrt > 0;
        float* out = p.scores + static_cast<int64_t>(t.b) * (p.M + 1) * pitch;
        const float rc = row_ok ? p.rowc[static_cast<size_t>(t.b) * p.M + grow] : 0.f;
        float rv = -INFINITY;
        uint32_t rci = kNone;
        for (int ctile = t.c0; ctile < t.c1; ++ctile) {
          const int j = ctile - t.c0, col0 = ctile * kTile;
          if (!live || j >= t.nreal) {
            // rows or columns entirely in the padding of a variable-count batch: -inf, the warpgroups take turns
            if ((pad++ & 3) != static_cast<uint32_t>(wg)) continue;
            for (int i = wq; i < kTile; i += 4)
              for (int q = 0; q < 4; ++q) {
                const int rr = row0 + i, cc = col0 + q * 32 + lane;
                if (rr < p.M && cc < p.N) out[rr * pitch + cc] = -INFINITY;
              }
            continue;
          }
          const uint32_t s = gstep + j;
          if ((s & 3) != static_cast<uint32_t>(wg)) continue;
          const int vcols = min(kTile, t.n - col0);        // columns that hold similarities
          const int scols = min(kTile, p.N - col0);        // columns that exist in the matrix
          s_colc[row] = row < vcols ? p.colc[static_cast<size_t>(t.b) * p.N + col0 + row] : 0.f;
          mbar_wait(&bar->acc_full[wg], (s >> 2) & 1);
          tc_fence_after();
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // column terms / key slots are in place
          // every element of this warp's 32 x 128 block valid and inside the matrix: no masks, vector stores
          const bool full = vcols == kTile && wq * 32 + 32 <= vrows;
          const int nld = (vcols + kChunk - 1) / kChunk, nst = (scols + kChunk - 1) / kChunk;
#pragma unroll 1
          for (int ch = 0; ch < nst; ++ch) {
            uint32_t v[kChunk];
            if (ch < nld) {
              tmem_ld16(t_acc + ch * kChunk, v);
              tmem_wait_ld();
              if (ch == nld - 1) {
                tc_fence_before();
                mbar_arrive(&bar->acc_empty[wg]);
              }
            } else {
#pragma unroll
              for (int i = 0; i < kChunk; ++i) v[i] = 0u;
            }
            float val[kChunk];
#pragma unroll
            for (int q = 0; q < kChunk / 4; ++q) {
              const float4 c4 = *reinterpret_cast<const float4*>(s_colc + ch * kChunk + 4 * q);
              // (s - lse_r) + (s - lse_c) + (ls_r + ls_c) as one fma: 2 s scale + ((ls_r - lse_r) + (ls_c - lse_c))
              // packed f32x2: the same fmaf(s, 2 scale, rc + c) per element, half the issue slots
              const float2 a = __ffma2_rn(make_float2(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1])),
                                          make_float2(s2, s2), __fadd2_rn(make_float2(rc, rc), make_float2(c4.x, c4.y)));
              const float2 b = __ffma2_rn(make_float2(__uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])),
                                          make_float2(s2, s2), __fadd2_rn(make_float2(rc, rc), make_float2(c4.z, c4.w)));
              val[4 * q] = a.x; val[4 * q + 1] = a.y; val[4 * q + 2] = b.x; val[4 * q + 3] = b.y;
            }
            if (!full) {
#pragma unroll
              for (int i = 0; i < kChunk; ++i) val[i] = (row_ok && ch * kChunk + i < vcols) ? val[i] : -INFINITY;
            }
            if (keys) {  // best column of the row so far: strict > keeps the lowest column on ties
              float cm = val[0];
#pragma unroll
              for (int i = 1; i + 1 < kChunk; i += 2) cm = max3(cm, val[i], val[i + 1]);
              cm = fmaxf(cm, val[kChunk - 1]);
              if (cm > rv || rci == kNone) {
                rv = cm;
                uint32_t first = kChunk - 1;
#pragma unroll
                for (int i = kChunk - 2; i >= 0; --i) first = val[i] == cm ? i : first;
                rci = col0 + ch * kChunk + first;
              }
            }
#pragma unroll
            for (int q = 0; q < kChunk / 4; ++q)
              st_shared_f4(stg + (lane * kStgPitch + 4 * q) * 4, val[4 * q], val[4 * q + 1], val[4 * q + 2], val[4 * q + 3]);
            __syncwarp();
            const int cbase = col0 + ch * kChunk;
            float cbv = -INFINITY;
            uint32_t cbi = kNone;
            if (full) {
              // lane -> (row rq of 8, 16-byte quarter f4 of the 64-byte segment): 8 rows x 64 B per instruction.
              // All shared-memory reads of the chunk (store path and key path) are issued before their first use.
              const int rq = lane >> 2, f4 = lane & 3;
              float* o = out + static_cast<int64_t>(row0 + wq * 32 + rq) * pitch + cbase + 4 * f4;
              float4 x4[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) x4[q] = ld_shared_f4(stg + ((q * 8 + rq) * kStgPitch + 4 * f4) * 4);
              float xs[kChunk];
              if (keys) {
#pragma unroll
                for (int jj = 0; jj < kChunk; ++jj)
                  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(xs[jj]) : "r"(stg + ((2 * jj + half) * kStgPitch + cl) * 4));
              }
#pragma unroll
              for (int q = 0; q < 4; ++q) *reinterpret_cast<float4*>(o + static_cast<int64_t>(q) * 8 * pitch) = x4[q];
              if (keys) {  // best row of the column among this warp's rows: lane == (row parity, column)
                // four independent chains of four rows, merged pairwise: strict > keeps the lowest row on ties
                float gv[4];
                uint32_t gi[4];
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                  gv[g] = xs[4 * g];
                  gi[g] = 2 * (4 * g) + half;
#pragma unroll
                  for (int k = 1; k < 4; ++k) {
                    const bool up = xs[4 * g + k] > gv[g];
                    gv[g] = up ? xs[4 * g + k] : gv[g];
                    gi[g] = up ? static_cast<uint32_t>(2 * (4 * g + k) + half) : gi[g];
                  }
                }
                const bool u1 = gv[1] > gv[0], u3 = gv[3] > gv[2];
                const float v01 = u1 ? gv[1] : gv[0], v23 = u3 ? gv[3] : gv[2];
                const uint32_t i01 = u1 ? gi[1] : gi[0], i23 = u3 ? gi[3] : gi[2];
                const bool u = v23 > v01;
                cbv = u ? v23 : v01;
                cbi = u ? i23 : i01;
              }
            } else {
              // lane == (row parity, column): scalar stores with bounds
              const int c = cbase + cl;
#pragma unroll 4
              for (int jj = 0; jj < kChunk; ++jj) {
                const int lr = 2 * jj + half, rr = wq * 32 + lr;  // row inside the warp's staging block / the tile
                float x;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(stg + (lr * kStgPitch + cl) * 4));
                if (c < p.N && row0 + rr < p.M) out[static_cast<int64_t>(row0 + rr) * pitch + c] = x;
              }
              if (keys) {
                const bool col_ok = ch * kChunk + cl < vcols;
#pragma unroll 4
                for (int jj = 0; jj < kChunk; ++jj) {
                  const int lr = 2 * jj + half;
                  float x;
                  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(stg + (lr * kStgPitch + cl) * 4));
                  const bool up = col_ok && wq * 32 + lr < vrows && (x > cbv || cbi == kNone);
                  cbv = up ? x : cbv;
                  cbi = up ? static_cast<uint32_t>(lr) : cbi;
                }
              }
            }
            if (keys) {
              // the other row parity of the same column: larger value, lower row on ties
              const float ov = __shfl_xor_sync(0xffffffffu, cbv, 16);
              const uint32_t oi = __shfl_xor_sync(0xffffffffu, cbi, 16);
              const bool take = oi != kNone && (cbi == kNone || ov > cbv || (ov == cbv && oi < cbi));
              cbv = take ? ov : cbv;
              cbi = take ? oi : cbi;
              if (half == 0)
                s_wkey[wq * kTile + ch * kChunk + cl] = cbi == kNone ? 0ull : make_key(cbv, row0 + wq * 32 + cbi);
            }
            __syncwarp();
          }
          if (keys) {
            asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            if (row < vcols) {  // thread == column: merge the four warps' candidates, publish
              unsigned long long k = s_wkey[row];
#pragma unroll
              for (int w4 = 1; w4 < 4; ++w4) {
                const unsigned long long o = s_wkey[w4 * kTile + row];
                k = o > k ? o : k;
              }
              if (k != 0ull) atomicMax(p.colkey + static_cast<size_t>(t.b) * p.N + col0 + row, k);
            }
          }
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");  // terms / keys may be overwritten for the next tile
        }
        gstep += t.nreal;
        if (keys && row_ok && rci != kNone) atomicMax(p.rowkey + static_cast<size_t>(t.b) * p.M + grow, make_key(rv, rci));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps + 1) tmem_dealloc<512>(tmem);
}

__device__ __forceinline__ float log_sigmoid(float x) {  // torch: min(x,0) - log1p(exp(-|x|))
  return fminf(x, 0.f) - log1pf(expf(-fabsf(x)));
}

// Partials -> log-sigmoid - logsumexp per row / column, dustbin row / column / corner (lightglue.py:289,294-295).
__global__ void assign_f16_merge_kernel(const float2* __restrict__ rpart, int rnp, const float2* __restrict__ cpart,
                                        int cnp, const float* __restrict__ z0, const float* __restrict__ z1,
                                        const int* __restrict__ m_len, const int* __restrict__ n_len,
                                        float* __restrict__ rowc, float* __restrict__ colc, float* __restrict__ scores,
                                        int64_t pitch, int M, int N) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = m_len ? min(m_len[b], M) : M, n = n_len ? min(n_len[b], N) : N;
  float* out = scores + static_cast<int64_t>(b) * (M + 1) * pitch;
  auto lse = [](const float2* part, int np, size_t stride, size_t idx) {
    float mx = -INFINITY, sum = 0.f;
    for (int t = 0; t < np; ++t) mx = fmaxf(mx, part[t * stride + idx].x);
    for (int t = 0; t < np; ++t) {
      const float2 pr = part[t * stride + idx];
      if (pr.y > 0.f) sum += pr.y * exp2f(pr.x - mx);
    }
    return (mx + log2f(sum)) * kLn2;
  };
  if (i < M) {
    const float z = z0[static_cast<size_t>(b) * M + i];
    rowc[static_cast<size_t>(b) * M + i] = log_sigmoid(z) - lse(rpart + static_cast<size_t>(b) * rnp * M, rnp, M, i);
    out[i * pitch + N] = i < m ? log_sigmoid(-z) : -INFINITY;
  }
  if (i < N) {
    const float z = z1[static_cast<size_t>(b) * N + i];
    colc[static_cast<size_t>(b) * N + i] = log_sigmoid(z) - lse(cpart + static_cast<size_t>(b) * cnp * N, cnp, N, i);
    out[static_cast<int64_t>(M) * pitch + i] = i < n ? log_sigmoid(-z) : -INFINITY;
  }
  if (i == 0) out[static_cast<int64_t>(M) * pitch + N] = 0.f;
}

}  // namespace af
}  // namespace lg

static inline size_t af_align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// column tiles per item: the fewer items, the less often the two row tiles (128 KB) are re-staged; the more items, the
// better the last wave fills the SMs.  Minimise (bytes per tile) / (fill of the last wave).
static int af_pick_cw(int B, int rt2, int ct, int sms, int rt_per_item) {
  const double stream_kb = 64.0 / rt_per_item;  // column-operand bytes per 128 x 128 tile
  int best = 1;
  double best_cost = 1e30;
  for (int cw = 1; cw <= ct; cw *= 2) {
    const long long items = static_cast<long long>(B) * rt2 * ((ct + cw - 1) / cw);
    const long long waves = (items + sms - 1) / sms;
    const double fill = static_cast<double>(items) / static_cast<double>(waves * sms);
    const double cost = (stream_kb + 64.0 / cw) / fill;
    if (cost < best_cost) { best_cost = cost; best = cw; }
    if (cw * 2 > ct && cw != ct) {  // also try "all column tiles in one item"
      const long long it2 = static_cast<long long>(B) * rt2;
      const long long w2 = (it2 + sms - 1) / sms;
      const double c2 = (stream_kb + 64.0 / ct) / (static_cast<double>(it2) / static_cast<double>(w2 * sms));
      if (c2 < best_cost) { best_cost = c2; best = ct; }
    }
  }
  return best;
}

extern "C" size_t lg_assign_tiles_ws_bytes(int B, int M, int N) {
  if (B <= 0 || M <= 0 || N <= 0) return 0;
  const size_t rt = (M + 127) / 128, ct = (N + 127) / 128;
  size_t bytes = 0;
  bytes += af_align256(static_cast<size_t>(B) * 2 * ct * M * sizeof(float2));  // row partials: <= 2 per column tile
  bytes += af_align256(static_cast<size_t>(B) * 2 * rt * N * sizeof(float2));  // column partials
  bytes += af_align256(static_cast<size_t>(B) * M * 4) + af_align256(static_cast<size_t>(B) * N * 4);
  return bytes;
}

extern "C" int lg_assign_tiles(const void* a_tiles, const void* b_tiles, int tpa, int tpb, const float* z0,
                               const float* z1, const int* m_len, const int* n_len, float scale, float* scores,
                               int64_t pitch, void* ws, void* keys_ws, int B, int M, int N, cudaStream_t stream) {
  using namespace lg::af;
  if (!a_tiles || !b_tiles || !z0 || !z1 || !scores || !ws) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0 || pitch < N + 1) return LG_ERR_SHAPE;
  const int rt = (M + 127) / 128, ct = (N + 127) / 128;
  if (tpa < rt || tpb < ct) return LG_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(a_tiles) | reinterpret_cast<uintptr_t>(b_tiles) | reinterpret_cast<uintptr_t>(scores)) & 15)
    return LG_ERR_ALIGN;
  if (pitch % 4) return LG_ERR_ALIGN;  // 16-byte vector stores; a multiple of 32 makes every 64-byte segment two whole sectors
  uint8_t* w = static_cast<uint8_t*>(ws);
  float2* rpart = reinterpret_cast<float2*>(w); w += af_align256(static_cast<size_t>(B) * 2 * ct * M * sizeof(float2));
  float2* cpart = reinterpret_cast<float2*>(w); w += af_align256(static_cast<size_t>(B) * 2 * rt * N * sizeof(float2));
  float* rowc = reinterpret_cast<float*>(w); w += af_align256(static_cast<size_t>(B) * M * 4);
  float* colc = reinterpret_cast<float*>(w);

  unsigned long long* rowkey = nullptr;
  unsigned long long* colkey = nullptr;
  if (keys_ws) {
    cudaError_t e = cudaMemsetAsync(keys_ws, 0, lg_filter_ws_bytes(B, M, N), stream);
    if (e != cudaSuccess) return static_cast<int>(e);
    rowkey = static_cast<unsigned long long*>(keys_ws);
    colkey = rowkey + static_cast<size_t>(B) * M;
  }
  static bool attr_done_dev[lg::kMaxDevices] = {};
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(assign_f16_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem1);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(assign_f16_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  const int sms = lg::device_sm_count();
  auto make = [&](bool swapped, int rt_per_item) {
    Params p{};
    p.a_tiles = static_cast<const uint8_t*>(swapped ? b_tiles : a_tiles);
    p.b_tiles = static_cast<const uint8_t*>(swapped ? a_tiles : b_tiles);
    p.tpa = swapped ? tpb : tpa; p.tpb = swapped ? tpa : tpb;
    p.m_len = swapped ? n_len : m_len; p.n_len = swapped ? m_len : n_len;
    p.scale = scale; p.B = B;
    p.M = swapped ? N : M; p.N = swapped ? M : N;
    p.rt = (p.M + 127) / 128; p.ct = (p.N + 127) / 128; p.rt2 = (p.rt + rt_per_item - 1) / rt_per_item;
    p.cw = af_pick_cw(B, p.rt2, p.ct, sms, rt_per_item);
    p.chunks = (p.ct + p.cw - 1) / p.cw;
    return p;
  };
  auto grid_of = [&](const Params& p) {
    const long long items = static_cast<long long>(p.B) * p.rt2 * p.chunks;
    return static_cast<int>(items < sms ? items : sms);
  };
  Params pr = make(false, kRT1), pc = make(true, kRT1), pw = make(false, kRT2);
  if (static_cast<long long>(B) * pr.rt2 * pr.chunks > 0x7fffffffLL || static_cast<long long>(B) * pc.rt2 * pc.chunks > 0x7fffffffLL ||
      static_cast<long long>(B) * pw.rt2 * pw.chunks > 0x7fffffffLL)
    return LG_ERR_SHAPE;
  pr.part = rpart;
  pc.part = cpart;
  assign_f16_kernel<1><<<grid_of(pr), kThreads, kSmem1, stream>>>(pr);
  assign_f16_kernel<1><<<grid_of(pc), kThreads, kSmem1, stream>>>(pc);
  const int pts = M > N ? M : N;
  assign_f16_merge_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(rpart, 2 * pr.chunks, cpart, 2 * pc.chunks, z0, z1,
                                                                         m_len, n_len, rowc, colc, scores, pitch, M, N);
  pw.rowc = rowc; pw.colc = colc; pw.scores = scores; pw.pitch = pitch; pw.rowkey = rowkey; pw.colkey = colkey;
  assign_f16_kernel<2><<<grid_of(pw), kThreads, kSmem2, stream>>>(pw);
  return static_cast<int>(cudaGetLastError());
}
