// lg_assign : MatchAssignment similarity GEMM on tensor cores fused with the double
// log-softmax + log-sigmoid matchability of sigmoid_log_double_softmax.
//
// Replaces MatchAssignment.forward from the einsum on (reference
// gluefactory/models/matchers/lightglue.py:311-315) and sigmoid_log_double_softmax (:284-296).
//
// Precision: the similarity needs fp32-class accuracy in BOTH precision modes (bf16 operands
// alone cost 1.9e-2 on log_assignment), so the GEMM runs as TF32x3 on
// tcgen05 (kind::tf32): x = hi + lo with hi = rn_tf32(x), sim = hi.hi + hi.lo + lo.hi, fp32
// accumulation in TMEM (dropped lo.lo term ~2^-22).
//
// Two passes over 128x128 similarity tiles by persistent CTAs (one per SM), the tile never leaves the SM:
//   pass 1 (stats): per-tile row and column (max, sum exp) partials -> tiny partial buffers;
//   reduce:         merge partials into row / column logsumexp, logsigmoid terms, dustbins;
//   pass 2 (write): recompute the tile (cheaper than storing + reloading 16.8 MB/pair of sim),
//                   write (s - lse_r) + (s - lse_c) + (logsig z0 + logsig z1) through a
//                   shared-memory transpose so every warp stores whole 128-byte row segments.
// Output row pitch is N+1 floats (odd): rows are not 16 B aligned, hence 4-byte coalesced stores.
#include "lg_common.cuh"
#include "lg_keys.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kAssignK = 256;                 // descriptor dim
constexpr int kKAtoms = kAssignK / 32;        // 8 swizzle atoms of 32 fp32 along K
constexpr int kAtomTile = 16384;              // [128 rows][32 fp32]
constexpr int kOperandStage = 2 * kAtomTile;  // hi | lo
constexpr int kAssignStages = 2;
constexpr int kAssignThreads = 192;
constexpr int kStagePitch = 129;              // fp32 transpose buffer pitch (conflict-free both ways)

// ---------------------------------------------------------------------------- operand pack
// md (B,N,256) f32 -> [B][tiles][8 atoms][hi 16 KB | lo 16 KB] SWIZZLE_128B tile images
__global__ void __launch_bounds__(1024) assign_pack_kernel(const float* __restrict__ md, const int* __restrict__ len,
                                                           uint8_t* __restrict__ tiles, int N, int ntiles,
                                                           float prescale) {
  const int tile = blockIdx.x, b = blockIdx.y;
  const int n_valid = len ? min(len[b], N) : N;
  uint8_t* dst = tiles + (static_cast<size_t>(b) * ntiles + tile) * (kKAtoms * kOperandStage);
  for (int task = threadIdx.x; task < kTile * 64; task += blockDim.x) {
    const int r = task >> 6, ch = task & 63;  // 64 chunks of 4 floats per row
    const int n = tile * kTile + r;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n < n_valid) x = __ldg(reinterpret_cast<const float4*>(md + (static_cast<size_t>(b) * N + n) * kAssignK) + ch);
    float v[4] = {x.x * prescale, x.y * prescale, x.z * prescale, x.w * prescale}, hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t u = __float_as_uint(v[i]);
      u += 0xFFFu + ((u >> 13) & 1u);  // round-to-nearest-even at the tf32 mantissa boundary
      u &= 0xFFFFE000u;
      hi[i] = __uint_as_float(u);
      lo[i] = v[i] - hi[i];            // exact
    }
    uint8_t* at = dst + (ch >> 3) * kOperandStage + sw128_offset(r, ch & 7);
    *reinterpret_cast<float4*>(at) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(at + kAtomTile) = make_float4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// ---------------------------------------------------------------------------- tile kernel
struct AssignParams {
  const uint8_t* a_tiles;   // image 0 operand tiles
  const uint8_t* b_tiles;   // image 1 operand tiles
  const int* m_len;
  const int* n_len;
  float2* row_part;         // pass 1: [B][ct][M] (max, sum)
  float2* col_part;         // pass 1: [B][rt][N]
  const float* row_lse;     // pass 2: [B][M]
  const float* col_lse;     // pass 2: [B][N]
  const float* row_ls;      // pass 2: logsigmoid(z0) [B][M]
  const float* col_ls;      // pass 2: logsigmoid(z1) [B][N]
  float* scores;            // pass 2: (B, M+1, N+1)
  unsigned long long* rowkey;  // pass 2, optional: [B][M] best (score, column) key of every row (lg_keys.cuh)
  unsigned long long* colkey;  // pass 2, optional: [B][N] best (score, row) key of every column
  int B, M, N, rt, ct;      // sizes, tile counts
  float* sim_out;           // pass 2, optional: raw similarity (B, M, N) (NearestNeighborMatcher's "similarity")
};

struct __align__(8) AssignBarriers {
  uint64_t full[kAssignStages];
  uint64_t empty[kAssignStages];
  uint64_t acc_full;
  uint64_t acc_empty;   // 128 arrivals: the epilogue has read the four accumulators of this tile
  uint32_t tmem_base;
};

constexpr int kTransposeBytes = (kTile * kStagePitch * 4 + 1023) & ~1023;
constexpr int kLenCache = 1024;  // per-pair lengths of the first kLenCache pairs are cached in shared memory
constexpr int kAssignSmem = 1024 + kAssignStages * 2 * kOperandStage + kTransposeBytes + 2 * kTile * 4 + kTile * 8 +
                            2 * kLenCache * 4 + static_cast<int>(sizeof(AssignBarriers));

// One similarity tile of the (b, rtile, ctile) grid, linearised with ctile fastest.
struct AssignTile {
  int b, rtile, ctile, row0, col0, vrows, vcols;
};

__device__ __forceinline__ AssignTile assign_tile(const AssignParams& p, const int* s_len, int it) {
  AssignTile t;
  t.ctile = it % p.ct;
  const int br = it / p.ct;
  t.rtile = br % p.rt;
  t.b = br / p.rt;
  int m = p.M, n = p.N;
  if (p.m_len) m = min(t.b < kLenCache ? s_len[2 * t.b] : p.m_len[t.b], p.M);
  if (p.n_len) n = min(t.b < kLenCache ? s_len[2 * t.b + 1] : p.n_len[t.b], p.N);
  t.row0 = t.rtile * kTile;
  t.col0 = t.ctile * kTile;
  t.vrows = max(0, min(kTile, m - t.row0));
  t.vcols = max(0, min(kTile, n - t.col0));
  return t;
}

__device__ __forceinline__ float exp_fast(float x) {  // x <= 0; 2 ulp of the exact exponential
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}

// Persistent: one CTA per SM walks tiles it, it + grid, ...  The operand ring runs ahead across
// tile boundaries and the tensor core starts tile k+1 as soon as the epilogue has pulled tile k
// out of TMEM, i.e. under the column statistics (pass 1) / the stores (pass 2) of tile k.
template <int kPass>
__global__ void __launch_bounds__(kAssignThreads, 1) assign_tile_kernel(const AssignParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // 1 KB aligned, still provably shared
  uint8_t* ring = base;
  float* stage = reinterpret_cast<float*>(base + kAssignStages * 2 * kOperandStage);
  float* s_rowterm = reinterpret_cast<float*>(base + kAssignStages * 2 * kOperandStage + kTransposeBytes);  // [2][128]
  unsigned long long* s_colkey = reinterpret_cast<unsigned long long*>(s_rowterm + 2 * kTile);          // [128]
  int* s_len = reinterpret_cast<int*>(s_colkey + kTile);                                                 // [kLenCache][2]
  AssignBarriers* bar = reinterpret_cast<AssignBarriers*>(s_len + 2 * kLenCache);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = p.B * p.rt * p.ct;
  const int64_t pitch = p.N + 1;

  for (int i = threadIdx.x; i < min(p.B, kLenCache); i += kAssignThreads) {
    s_len[2 * i] = p.m_len ? p.m_len[i] : p.M;
    s_len[2 * i + 1] = p.n_len ? p.n_len[i] : p.N;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < kAssignStages; ++s) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
    mbar_init(&bar->acc_full, 1);
    mbar_init(&bar->acc_empty, kTile);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp == 4) {
    // ------------------------------------------------------------------ producer: bulk copies of operand atoms
    if (lane == 0) {
      uint32_t g = 0;  // running ring step
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const AssignTile t = assign_tile(p, s_len, it);
        if (t.vrows == 0 || t.vcols == 0) continue;
        const uint8_t* ga = p.a_tiles + (static_cast<size_t>(t.b) * p.rt + t.rtile) * (kKAtoms * kOperandStage);
        const uint8_t* gb = p.b_tiles + (static_cast<size_t>(t.b) * p.ct + t.ctile) * (kKAtoms * kOperandStage);
        for (int ka = 0; ka < kKAtoms; ++ka, ++g) {
          const int s = g % kAssignStages;
          if (g >= kAssignStages) mbar_wait(&bar->empty[s], ((g / kAssignStages) - 1) & 1);
          mbar_arrive_expect_tx(&bar->full[s], 2 * kOperandStage);
          bulk_g2s(ring + s * 2 * kOperandStage, ga + ka * kOperandStage, kOperandStage, &bar->full[s]);
          bulk_g2s(ring + s * 2 * kOperandStage + kOperandStage, gb + ka * kOperandStage, kOperandStage, &bar->full[s]);
        }
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------------ tensor core issue (warp-uniform, one elected lane)
    constexpr uint32_t idesc = umma_idesc(kFmtTF32, 128, 128);
    const uint32_t ring_a = smem_u32(ring);
    const bool leader = elect_one();
    uint32_t g = 0, done = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const AssignTile t = assign_tile(p, s_len, it);
      if (t.vrows == 0 || t.vcols == 0) continue;
      if (done > 0) {  // the epilogue has read the previous tile out of TMEM
        mbar_wait(&bar->acc_empty, (done - 1) & 1);
        tc_fence_after();
      }
      for (int ka = 0; ka < kKAtoms; ++ka, ++g) {
        const int s = g % kAssignStages;
        mbar_wait(&bar->full[s], (g / kAssignStages) & 1);
        tc_fence_after();
        const uint32_t a_hi = ring_a + s * 2 * kOperandStage, a_lo = a_hi + kAtomTile;
        const uint32_t b_hi = a_hi + kOperandStage, b_lo = b_hi + kAtomTile;
        const uint64_t dah = umma_desc_kmajor_sw128(a_hi), dal = umma_desc_kmajor_sw128(a_lo);
        const uint64_t dbh = umma_desc_kmajor_sw128(b_hi), dbl = umma_desc_kmajor_sw128(b_lo);
        // The tensor core truncates (does not round) when it adds into the fp32 accumulator, so a
        // single 96-step chain loses ~5e-5 on |sim|~10.  The dominant hi.hi products are therefore
        // spread over three TMEM accumulators (<= 12 steps each) and the 2^-11-small cross terms
        // go to a fourth; the epilogue adds the four with round-to-nearest.
        const uint32_t acc_hh = tmem + (ka % 3) * 128, acc_x = tmem + 384;
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {  // 32 fp32 per atom = 4 x K8, 32 B per step
            umma_tf32_ss(acc_x, dal + 2 * kk, dbh + 2 * kk, idesc, (ka > 0 || kk > 0));
            umma_tf32_ss(acc_x, dah + 2 * kk, dbl + 2 * kk, idesc, 1);
            umma_tf32_ss(acc_hh, dah + 2 * kk, dbh + 2 * kk, idesc, (ka >= 3 || kk > 0));
          }
          umma_commit(&bar->empty[s]);
        }
        __syncwarp();
      }
      if (leader) umma_commit(&bar->acc_full);
      __syncwarp();
      ++done;
    }
  } else {
    // ------------------------------------------------------------------ epilogue (128 threads, thread == row)
    const int r = warp * 32 + lane;
    const uint32_t tS = tmem + (static_cast<uint32_t>(warp * 32) << 16);
    uint32_t done = 0;
    for (int it = blockIdx.x; it < items; it += gridDim.x) {
      const AssignTile t = assign_tile(p, s_len, it);
      const int b = t.b, row0 = t.row0, col0 = t.col0, vrows = t.vrows, vcols = t.vcols;
      if (vrows == 0 || vcols == 0) {
        // tile entirely in the padding of a variable-count batch
        if (kPass == 1) {
          if (row0 + r < p.M)
            p.row_part[(static_cast<size_t>(b) * p.ct + t.ctile) * p.M + row0 + r] = make_float2(-INFINITY, 0.f);
          if (col0 + r < p.N)
            p.col_part[(static_cast<size_t>(b) * p.rt + t.rtile) * p.N + col0 + r] = make_float2(-INFINITY, 0.f);
        } else {
          float* out = p.scores + static_cast<int64_t>(b) * (p.M + 1) * pitch;
          for (int i = r; i < kTile * kTile; i += kTile) {
            const int rr = row0 + (i >> 7), c = col0 + (i & 127);
            if (rr < p.M && c < p.N) out[rr * pitch + c] = -INFINITY;
          }
        }
        continue;
      }
      float rl = 0.f, ra = 0.f;
      if (kPass == 2 && row0 + r < p.M) {  // this row's logsumexp / log-sigmoid term, fetched under the MMA
        rl = p.row_lse[static_cast<size_t>(b) * p.M + row0 + r];
        ra = p.row_ls[static_cast<size_t>(b) * p.M + row0 + r];
      }
      mbar_wait(&bar->acc_full, done & 1);
      tc_fence_after();
      asm volatile("bar.sync 1, 128;" ::: "memory");  // every warp is done with the previous tile's transpose buffer
      if (kPass == 2) { s_rowterm[r] = rl; s_rowterm[kTile + r] = ra; s_colkey[r] = 0ull; }
      uint32_t v[32], w0[32], w1[32], w2[32];
      float rmax = -INFINITY, rsum = 0.f;
#pragma unroll
      for (int cb = 0; cb < 4; ++cb) {
        tmem_ld32(tS + cb * 32, v);
        tmem_ld32(tS + 128 + cb * 32, w0);
        tmem_ld32(tS + 256 + cb * 32, w1);
        tmem_ld32(tS + 384 + cb * 32, w2);
        tmem_wait_ld();
        if (cb == 3) {  // accumulators are in registers: the tensor core may start the next tile
          tc_fence_before();
          mbar_arrive(&bar->acc_empty);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = __uint_as_float(v[i]) + __uint_as_float(w0[i]);
          s += __uint_as_float(w1[i]);
          s += __uint_as_float(w2[i]);
          v[i] = __float_as_uint(s);
        }
        if (kPass == 1) {
          float cmax = -INFINITY;
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (cb * 32 + i < vcols) cmax = fmaxf(cmax, __uint_as_float(v[i]));
          if (cmax > rmax) { rsum *= exp_fast(rmax - cmax); rmax = cmax; }
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (cb * 32 + i < vcols) rsum += exp_fast(__uint_as_float(v[i]) - rmax);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) stage[r * kStagePitch + cb * 32 + i] = __uint_as_float(v[i]);
      }
      ++done;
      if (kPass == 1 && row0 + r < p.M)
        p.row_part[(static_cast<size_t>(b) * p.ct + t.ctile) * p.M + row0 + r] =
            r < vrows ? make_float2(rmax, rsum) : make_float2(-INFINITY, 0.f);
      asm volatile("bar.sync 1, 128;" ::: "memory");  // epilogue warps only
      if (kPass == 1) {
        // thread == column: (max, sum exp) down the valid rows
        const int c = r;
        float cmax = -INFINITY, csum = 0.f;
        for (int i = 0; i < vrows; ++i) cmax = fmaxf(cmax, stage[i * kStagePitch + c]);
        for (int i = 0; i < vrows; ++i) csum += exp_fast(stage[i * kStagePitch + c] - cmax);
        if (col0 + c < p.N)
          p.col_part[(static_cast<size_t>(b) * p.rt + t.rtile) * p.N + col0 + c] =
              c < vcols ? make_float2(cmax, csum) : make_float2(-INFINITY, 0.f);
      } else {
        float* out = p.scores + static_cast<int64_t>(b) * (p.M + 1) * pitch;
        float lc[4], lb[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = q * 32 + lane;
          const bool ok = c < vcols;
          lc[q] = ok ? p.col_lse[static_cast<size_t>(b) * p.N + col0 + c] : 0.f;
          lb[q] = ok ? p.col_ls[static_cast<size_t>(b) * p.N + col0 + c] : 0.f;
        }
        const int rows_here = min(kTile, p.M - row0);
        const bool keys = p.rowkey != nullptr;
        // filter_matches (lightglue.py:321-337) needs the best column of every row and the best row of every column of
        // the valid block: gathered here, while the scores are on chip, instead of re-reading the matrix from HBM
        float cbv[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        uint32_t cbi[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
#pragma unroll 4
        for (int i = warp; i < rows_here; i += 4) {  // a warp stores one 512 B row segment per iteration
          const float rli = s_rowterm[i], rai = s_rowterm[kTile + i];
          float rv = -INFINITY;
          uint32_t rc = 0xFFFFFFFFu;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int c = q * 32 + lane;
            if (col0 + c < p.N) {
              const float s = stage[i * kStagePitch + c];
              const bool in = i < vrows && c < vcols;
              const float val = in ? ((s - rli) + (s - lc[q])) + (rai + lb[q]) : -INFINITY;
              out[static_cast<int64_t>(row0 + i) * pitch + col0 + c] = val;
              if (p.sim_out) p.sim_out[(static_cast<int64_t>(b) * p.M + row0 + i) * p.N + col0 + c] = s;
              if (keys && in) {  // strict > keeps the lowest index on ties: columns ascend with q, rows with i
                if (val > rv || rc == 0xFFFFFFFFu) { rv = val; rc = col0 + c; }
                if (val > cbv[q] || cbi[q] == 0xFFFFFFFFu) { cbv[q] = val; cbi[q] = row0 + i; }
              }
            }
          }
          if (keys && i < vrows) {
            // warp argmax in two redux steps: largest value, then the lowest column among the lanes that hold it
            const uint32_t u = rc == 0xFFFFFFFFu ? 0u : orderable_f32(rv);
            const uint32_t um = __reduce_max_sync(0xffffffffu, u);
            const uint32_t cm = __reduce_min_sync(0xffffffffu, u == um ? rc : 0xFFFFFFFFu);
            if (lane == 0 && cm != 0xFFFFFFFFu)
              atomicMax(p.rowkey + static_cast<size_t>(b) * p.M + row0 + i,
                        (static_cast<unsigned long long>(um) << 32) | (0xFFFFFFFFu - cm));
          }
        }
        if (keys) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (cbi[q] != 0xFFFFFFFFu) atomicMax(&s_colkey[q * 32 + lane], make_key(cbv[q], cbi[q]));
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (r < vcols) atomicMax(p.colkey + static_cast<size_t>(b) * p.N + col0 + r, s_colkey[r]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc<512>(tmem);
}
}  // namespace lg

namespace lg {

__device__ __forceinline__ float log_sigmoid(float x) {  // torch: min(x,0) - log1p(exp(-|x|))
  return fminf(x, 0.f) - log1pf(expf(-fabsf(x)));
}

// Merge the per-tile partials into logsumexp vectors, evaluate the log-sigmoid terms and write
// the dustbin row / column / corner of the (M+1)x(N+1) matrix (lightglue.py:289,294-295).
// grid (ceil(max(M,N)+1 / 256), B)
__global__ void assign_reduce_kernel(const float2* __restrict__ row_part, const float2* __restrict__ col_part,
                                     const float* __restrict__ z0, const float* __restrict__ z1,
                                     const int* __restrict__ m_len, const int* __restrict__ n_len,
                                     float* __restrict__ row_lse, float* __restrict__ col_lse,
                                     float* __restrict__ row_ls, float* __restrict__ col_ls,
                                     float* __restrict__ scores, int M, int N, int rt, int ct, int plain) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = m_len ? min(m_len[b], M) : M, n = n_len ? min(n_len[b], N) : N;
  const int64_t pitch = N + 1;
  float* out = scores + static_cast<int64_t>(b) * (M + 1) * pitch;
  if (i < M) {
    float mx = -INFINITY, sum = 0.f;
    for (int t = 0; t < ct; ++t) mx = fmaxf(mx, row_part[(static_cast<size_t>(b) * ct + t) * M + i].x);
    for (int t = 0; t < ct; ++t) {
      const float2 pr = row_part[(static_cast<size_t>(b) * ct + t) * M + i];
      if (pr.y > 0.f) sum += pr.y * expf(pr.x - mx);
    }
    // plain: no matchability terms, zero dustbins (NearestNeighborMatcher's log_assignment)
    const float z = plain ? 0.f : z0[static_cast<size_t>(b) * M + i];
    row_lse[static_cast<size_t>(b) * M + i] = mx + logf(sum);
    row_ls[static_cast<size_t>(b) * M + i] = plain ? 0.f : log_sigmoid(z);
    out[i * pitch + N] = i < m ? (plain ? 0.f : log_sigmoid(-z)) : -INFINITY;
  }
  if (i < N) {
    float mx = -INFINITY, sum = 0.f;
    for (int t = 0; t < rt; ++t) mx = fmaxf(mx, col_part[(static_cast<size_t>(b) * rt + t) * N + i].x);
    for (int t = 0; t < rt; ++t) {
      const float2 pr = col_part[(static_cast<size_t>(b) * rt + t) * N + i];
      if (pr.y > 0.f) sum += pr.y * expf(pr.x - mx);
    }
    const float z = plain ? 0.f : z1[static_cast<size_t>(b) * N + i];
    col_lse[static_cast<size_t>(b) * N + i] = mx + logf(sum);
    col_ls[static_cast<size_t>(b) * N + i] = plain ? 0.f : log_sigmoid(z);
    out[static_cast<int64_t>(M) * pitch + i] = i < n ? (plain ? 0.f : log_sigmoid(-z)) : -INFINITY;
  }
  if (i == 0) out[static_cast<int64_t>(M) * pitch + N] = 0.f;
}

}  // namespace lg

static inline size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

namespace lg {
// host launcher shared with lg_assign_strip.cu (declared in lg_assign.cuh)
void launch_assign_reduce(const float2* row_part, const float2* col_part, const float* z0, const float* z1,
                          const int* m_len, const int* n_len, float* row_lse, float* col_lse, float* row_ls,
                          float* col_ls, float* scores, int B, int M, int N, int rt, int ct, cudaStream_t stream,
                          int plain) {
  const int pts = (M > N ? M : N);
  assign_reduce_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(row_part, col_part, z0, z1, m_len, n_len,
                                                                      row_lse, col_lse, row_ls, col_ls, scores, M, N,
                                                                      rt, ct, plain);
}
}  // namespace lg

extern "C" size_t lg_assign_ws_bytes(int B, int M, int N) {
  if (B <= 0 || M <= 0 || N <= 0) return 0;
  const size_t rt = (M + 127) / 128, ct = (N + 127) / 128;
  size_t bytes = 0;
  bytes += align256(B * rt * lg::kKAtoms * static_cast<size_t>(lg::kOperandStage));  // a tiles
  bytes += align256(B * ct * lg::kKAtoms * static_cast<size_t>(lg::kOperandStage));  // b tiles
  bytes += align256(B * ct * static_cast<size_t>(M) * sizeof(float2));               // row partials
  bytes += align256(B * rt * static_cast<size_t>(N) * sizeof(float2));               // col partials
  bytes += 2 * align256(static_cast<size_t>(B) * M * 4) + 2 * align256(static_cast<size_t>(B) * N * 4);
  return bytes;
}

static int assign_impl(const float* md0, const float* md1, const float* z0, const float* z1, const int* m_len,
                       const int* n_len, float prescale, float* scores, void* ws, unsigned long long* rowkey,
                       unsigned long long* colkey, int B, int M, int N, cudaStream_t stream, float* sim_out = nullptr,
                       int plain = 0) {
  if (!md0 || !md1 || !scores || !ws) return LG_ERR_NULL;
  if (!plain && (!z0 || !z1)) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0) return LG_ERR_SHAPE;
  const int rt = (M + 127) / 128, ct = (N + 127) / 128;
  uint8_t* w = static_cast<uint8_t*>(ws);
  uint8_t* a_tiles = w; w += align256(static_cast<size_t>(B) * rt * lg::kKAtoms * lg::kOperandStage);
  uint8_t* b_tiles = w; w += align256(static_cast<size_t>(B) * ct * lg::kKAtoms * lg::kOperandStage);
  float2* row_part = reinterpret_cast<float2*>(w); w += align256(static_cast<size_t>(B) * ct * M * sizeof(float2));
  float2* col_part = reinterpret_cast<float2*>(w); w += align256(static_cast<size_t>(B) * rt * N * sizeof(float2));
  float* row_lse = reinterpret_cast<float*>(w); w += align256(static_cast<size_t>(B) * M * 4);
  float* row_ls = reinterpret_cast<float*>(w); w += align256(static_cast<size_t>(B) * M * 4);
  float* col_lse = reinterpret_cast<float*>(w); w += align256(static_cast<size_t>(B) * N * 4);
  float* col_ls = reinterpret_cast<float*>(w);

  static bool attr_done_dev[lg::kMaxDevices] = {};  // per device; a benign race sets it twice
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(lg::assign_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         lg::kAssignSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(lg::assign_tile_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, lg::kAssignSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  lg::assign_pack_kernel<<<dim3(rt, B), 1024, 0, stream>>>(md0, m_len, a_tiles, M, rt, prescale);
  lg::assign_pack_kernel<<<dim3(ct, B), 1024, 0, stream>>>(md1, n_len, b_tiles, N, ct, prescale);
  lg::AssignParams p{a_tiles, b_tiles, m_len, n_len, row_part, col_part, row_lse, col_lse, row_ls, col_ls, scores,
                     rowkey, colkey, B, M, N, rt, ct, sim_out};
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  const long long items = static_cast<long long>(B) * rt * ct;
  if (items > 0x7fffffffLL) return LG_ERR_SHAPE;
  const int grid = items < sms ? static_cast<int>(items) : sms;  // persistent: one CTA per SM
  lg::assign_tile_kernel<1><<<grid, lg::kAssignThreads, lg::kAssignSmem, stream>>>(p);
  const int pts = (M > N ? M : N);
  lg::assign_reduce_kernel<<<dim3((pts + 255) / 256, B), 256, 0, stream>>>(row_part, col_part, z0, z1, m_len, n_len,
                                                                          row_lse, col_lse, row_ls, col_ls, scores, M,
                                                                          N, rt, ct, plain);
  lg::assign_tile_kernel<2><<<grid, lg::kAssignThreads, lg::kAssignSmem, stream>>>(p);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int lg_assign(const float* md0, const float* md1, const float* z0, const float* z1, const int* m_len,
                         const int* n_len, float prescale, float* scores, void* ws, int B, int M, int N,
                         cudaStream_t stream) {
  return assign_impl(md0, md1, z0, z1, m_len, n_len, prescale, scores, ws, nullptr, nullptr, B, M, N, stream);
}

extern "C" int lg_assign_keys(const float* md0, const float* md1, const float* z0, const float* z1, const int* m_len,
                              const int* n_len, float prescale, float* scores, void* ws, void* keys_ws, int B, int M,
                              int N, cudaStream_t stream) {
  if (!keys_ws) return LG_ERR_NULL;
  if (B <= 0 || M <= 0 || N <= 0) return LG_ERR_SHAPE;
  cudaError_t e = cudaMemsetAsync(keys_ws, 0, lg_filter_ws_bytes(B, M, N), stream);
  if (e != cudaSuccess) return static_cast<int>(e);
  auto* rowkey = static_cast<unsigned long long*>(keys_ws);
  return assign_impl(md0, md1, z0, z1, m_len, n_len, prescale, scores, ws, rowkey, rowkey + static_cast<size_t>(B) * M,
                     B, M, N, stream);
}

// NearestNeighborMatcher._forward (reference gluefactory/models/matchers/nearest_neighbor_matcher.py:52-72): similarity
// = d0 . d1^T (TF32x3) written out, log_assignment = log_softmax(sim, -1) + log_softmax(sim, -2) with zero dustbins.
extern "C" int lg_nn_assign(const float* d0, const float* d1, const int* m_len, const int* n_len, float* sim_out,
                            float* scores, void* ws, int B, int M, int N, cudaStream_t stream) {
  if (!sim_out) return LG_ERR_NULL;
  return assign_impl(d0, d1, nullptr, nullptr, m_len, n_len, 1.0f, scores, ws, nullptr, nullptr, B, M, N, stream,
                     sim_out, 1);
}
