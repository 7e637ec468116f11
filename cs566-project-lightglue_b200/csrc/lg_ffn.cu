// lg_ffn : first half of the block MLP in one kernel,
//     h = GELU(LayerNorm([x | ctx] . W0'^T + b0'))         (512 -> 512, LayerNorm over the 512 outputs)
// i.e. ffn.0 (with out_proj / to_out folded in), ffn.1 and ffn.2 of SelfBlock / CrossBlock
// (reference gluefactory/models/matchers/lightglue.py:171-176,190-191 and :205-210,246-248).
//
// The pre-LayerNorm activations never leave the SM: the stand-alone path (lg_linear_f32 + lg_ln_gelu_tiles) wrote
// 268 MB of fp32 rows per launch with thread-per-row 16-byte stores and read them back.
//
// A CTA PAIR (cluster of 2, one TPC) owns a row tile of 128 tokens; CTA `rank` computes output columns
// [256 rank, 256 rank + 256) with its own single-CTA tcgen05 pipeline (A atom 32 KB + W atom 64 KB per stage, 2 stages,
// two 256-column TMEM accumulators so the MMAs of row tile i+1 run under the epilogue of tile i).  LayerNorm needs
// the statistics of the whole 512-wide row: every CTA reduces its half to (mean, M2 = sum (x - mean)^2), ships the
// pair of floats per row to its peer through distributed shared memory (st.async + mbarrier complete_tx, no
// cluster barrier), and merges the two halves exactly (Chan et al.):
//     mean = (m0 + m1) / 2,   M2 = M2_0 + M2_1 + (m0 - m1)^2 * 128.
// 320 threads: warps 0-7 epilogue (warp w: TMEM lanes 32 (w % 4).., columns 32 (w / 4).. of every 64-column atom),
// warp 8 producer, warp 9 MMA issue.  Output: split tiles (hi | lo bf16 images) staged in shared memory and written
// with one 32 KB bulk store per atom.
#include "lg_common.cuh"
#include "lg_tiles.cuh"
#include "lg_b200.h"

namespace lg {

constexpr int kFfnThreads = 320;
constexpr int kFfnStages = 2;
constexpr int kFfnNt = 256;                       // output columns per CTA
constexpr int kFfnWAtom = kFfnNt * 256;           // hi | lo of one K atom of the CTA's weight tile
constexpr int kFfnStage = kSplitAtom + kFfnWAtom;  // 96 KB

struct FfnParams {
  const uint8_t* a0;     // [row_tiles][a0_atoms][32 KB]
  const uint8_t* a1;     // [row_tiles][a1_atoms][32 KB], concatenated along K after a0 (may be null)
  int a0_atoms, a1_atoms;
  const uint8_t* w;      // lg_pack_weight(W (512, K), nt = 256): [2][k_atoms][hi | lo][256 x 128 B]
  const float* bias;     // [512]
  const float* gamma;    // [512] LayerNorm weight
  const float* beta;     // [512] LayerNorm bias
  float eps;
  uint8_t* out_tiles;    // [row_tiles][8][32 KB]
  int row_tiles;
};

struct __align__(8) FfnBarriers {
  uint64_t full[kFfnStages];
  uint64_t empty[kFfnStages];
  uint64_t acc_full[2];
  uint64_t acc_empty[2];
  uint64_t xchg_full[2];   // the peer's row statistics of this item have landed in xchg[buf]
  uint32_t tmem_base;
};

constexpr int kFfnXchgBytes = 2 * kTile * 8;  // [buf][row] (mean, M2)
constexpr int kFfnSmem = kFfnStages * kFfnStage + kSplitAtom + kFfnXchgBytes + static_cast<int>(sizeof(FfnBarriers));
static_assert(kFfnSmem <= 232448, "227 KB per CTA");

#ifdef LG_FFN_TRACE
__device__ unsigned long long g_ffn_trace[16 * 64];
__device__ __forceinline__ unsigned long long ffn_clock() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t));
  return t;
}
#define FTRACE(slot, idx) do { if (blockIdx.x == 8 && (idx) < 64) g_ffn_trace[(slot) * 64 + (idx)] = ffn_clock(); } while (0)
#else
#define FTRACE(slot, idx) do {} while (0)
#endif

__device__ __forceinline__ void epi_bar_sync_256() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__device__ __forceinline__ void st_async_peer(void* local_slot, float value, uint64_t* local_bar, uint32_t peer) {
  st_async_remote(local_slot, __float_as_uint(value), local_bar, peer);
}

// GELU(z) = z/2 (1 + erf(z / sqrt 2)), erf by Abramowitz-Stegun 7.1.26 (|err| <= 1.5e-7) with the 1/sqrt 2 folded
// into the constants: branch-free, 2 MUFU + 12 FP32 instructions.  Max abs error against fp64 GELU over [-9, 9]:
// 4.7e-7 (torch's own fp32 erf GELU: 1.2e-6) -- below the 2^-17 relative step of the hi|lo bf16 output format.
__device__ __forceinline__ float gelu_fast(float z) {
  const float az = fabsf(z);
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(az, 0.23164189f, 1.f)));
  float pl = fmaf(1.061405429f, t, -1.453152027f);
  pl = fmaf(pl, t, 1.421413741f);
  pl = fmaf(pl, t, -0.284496736f);
  pl = fmaf(pl, t, 0.254829592f);
  pl *= t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * z * -0.72134752044448170f));
  const float er = copysignf(fmaf(-pl, e, 1.f), z);
  const float hz = 0.5f * z;
  return fmaf(hz, er, hz);
}

// two elements at a time: the FP32 part as packed f32x2 instructions
__device__ __forceinline__ float2 gelu_fast2(float2 z) {
  const float2 d = __ffma2_rn(make_float2(fabsf(z.x), fabsf(z.y)), make_float2(0.23164189f, 0.23164189f),
                              make_float2(1.f, 1.f));
  float2 t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(d.y));
  // -(polynomial): the sign of 1 - pl * e is folded into the coefficients
  float2 pl = __ffma2_rn(make_float2(-1.061405429f, -1.061405429f), t, make_float2(1.453152027f, 1.453152027f));
  pl = __ffma2_rn(pl, t, make_float2(-1.421413741f, -1.421413741f));
  pl = __ffma2_rn(pl, t, make_float2(0.284496736f, 0.284496736f));
  pl = __ffma2_rn(pl, t, make_float2(-0.254829592f, -0.254829592f));
  pl = __fmul2_rn(pl, t);
  const float2 a = __fmul2_rn(__fmul2_rn(z, z), make_float2(-0.72134752044448170f, -0.72134752044448170f));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(a.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(a.y));
  float2 er = __ffma2_rn(pl, e, make_float2(1.f, 1.f));
  er.x = copysignf(er.x, z.x);
  er.y = copysignf(er.y, z.y);
  const float2 hz = __fmul2_rn(z, make_float2(0.5f, 0.5f));
  return __ffma2_rn(hz, er, hz);
}

__global__ void __launch_bounds__(kFfnThreads, 1) ffn_ln_gelu_kernel(const FfnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* stg = ring + kFfnStages * kFfnStage;                  // one split atom of output
  float2* xchg = reinterpret_cast<float2*>(stg + kSplitAtom);   // [2][128] peer CTA's (mean, M2) per buffer
  FfnBarriers* bar = reinterpret_cast<FfnBarriers*>(stg + kSplitAtom + kFfnXchgBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_atoms = p.a0_atoms + p.a1_atoms;
  const uint32_t rank = cluster_ctarank();
  const int first = blockIdx.x >> 1, stride = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023u) __trap();  // SWIZZLE_128B tiles need the 1 KB alignment the declaration asks for
    for (int s = 0; s < kFfnStages; ++s) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar->acc_full[i], 1);
      mbar_init(&bar->acc_empty[i], 256);
      mbar_init(&bar->xchg_full[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 9) tmem_alloc<512>(&bar->tmem_base);
  tc_fence_before();
  cluster_sync_all();  // the peer's barriers exist before the first remote store
  pdl_launch_dependents();
  pdl_wait();  // the activations of the previous kernel are complete and visible from here on
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp == 8) {
    // ------------------------------------------------------------------ producer
    if (lane == 0) {
      const uint8_t* gw = p.w + static_cast<size_t>(rank) * k_atoms * kFfnWAtom;
      uint32_t it = 0;
      for (int rt = first; rt < p.row_tiles; rt += stride) {
        for (int ka = 0; ka < k_atoms; ++ka, ++it) {
          const int s = it % kFfnStages;
          if (it >= kFfnStages) mbar_wait(&bar->empty[s], ((it / kFfnStages) - 1) & 1);
          const uint8_t* ga = ka < p.a0_atoms
                                  ? p.a0 + (static_cast<size_t>(rt) * p.a0_atoms + ka) * kSplitAtom
                                  : p.a1 + (static_cast<size_t>(rt) * p.a1_atoms + (ka - p.a0_atoms)) * kSplitAtom;
          mbar_arrive_expect_tx(&bar->full[s], kFfnStage);
          bulk_g2s(ring + s * kFfnStage, ga, kSplitAtom, &bar->full[s]);
          bulk_g2s(ring + s * kFfnStage + kSplitAtom, gw + static_cast<size_t>(ka) * kFfnWAtom, kFfnWAtom,
                   &bar->full[s]);
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------------ MMA issue (warp-uniform, one elected lane)
    const bool leader = elect_one();
    constexpr uint32_t idesc = umma_idesc(kFmtBF16, 128, kFfnNt);
    const uint32_t ring_a = smem_u32(ring);
    uint32_t it = 0;
    int local = 0;
    for (int rt = first; rt < p.row_tiles; rt += stride, ++local) {
      const int buf = local & 1;
      if (local >= 2) { mbar_wait(&bar->acc_empty[buf], ((local >> 1) - 1) & 1); tc_fence_after(); }
      const uint32_t acc = tmem + buf * 256;
      if (leader) FTRACE(0, local);
      for (int ka = 0; ka < k_atoms; ++ka, ++it) {
        const int s = it % kFfnStages;
        mbar_wait(&bar->full[s], (it / kFfnStages) & 1);
        tc_fence_after();
        const uint32_t a_hi = ring_a + s * kFfnStage, a_lo = a_hi + kHalfAtom;
        const uint32_t w_hi = a_hi + kSplitAtom, w_lo = w_hi + kFfnNt * 128;
        const uint64_t dah = umma_desc_kmajor_sw128(a_hi), dal = umma_desc_kmajor_sw128(a_lo);
        const uint64_t dwh = umma_desc_kmajor_sw128(w_hi), dwl = umma_desc_kmajor_sw128(w_lo);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {  // 64 bf16 per atom = 4 x K16, 32 B per step
          if (leader) umma_bf16_ss(acc, dal + 2 * kk, dwh + 2 * kk, idesc, (ka > 0 || kk > 0));
          if (leader) umma_bf16_ss(acc, dah + 2 * kk, dwl + 2 * kk, idesc, 1);
          if (leader) umma_bf16_ss(acc, dah + 2 * kk, dwh + 2 * kk, idesc, 1);
        }
        if (leader) umma_commit(&bar->empty[s]);
      }
      if (leader) umma_commit(&bar->acc_full[buf]);
      if (leader) FTRACE(1, local);
    }
  } else {
    // ------------------------------------------------------------------ epilogue: 2 threads per row
    const int q = warp & 3, g = warp >> 2;
    const int r = q * 32 + lane;
    const bool leader = threadIdx.x == 0;
    const float* bias = p.bias + rank * kFfnNt;
    const float* gamma = p.gamma + rank * kFfnNt;
    const float* beta = p.beta + rank * kFfnNt;
    uint32_t v[32];
    int local = 0;
    for (int rt = first; rt < p.row_tiles; rt += stride, ++local) {
      const int buf = local & 1;
      const uint32_t par = (local >> 1) & 1;
      const uint32_t tacc = tmem + (static_cast<uint32_t>(q * 32) << 16) + buf * 256;
      if (leader) mbar_arrive_expect_tx(&bar->xchg_full[buf], kTile * 8);
      if (leader) FTRACE(2, local);
      mbar_wait(&bar->acc_full[buf], par);
      tc_fence_after();
      if (leader) FTRACE(3, local);

      // ---- statistics: each of the two threads of a row reduces 128 of this CTA's 256 columns (shifted by its first
      //      element, packed f32x2 arithmetic), the halves meet through the idle staging buffer, the CTA's (mean, M2)
      //      goes to the peer
      const float* bias_h = bias + g * 128;
      float shift;
      float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);
      {
        uint32_t u[32];
        tmem_ld32(tacc + g * 128, v);
        tmem_wait_ld();
        shift = __uint_as_float(v[0]) + __ldg(bias_h);
        const float2 nshift = make_float2(-shift, -shift);
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
#pragma unroll
          for (int i = 0; i < 32; ++i) u[i] = v[i];
          if (cb < 3) tmem_ld32(tacc + g * 128 + (cb + 1) * 32, v);
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias_h + cb * 32) + i4);
            const float2 bs0 = __fadd2_rn(make_float2(b4.x, b4.y), nshift);
            const float2 bs1 = __fadd2_rn(make_float2(b4.z, b4.w), nshift);
            const float2 d0 = __fadd2_rn(make_float2(__uint_as_float(u[4 * i4]), __uint_as_float(u[4 * i4 + 1])), bs0);
            const float2 d1 =
                __fadd2_rn(make_float2(__uint_as_float(u[4 * i4 + 2]), __uint_as_float(u[4 * i4 + 3])), bs1);
            s1 = __fadd2_rn(s1, d0);
            s2 = __ffma2_rn(d0, d0, s2);
            s1 = __fadd2_rn(s1, d1);
            s2 = __ffma2_rn(d1, d1, s2);
          }
          if (cb < 3) tmem_wait_ld();
        }
      }
      tmem_ld32(tacc + g * 32, v);  // pass B's first atom, in flight under the exchanges
      const float s1h = s1.x + s1.y, s2h = s2.x + s2.y;
      const float dmh = s1h * (1.f / 128.f);
      const float mean_h = shift + dmh;
      const float m2_h = fmaxf(s2h - s1h * dmh, 0.f);
      float2* halves = reinterpret_cast<float2*>(stg);  // [2][128]
      if (leader) bulk_wait_read<0>();                  // the previous tile's last store has left the buffer
      epi_bar_sync_256();
      halves[g * kTile + r] = make_float2(mean_h, m2_h);
      epi_bar_sync_256();
      const float2 other = halves[(g ^ 1) * kTile + r];
      const float dh = mean_h - other.x;
      const float mean_c = 0.5f * (mean_h + other.x);
      const float m2_c = m2_h + other.y + dh * dh * 64.f;
      if (g == 0) {
        st_async_peer(&xchg[buf * kTile + r].x, mean_c, &bar->xchg_full[buf], rank ^ 1);
        st_async_peer(&xchg[buf * kTile + r].y, m2_c, &bar->xchg_full[buf], rank ^ 1);
      }
      if (leader) FTRACE(4, local);
      mbar_wait(&bar->xchg_full[buf], par);
      if (leader) FTRACE(5, local);
      const float2 peer = xchg[buf * kTile + r];
      const float mean = 0.5f * (mean_c + peer.x);
      const float dmm = mean_c - peer.x;
      const float var = (m2_c + peer.y + dmm * dmm * 128.f) * (1.f / 512.f);
      const float rstd = rsqrtf(var + p.eps);
      const float nmr = -mean * rstd;
      const float2 rstd2 = make_float2(rstd, rstd), nmr2 = make_float2(nmr, nmr);

      // ---- normalise, GELU, split, stage, store: 4 atoms of 64 columns; this thread's 32 columns of each
#pragma unroll 1
      for (int j = 0; j < 4; ++j) {
        uint32_t u[32];
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) u[i] = v[i];
        if (j < 3) {
          tmem_ld32(tacc + (j + 1) * 64 + g * 32, v);  // next atom's columns, in flight under this atom's math
        } else {
          tc_fence_before();
          mbar_arrive(&bar->acc_empty[buf]);
        }
        uint4 hi[4], lo[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int col = j * 64 + g * 32 + c * 8;
          float2 y[4];
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + col) + h2);
            const float4 g4 = __ldg(reinterpret_cast<const float4*>(gamma + col) + h2);
            const float4 t4 = __ldg(reinterpret_cast<const float4*>(beta + col) + h2);
            const uint32_t* uu = u + c * 8 + h2 * 4;
            const float2 x0 = __fadd2_rn(make_float2(__uint_as_float(uu[0]), __uint_as_float(uu[1])), make_float2(b4.x, b4.y));
            const float2 x1 = __fadd2_rn(make_float2(__uint_as_float(uu[2]), __uint_as_float(uu[3])), make_float2(b4.z, b4.w));
            const float2 z0 = __ffma2_rn(__ffma2_rn(x0, rstd2, nmr2), make_float2(g4.x, g4.y), make_float2(t4.x, t4.y));
            const float2 z1 = __ffma2_rn(__ffma2_rn(x1, rstd2, nmr2), make_float2(g4.z, g4.w), make_float2(t4.z, t4.w));
            y[h2 * 2] = gelu_fast2(z0);
            y[h2 * 2 + 1] = gelu_fast2(z1);
          }
          split8(y, hi[c], lo[c]);
        }
        if (leader) bulk_wait_read<0>();  // the previous atom's store has drained the staging buffer
        epi_bar_sync_256();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint8_t* at = stg + sw128_offset(r, g * 4 + c);
          *reinterpret_cast<uint4*>(at) = hi[c];
          *reinterpret_cast<uint4*>(at + kHalfAtom) = lo[c];
        }
        fence_proxy_async_smem();
        epi_bar_sync_256();
        if (leader) {
          bulk_s2g(p.out_tiles + (static_cast<size_t>(rt) * 8 + rank * 4 + j) * kSplitAtom, stg, kSplitAtom);
          bulk_commit();
          FTRACE(6 + j, local);
        }
      }
    }
    if (leader) bulk_wait_all();
  }
  tc_fence_before();
  cluster_sync_all();  // no CTA exits while its peer may still store into its shared memory
  if (warp == 9) tmem_dealloc<512>(tmem);
}

}  // namespace lg

static bool ffn_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int lg_ffn_ln_gelu(const void* a0, int a0_atoms, const void* a1, int a1_atoms, const void* w,
                              const float* bias, const float* gamma, const float* beta, float eps, void* out_tiles,
                              int row_tiles, cudaStream_t stream) {
  if (!a0 || !w || !bias || !gamma || !beta || !out_tiles) return LG_ERR_NULL;
  if (row_tiles <= 0 || a0_atoms <= 0 || a1_atoms < 0 || (a1_atoms > 0 && !a1)) return LG_ERR_SHAPE;
  if (!ffn_aligned16(a0) || !ffn_aligned16(a1) || !ffn_aligned16(w) || !ffn_aligned16(out_tiles) ||
      !ffn_aligned16(bias) || !ffn_aligned16(gamma) || !ffn_aligned16(beta))
    return LG_ERR_ALIGN;
  lg::FfnParams p{};
  p.a0 = static_cast<const uint8_t*>(a0);
  p.a1 = static_cast<const uint8_t*>(a1);
  p.a0_atoms = a0_atoms;
  p.a1_atoms = a1_atoms;
  p.w = static_cast<const uint8_t*>(w);
  p.bias = bias;
  p.gamma = gamma;
  p.beta = beta;
  p.eps = eps;
  p.out_tiles = static_cast<uint8_t*>(out_tiles);
  p.row_tiles = row_tiles;

  static bool attr_done_dev[lg::kMaxDevices] = {};  // per device; a benign race sets it twice
  bool& attr_done = attr_done_dev[lg::current_device_slot()];
  const int sms_dev = lg::device_sm_count();
  const int tpcs = sms_dev > 1 ? sms_dev / 2 : 74;
  if (!attr_done) {
    cudaError_t e =
        cudaFuncSetAttribute(lg::ffn_ln_gelu_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, lg::kFfnSmem);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(2 * (row_tiles < tpcs ? row_tiles : tpcs)));  // persistent: one pair per TPC
  cfg.blockDim = dim3(lg::kFfnThreads);
  cfg.dynamicSmemBytes = lg::kFfnSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = lg::add_pdl_attr(attr, 1);
  return static_cast<int>(cudaLaunchKernelEx(&cfg, lg::ffn_ln_gelu_kernel, p));
}

#ifdef LG_FFN_TRACE
extern "C" int lg_debug_ffn_trace(unsigned long long* host_out, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host_out, lg::g_ffn_trace, sizeof(unsigned long long) * n));
}
#endif
