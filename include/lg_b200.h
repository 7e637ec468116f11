/* lg_b200.h -- C ABI of the B200-native LightGlue matcher forward (liblg_b200.so).
 *
 * Drop-in boundary for the hot path of the reference
 *   gluefactory/models/matchers/lightglue.py  (LightGlue.forward, :444-579)
 * The reference is pure Python/PyTorch and has no FFI of its own; these entry points are
 * what a ctypes binding for that path binds (INTEGRATION.md shows the stub).  Plain
 * pointers and sizes only -- no torch types.  All pointers are DEVICE pointers unless
 * stated; every call is asynchronous on `stream` and never synchronises.
 * Return value: 0 = ok, < 0 = argument error (LG_ERR_*), > 0 = cudaError_t.
 */
#ifndef LG_B200_H_
#define LG_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

#define LG_ERR_NULL (-1)   /* required pointer is null */
#define LG_ERR_SHAPE (-2)  /* unsupported / inconsistent sizes */
#define LG_ERR_ALIGN (-3)  /* pointer not 16-byte aligned */
#define LG_ERR_TOO_LARGE (-4)  /* problem exceeds the fixed capacity of a non-allocating entry point */

#define LG_TILE 128        /* rows per operand tile */
#define LG_TILE_BYTES 16384

/* Version of this ABI (major*100 + minor). */
int lg_version(void);

/* First 16 hex digits of the SHA-256 over the sources this library was compiled from (csrc/lg_*.cu, csrc/lg_*.cuh in
 * name order, then this header), set by the Makefile.  The Python binding recomputes it from the tree and refuses a
 * stale library.  No counterpart in the reference (pure Python). */
const char* lg_source_id(void);

/* (1) normalize_keypoints (lightglue.py:21-33) + ConditionalLearnableFourierPositionalEncoding
 * (lightglue.py:50-77, fed at :490-494).  kpts (B,N,2) f32 pixels; size (B,2) f32 (w,h), required
 * (for the reference's size=None branch fill it with lg_keypoint_extent); wr (32,2); wc,bc (32) condition_modulation; cond =
 * padded keypoint count.  Writes the compact tables cos_t, sin_t (B,N,32) f32 (frequency j is
 * lanes 2j,2j+1 of the reference's (2,B,1,N,64) tensor). */
int lg_posenc(const float* kpts, const float* size, const float* wr, const float* wc, const float* bc,
              float cond, float* cos_t, float* sin_t, int B, int N, cudaStream_t stream);

/* The same with add_scale_ori = True (lightglue.py:458-476): positional input [x, y, scale, ori], scales / oris (B,N)
 * f32, wr (32,4). */
int lg_posenc_scale_ori(const float* kpts, const float* scales, const float* oris, const float* size, const float* wr,
                        const float* wc, const float* bc, float cond, float* cos_t, float* sin_t, int B, int N,
                        cudaStream_t stream);

/* size_out (B,2) f32 = 1 + max - min over each image's keypoints (lightglue.py:25-26). */
int lg_keypoint_extent(const float* kpts, float* size_out, int B, int N, cudaStream_t stream);
/* the same over the first lens[b] points of image b (lens (B) int32 or NULL): padded batches of unequal images */
int lg_keypoint_extent_valid(const float* kpts, const int* lens, float* size_out, int B, int N, cudaStream_t stream);
/* lg_posenc / lg_posenc_scale_ori (scales, oris may be NULL) with ONE CONDITION PER IMAGE: cond_b (B) f32 replaces the
 * scalar keypoint count of lightglue.py:490-491, so that every image of a padded batch gets the encodings of the
 * reference's one-pair-per-call forward on that image alone (conf b200.condition = "valid"). */
int lg_posenc_per_image(const float* kpts, const float* scales, const float* oris, const float* size, const float* wr,
                        const float* wc, const float* bc, const float* cond_b, float* cos_t, float* sin_t, int B, int N,
                        cudaStream_t stream);

/* (1)+(2) SelfBlock operand producer (lightglue.py:185-188): de-interleave Wqkv output
 * qkv (B,N,H*192) [f = h*192 + d*3 + {q,k,v}], rotary on q,k (lightglue.py:36-43) from the compact
 * tables (NULL,NULL = no rotary), bf16 cast, pre-swizzled operand tiles [B][H][tiles_pad][16 KB].
 * len: per-image valid rows (NULL = N); rows beyond are zero.  tiles_pad even, tiles_pad*128 >= N. */
int lg_pack_self_qkv(const float* qkv, const float* cos_t, const float* sin_t, const int* len, void* q_tiles,
                     void* k_tiles, void* v_tiles, int B, int N, int H, int tiles_pad, cudaStream_t stream);

/* (2) CrossBlock operand producer (lightglue.py:223-228): qk, v (B,N,H*64) [f = h*64 + d]. */
int lg_pack_cross_qkv(const float* qk, const float* v, const int* len, void* qk_tiles, void* v_tiles, int B, int N,
                      int H, int tiles_pad, cudaStream_t stream);

/* (2) Attention.forward (lightglue.py:139-156): out (B,nq,H*64) f32 = softmax(q k^T * scale, keys >=
 * k_len masked) v ; rows >= q_len are written as 0 (reference: nan_to_num).  q_len/k_len NULL = full. */
int lg_attention(const void* q_tiles, const void* k_tiles, const void* v_tiles, float* out, const int* q_len,
                 const int* k_len, int B, int H, int nq, int nk, int tq_pad, int tk_pad, float softmax_scale,
                 cudaStream_t stream);

/* ---- split-tile activations and the per-token linears (lg_linear.cu) ----------------------------
 * "Split tiles": an activation matrix [rows][K] is held as [row tile of 128][K/64 atoms][hi 16 KB | lo 16 KB],
 * each 16 KB block the byte image of a SWIZZLE_128B K-major tile of bf16 (x = hi + lo, ~16 mantissa bits).
 * A batch of B images with N points each occupies B*tp row tiles (tp*128 >= N, rows >= N zero). */

/* One-off weight packing: w (n_out, k) f32 row-major (nn.Linear.weight) -> [n_out/nt][k/64][hi | lo][nt*128 B].
 * nt = 256 (192 for lg_linear_qkv_self); out holds n_out*k*4 bytes. */
int lg_pack_weight(const float* w, void* out, int n_out, int k, int nt, cudaStream_t stream);

/* x (B,N,K) f32 contiguous <-> split tiles [B*tp][K/64][32 KB]  (the .contiguous() copies of
 * lightglue.py:478-479 become this format change). */
int lg_rows_to_tiles(const float* x, void* tiles, int B, int N, int K, int tp, cudaStream_t stream);
int lg_tiles_to_rows(const void* tiles, float* x, int B, int N, int K, int tp, cudaStream_t stream);

/* Compact rotary tables (B,N,32) f32 from lg_posenc -> [B*tp][cos 16 KB | sin 16 KB] row tiles. */
int lg_rotary_tiles(const float* cos_t, const float* sin_t, void* rot_tiles, int B, int N, int tp,
                    cudaStream_t stream);

/* FFN middle (lightglue.py:173-174): h (rows,512) f32 -> GELU(LayerNorm(h; gamma, beta, eps)) as split
 * tiles [rows/128][8][32 KB]; rows is a multiple of 128. */
int lg_ln_gelu_tiles(const float* h, const float* gamma, const float* beta, float eps, void* tiles, int rows,
                     cudaStream_t stream);

/* Fused first half of the block MLP (lightglue.py:171-174 with :190-191 / :246-247 folded in):
 *     out = GELU(LayerNorm([a0 | a1] . w^T + bias; gamma, beta, eps))        512 outputs, LayerNorm over them
 * a0 / a1 split tiles as for lg_linear_*; w = lg_pack_weight(W (512, K), nt=256); out: split tiles
 * [row_tiles][8][32 KB].  One CTA pair per row tile (each CTA 256 of the 512 columns, row statistics exchanged
 * through distributed shared memory); equals lg_linear_f32 followed by lg_ln_gelu_tiles up to fp32 rounding. */
int lg_ffn_ln_gelu(const void* a0, int a0_atoms, const void* a1, int a1_atoms, const void* w, const float* bias,
                   const float* gamma, const float* beta, float eps, void* out_tiles, int row_tiles,
                   cudaStream_t stream);

/* y = [a0 | a1] . w^T + bias as a 3-term split-bf16 tcgen05 GEMM (nn.Linear, lightglue.py:171-176,190,246,310).
 * a0 / a1: split tiles with a0_atoms / a1_atoms K atoms (a1 NULL, 0 = none: torch.cat([x, msg], -1) of
 * lightglue.py:191,247 without the copy); a1_hi_only != 0: a1 holds plain bf16 tiles [row_tiles][a1_atoms][16 KB]
 * (x = hi exactly, e.g. an operand that is bf16 by construction) and contributes a 2-term product; w from lg_pack_weight(nt=256);
 * n_out multiple of 256.
 * lg_linear_split: out (and optional residual, added in fp32: x + ffn(...)) are split tiles
 *   [row_tiles][n_out/64][32 KB].   lg_linear_f32: out f32 rows with pitch ld floats. */
int lg_linear_split(const void* a0, int a0_atoms, const void* a1, int a1_atoms, int a1_hi_only, const void* w, const float* bias,
                    const void* resid_tiles, void* out_tiles, int row_tiles, int n_out, cudaStream_t stream);
int lg_linear_f32(const void* a0, int a0_atoms, const void* a1, int a1_atoms, int a1_hi_only, const void* w, const float* bias,
                  float* out, int64_t ld, int row_tiles, int n_out, cudaStream_t stream);
/* the same with the output as plain fp16 tiles [row_tiles][n_out/64][16 KB] (values clamped to the fp16 range): the
 * operand format of lg_assign_tiles (final_proj, lightglue.py:310) */
int lg_linear_f16(const void* a0, int a0_atoms, const void* a1, int a1_atoms, int a1_hi_only, const void* w, const float* bias,
                  void* out_tiles, int row_tiles, int n_out, cudaStream_t stream);

/* SelfBlock Wqkv + head split + rotary (lightglue.py:184-188) -> lg_attention operand tiles.  a: split tiles
 * of x, K = 256, rows = image-0 batch (B*tp0 tiles) then image-1 batch (B*tp1 tiles; tp1 may be 0).
 * w: Wqkv rows permuted to [head][q 64 | k 64 | v 64] (row h*192 + d*3 + part of the reference ->
 * h*192 + part*64 + d), packed with nt = 192; bias permuted alike.  q/k/v tiles per side: [B][H][tp][16 KB]. */
int lg_linear_qkv_self(const void* a, const void* w, const float* bias, const void* rot_tiles, void* q0, void* k0,
                       void* v0, void* q1, void* k1, void* v1, int B, int H, int tp0, int tp1, cudaStream_t stream);

/* CrossBlock to_qk | to_v (lightglue.py:223-228): w = rows [to_qk (256) ; to_v (256)], nt = 256. */
int lg_linear_qkv_cross(const void* a, const void* w, const float* bias, void* qk0, void* v0, void* qk1, void* v1,
                        int B, int H, int tp0, int tp1, cudaStream_t stream);

/* lg_attention writing the context as split tiles (K atom = head) for lg_linear: row tile of
 * (image b, query tile t) = out_rt_base + b*tq_pad + t. */
int lg_attention_tiles(const void* q_tiles, const void* k_tiles, const void* v_tiles, void* out_tiles,
                       int out_rt_base, const int* q_len, const int* k_len, int B, int H, int nq, int nk, int tq_pad,
                       int tk_pad, float softmax_scale, cudaStream_t stream);

/* Two attention problems of one block in ONE launch: both images of a SelfBlock (lightglue.py:268-269; side 0 =
 * image 0 on itself, side 1 = image 1), or both directions of a CrossBlock (:229-244; side 0 = queries of image 0
 * over keys / values of image 1, side 1 = the reverse).  Arguments per side as in lg_attention_tiles. */
int lg_attention_tiles2(const void* q0, const void* k0, const void* v0, int rt_base0, const int* q_len0,
                        const int* k_len0, int nq0, int nk0, int tq_pad0, int tk_pad0, const void* q1, const void* k1,
                        const void* v1, int rt_base1, const int* q_len1, const int* k_len1, int nq1, int nk1,
                        int tq_pad1, int tk_pad1, void* out_tiles, int B, int H, float softmax_scale,
                        cudaStream_t stream);

/* (3) MatchAssignment.forward from the similarity on (lightglue.py:310-315) + sigmoid_log_double_softmax
 * (lightglue.py:284-296).  md0 (B,M,256), md1 (B,N,256) f32 = final_proj outputs; prescale multiplies
 * both (d^-1/4 = 0.25, lightglue.py:310); z0 (B,M), z1 (B,N) f32 matchability logits.  Writes scores
 * (B,M+1,N+1) f32 contiguous.  Similarity GEMM: TF32x3 on tcgen05, fp32 accumulate.  Padded entries
 * of a variable-count pair (row >= m_len or col >= n_len) are -inf; the dustbin row/column stay at
 * index M / N.  ws: lg_assign_ws_bytes(B,M,N) bytes of scratch. */
size_t lg_assign_ws_bytes(int B, int M, int N);
int lg_assign(const float* md0, const float* md1, const float* z0, const float* z1, const int* m_len,
              const int* n_len, float prescale, float* scores, void* ws, int B, int M, int N, cudaStream_t stream);

/* (4) filter_matches (lightglue.py:321-337) on scores (B, M+1, N+1) f32, row pitch `pitch` floats
 * (>= N+1), image pitch (M+1)*pitch.  m_len/n_len: per-pair valid M/N (NULL = M,N).
 * Outputs m0 (B,M) i64, m1 (B,N) i64, ms0 (B,M) f32, ms1 (B,N) f32.  ws: lg_filter_ws_bytes(B,M,N)
 * bytes of scratch. */
size_t lg_filter_ws_bytes(int B, int M, int N);
int lg_filter_matches(const float* scores, int64_t pitch, const int* m_len, const int* n_len, float th, int64_t* m0,
                      int64_t* m1, float* ms0, float* ms1, void* ws, int B, int M, int N, cudaStream_t stream);

/* lg_filter_matches without a workspace argument and without any allocation (no cudaMalloc / cudaFree, no allocating
 * library routine): the key scratch is a file-scope `static __device__` array of LG_FILTER_STATIC_MAX_KEYS 64-bit
 * keys (1 MiB, held by every GPU that loads the library until exit), for callers that cannot allocate -- CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Same outputs as lg_filter_matches (bit-identical).  Limit: B*(M+N) <=
 * LG_FILTER_STATIC_MAX_KEYS, else LG_ERR_TOO_LARGE.  Every call on a device shares the one array: it is cleared on
 * `stream` at the start of each call, and calls that overlap on the GPU (different streams) race on it.  The first
 * call on a device resolves the array's address (cudaGetSymbolAddress) -- make it outside a graph capture. */
#define LG_FILTER_STATIC_MAX_KEYS 131072
int lg_filter_matches_static(const float* scores, int64_t pitch, const int* m_len, const int* n_len, float th,
                             int64_t* m0, int64_t* m1, float* ms0, float* ms1, int B, int M, int N,
                             cudaStream_t stream);

/* (3)+(4) fused: lg_assign that also gathers, while the scores are still on chip, the best column of every row
 * and the best row of every column of the valid block as order-preserving keys (largest score, lowest index on
 * ties -- torch.max's choice, lightglue.py:323-324) into keys_ws (lg_filter_ws_bytes(B,M,N) bytes, zeroed by the
 * call); lg_filter_from_keys then finishes filter_matches (mutual check, exp, threshold :325-336) without reading
 * the (M+1)x(N+1) matrix again.  Results are bit-identical to lg_assign + lg_filter_matches. */
int lg_assign_keys(const float* md0, const float* md1, const float* z0, const float* z1, const int* m_len,
                   const int* n_len, float prescale, float* scores, void* ws, void* keys_ws, int B, int M, int N,
                   cudaStream_t stream);
int lg_filter_from_keys(const void* keys_ws, const int* m_len, const int* n_len, float th, int64_t* m0, int64_t* m1,
                        float* ms0, float* ms1, int B, int M, int N, cudaStream_t stream);

/* (3)+(4) from fp16 tiles, the matcher's path: MatchAssignment (lightglue.py:306-315) + sigmoid_log_double_softmax
 * (:284-296) (+ filter keys when keys_ws != NULL, lg_filter_ws_bytes(B,M,N) bytes, zeroed by the call).  Operands are
 * the fp16 tiles lg_linear_f16 writes for final_proj ([row tile][4 atoms][16 KB]; image b of side 0 starts at row
 * tile b*tpa of a_tiles, side 1 at b*tpb of b_tiles): no fp32 rows, no pack pass, ONE tcgen05 kind::f16 product per
 * K step (11-bit operands: ~1e-3 on log_assignment inside the bf16 path's 2e-2 bound; lg_assign keeps TF32x3 for the
 * fp32 path).  similarity = scale * <a_row, b_row> (scale = d^-1/2 for the reference's two d^-1/4 factors,
 * lightglue.py:311-312).  z0 (B,M), z1 (B,N) = matchability logits (lg_logit_rows).  scores: (B, M+1, pitch) f32,
 * row pitch in floats >= N+1 and a multiple of 4 (16-byte vector stores; a multiple of 32 makes every stored 64-byte
 * segment two whole sectors); columns N+1 .. pitch-1 are never written.  ws: lg_assign_tiles_ws_bytes(B,M,N) bytes. */
size_t lg_assign_tiles_ws_bytes(int B, int M, int N);
int lg_assign_tiles(const void* a_tiles, const void* b_tiles, int tpa, int tpb, const float* z0, const float* z1,
                    const int* m_len, const int* n_len, float scale, float* scores, int64_t pitch, void* ws,
                    void* keys_ws, int B, int M, int N, cudaStream_t stream);

/* Linear(256 -> 1) logits of every row: out[i] = <x[i,:], w> + bias[0]  (MatchAssignment.matchability,
 * lightglue.py:304,313-314, before the log-sigmoid).  x (rows, 256) f32 contiguous. */
int lg_logit_rows(const float* x, const float* w, const float* bias, float* out, int64_t rows, cudaStream_t stream);

/* NearestNeighborMatcher (reference gluefactory/models/matchers/nearest_neighbor_matcher.py:52-72).
 * lg_nn_assign: similarity = d0 . d1^T (TF32x3 on tcgen05, the lg_assign tile kernels) written to sim_out (B,M,N)
 *   and log_assignment (B,M+1,N+1) = log_softmax(sim,-1) + log_softmax(sim,-2) with zero dustbin row / column (:63-64).
 *   d0 (B,M,256), d1 (B,N,256) f32; ws: lg_assign_ws_bytes(B,M,N).
 * lg_nn_match: find_nn (:16-27) along rows and columns of sim (top-2, dist = 2(1-sim), ratio / distance tests; a
 *   threshold <= 0 disables its test like the reference's None) and mutual_check (:30-39) when `mutual`;
 *   m0 (B,M) / m1 (B,N) i64 (-1 = unmatched), ms0 / ms1 = (m > -1) as f32 (:65-66).  ws: lg_nn_ws_bytes(B,M,N). */
int lg_nn_assign(const float* d0, const float* d1, const int* m_len, const int* n_len, float* sim_out, float* scores,
                 void* ws, int B, int M, int N, cudaStream_t stream);
size_t lg_nn_ws_bytes(int B, int M, int N);
int lg_nn_match(const float* sim, const int* m_len, const int* n_len, float ratio_thresh, float distance_thresh,
                int mutual, int64_t* m0, int64_t* m1, float* ms0, float* ms1, void* ws, int B, int M, int N,
                cudaStream_t stream);

/* fp32 path (b200.precision = "fp32"): IEEE fp32 FMA arithmetic on the CUDA cores for the 1e-4 bound on
 * log_assignment (the tensor-core path is bounded by 2e-2).
 * lg_gemm_f32: C = alpha * A . op(B) + bias + resid, strided-batched.  A (M,K) lda; B (N,K) ldb if b_kn == 0
 *   (Linear layers lightglue.py:168-176,184,203-211,223-228,310,370; q.k^T :151,235) or (K,N) ldb if b_kn == 1
 *   (P.v :155,241-244); bias (N) / resid (M,N, ldr; may alias C) optional.  Batch index z in [0, batch) maps to
 *   (outer, inner) = (z / inner, z % inner); element offsets outer*strides[2i] + inner*strides[2i+1] for
 *   i = A, B, C, resid (strides: 8 host int64).  m_len / k_len: optional valid M / K per OUTER index.
 * lg_softmax_rows_f32: in-place softmax over the first n_len[b] columns of s (B,H,M,pitch) (others -> 0; rows
 *   >= m_len[b] -> 0 like the reference's nan_to_num of fully masked rows, :145,149).
 * lg_rotary_rows_f32: apply_cached_rotary_emb (:36-43) in place on the q block (columns q_off + h*64 ..) and the
 *   k block (k_off + h*64 ..) of x (rows, ld) with the compact (rows, 32) cos / sin tables of lg_posenc.
 * lg_ln_gelu_rows_f32: LayerNorm(width, eps, affine) + GELU(erf) in place on x (rows, width) (:173-174). */
int lg_gemm_f32(const float* A, const float* B, float* C, const float* bias, const float* resid, int M, int N, int K,
                int64_t lda, int64_t ldb, int64_t ldc, int64_t ldr, int b_kn, float alpha, int batch, int inner,
                const int64_t* strides, const int* m_len, const int* k_len, cudaStream_t stream);
int lg_softmax_rows_f32(float* s, const int* m_len, const int* n_len, int B, int H, int M, int N, int64_t pitch,
                        cudaStream_t stream);
int lg_rotary_rows_f32(float* x, const float* cos_t, const float* sin_t, int64_t rows, int64_t ld, int q_off, int k_off,
                       int H, cudaStream_t stream);
int lg_ln_gelu_rows_f32(float* x, const float* gamma, const float* beta, float eps, int64_t rows, int width,
                        cudaStream_t stream);

/* (5) adaptive depth / width (lightglue.py:527-547).
 * lg_token_scores: conf (B,N) = sigmoid(w_tok.desc + b_tok) (TokenConfidence.forward :102-107; w_tok NULL
 *   => skipped) and match (B,N) = sigmoid(w_match.desc + b_match) (get_matchability :317-318), one read
 *   of desc (B,N,256).
 * lg_prune_plan: keep = match > min_match | conf <= thr (get_pruning_mask :586-593; conf NULL => first
 *   term only); keep_idx (B,N) i32 = ascending kept positions (== torch.where(mask)[1]); counts (B,2) i32
 *   = {kept, #(conf < thr)} (the count check_if_stop :595-606 needs).
 * lg_prune_gather: index_select of desc / cos / sin / ind by keep_idx into buffers of N_out rows per
 *   image, and prune[b, ind] += 1 (:537-540); prune is (B, N_orig) i64. */
int lg_token_scores(const float* desc, const float* w_tok, const float* b_tok, const float* w_match,
                    const float* b_match, const int* len, float* conf, float* match, int B, int N,
                    cudaStream_t stream);
int lg_prune_plan(const float* conf, const float* match, const int* len, float thr, float min_match, int* keep_idx,
                  int* counts, int B, int N, cudaStream_t stream);
int lg_prune_gather(const float* desc, const float* cos_t, const float* sin_t, const int64_t* ind,
                    const int* keep_idx, const int* counts, float* desc_o, float* cos_o, float* sin_o, int64_t* ind_o,
                    int64_t* prune, int B, int N, int N_out, int N_orig, cudaStream_t stream);

/* (5) fused, on the engine's split tiles: the adaptive step of one layer for a whole batch in ONE launch --
 * token confidence + matchability (lightglue.py:102-107, :317-318), keep mask (:586-593), the count check_if_stop
 * needs (:595-606), exclusive scan and the order-preserving compaction (:536-547) of descriptor rows (x_tiles ->
 * x_out), rotary rows (rot_tiles -> rot_out) and original indices (ind -> ind_out; ind NULL = identity), with
 * prune_out = prune + 1 on the kept points (prune itself untouched: a pair that exits at this layer keeps its old
 * counters, the reference breaks before pruning).  Tiles: side 0 images at row tile b*tp0, side 1 at B*tp0 + b*tp1.
 * counts (B,4) i32 = {kept0, #(conf0 < thr), kept1, #(conf1 < thr)}; len_out = kept.  do_prune = 0: scores and
 * counts only (depth-only mode; the *_out / rot / prune arguments may then be NULL).  w_tok NULL: no confidence head.
 * Rows of x_out / rot_out behind the kept ones are zero-filled up to tp*128.  One CTA per row tile of the batch; ws =
 * lg_score_prune_ws_bytes(B, tp0, tp1) bytes of scratch, 8-byte aligned (cleared by the call: a ticket and one word
 * per row tile for the prefix exchange between the tiles of an image); prune_out is filled by a device-to-device copy
 * of prune on the same stream before the kernel runs. */
size_t lg_score_prune_ws_bytes(int B, int tp0, int tp1);
int lg_score_prune_tiles(const void* x_tiles, const void* rot_tiles, void* x_out, void* rot_out, const int64_t* ind0,
                         const int64_t* ind1, int64_t* ind0_out, int64_t* ind1_out, const int64_t* prune0,
                         const int64_t* prune1, int64_t* prune0_out, int64_t* prune1_out, const int* len0,
                         const int* len1, int* len0_out, int* len1_out, int* counts, const float* w_tok,
                         const float* b_tok, const float* w_match, const float* b_match, float thr, float min_match,
                         int do_prune, void* ws, int B, int n0, int n1, int tp0, int tp1, int n0_orig, int n1_orig,
                         cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LG_B200_H_ */
