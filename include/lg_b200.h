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

#define LG_TILE 128        /* rows per operand tile */
#define LG_TILE_BYTES 16384

/* Version of this ABI (major*100 + minor). */
int lg_version(void);

/* (1) normalize_keypoints (lightglue.py:21-33) + ConditionalLearnableFourierPositionalEncoding
 * (lightglue.py:50-77, fed at :490-494).  kpts (B,N,2) f32 pixels; size (B,2) f32 (w,h), required
 * (for the reference's size=None branch fill it with lg_keypoint_extent); wr (32,2); wc,bc (32) condition_modulation; cond =
 * padded keypoint count.  Writes the compact tables cos_t, sin_t (B,N,32) f32 (frequency j is
 * lanes 2j,2j+1 of the reference's (2,B,1,N,64) tensor). */
int lg_posenc(const float* kpts, const float* size, const float* wr, const float* wc, const float* bc,
              float cond, float* cos_t, float* sin_t, int B, int N, cudaStream_t stream);

/* size_out (B,2) f32 = 1 + max - min over each image's keypoints (lightglue.py:25-26). */
int lg_keypoint_extent(const float* kpts, float* size_out, int B, int N, cudaStream_t stream);

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

#ifdef __cplusplus
}
#endif
#endif /* LG_B200_H_ */
