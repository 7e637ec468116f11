// Pieces of the assignment path shared by lg_assign.cu (TF32x3 tiles) and lg_assign_strip.cu (split-bf16 strips).
#pragma once
#include <cuda_runtime.h>

namespace lg {
// merge the per-tile (max, sum exp) partials into row / column logsumexp, log-sigmoid terms and the dustbins
void launch_assign_reduce(const float2* row_part, const float2* col_part, const float* z0, const float* z1,
                          const int* m_len, const int* n_len, float* row_lse, float* col_lse, float* row_ls,
                          float* col_ls, float* scores, int B, int M, int N, int rt, int ct, cudaStream_t stream,
                          int plain = 0);
// pass 2 (values + filter keys) of the TMEM-resident variant (lg_assign_tmem.cu) on statistics some pass 1 + reduce
// has left in row_lse / col_lse / row_ls / col_ls; rowkey / colkey may be null
int launch_assign_tmem_pass2(const void* a_tiles, const void* b_tiles, int tpa, int tpb, const int* m_len,
                             const int* n_len, const float* row_lse, const float* col_lse, const float* row_ls,
                             const float* col_ls, float* scores, unsigned long long* rowkey,
                             unsigned long long* colkey, float scale, int B, int M, int N, cudaStream_t stream);
}  // namespace lg
