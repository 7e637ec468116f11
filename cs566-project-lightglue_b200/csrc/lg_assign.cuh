// Pieces of the TF32x3 assignment path (lg_assign.cu) that lg_nn.cu shares.
#pragma once
#include <cuda_runtime.h>

namespace lg {
// merge the per-tile (max, sum exp) partials into row / column logsumexp, log-sigmoid terms and the dustbins
void launch_assign_reduce(const float2* row_part, const float2* col_part, const float* z0, const float* z1,
                          const int* m_len, const int* n_len, float* row_lse, float* col_lse, float* row_ls,
                          float* col_ls, float* scores, int B, int M, int N, int rt, int ct, cudaStream_t stream,
                          int plain = 0);
}  // namespace lg
