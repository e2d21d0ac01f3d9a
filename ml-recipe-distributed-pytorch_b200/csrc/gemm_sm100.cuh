// gemm_sm100.cuh — host-side plan/launch interface of the tcgen05 GEMM (gemm_sm100.cu).
//
// One kernel family covers the three contractions of a Linear layer
// (site-packages/transformers/models/bert/modeling_bert.py:179-181,294-298,339-356 are the
// reference call sites; their backward is autograd's AddmmBackward):
//   forward  Y[M,N]  = X[M,K]  . W[N,K]^T      A K-major , B K-major
//   dgrad    dX[M,K'] = dY[M,N'] . W[N',K']     A K-major , B MN-major
//   wgrad    dW[N',K'] += dY[M,N']^T . X[M,K']  A MN-major, B MN-major, fp32 reduce-add, split-K
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "bqa_common.cuh"

namespace bqa {

enum GemmEpilogue : int {
  EPI_LINEAR = 0,    // C = dropout(acc + bias) + residual   (each part optional)   -> bf16
  EPI_GELU_FWD = 1,  // u = acc + bias: C = gelu_erf'(u), C2 = gelu_erf(u)             -> bf16 x2
  EPI_GELU_BWD = 2,  // C = acc * aux, aux = the C of EPI_GELU_FWD (gelu')            -> bf16
  EPI_F32_ADD = 3,   // C += acc  (TMA reduce-add, supports split-K)                 -> fp32
};

struct GemmProblem {
  int M, N, K;
  const void* A;  // bf16
  int lda;        // elements; K-major: A is [M][lda]; MN-major: A is [K][lda]
  int a_mn_major;
  const void* B;  // bf16
  int ldb;        // K-major: B is [N][ldb]; MN-major: B is [K][ldb]
  int b_mn_major;
  void* C;  // bf16 (fp32 for EPI_F32_ADD), [M][ldc]
  int ldc;
  void* C2;  // EPI_GELU_FWD second output, same shape/ld as C
  int epilogue;
  const float* bias;     // [N] fp32 or null
  const void* residual;  // bf16 [M][ld_res] or null   (EPI_LINEAR)
  int ld_res;
  const void* aux;  // bf16 [M][ld_aux], gelu'(u)       (EPI_GELU_BWD)
  int ld_aux;
  int k_splits;  // >= 1; > 1 only with EPI_F32_ADD; 0 = choose automatically (EPI_F32_ADD)
  int force_single;  // 1 = never use CTA pairs (tests)
  DropKey drop;  // thresh == 0 -> no dropout           (EPI_LINEAR)
  // EPI_LINEAR without residual: also emit, per row m and 64-column block j, the dot product of the
  // (bf16-rounded) output block with the same block of `dot_with` (bf16 [M][ld_dot]):
  //   dot_out[((m / dot_rows) * (N / 64) + j) * dot_rows + m % dot_rows]
  // — the attention backward's D = rowsum(dO * O) per (batch, head, query), with M = B * S, dot_rows = S.
  const void* dot_with;
  int ld_dot;
  float* dot_out;
  int dot_rows;
  // EPI_GELU_BWD: also accumulate the column sums of the output into colsum_out[N] (fp32 atomics) — the bias
  // gradient of the layer whose pre-activation gradient this GEMM produces.
  float* colsum_out;
};

struct GemmArgs {
  int M, N, K;
  int m_tiles, n_tiles, k_splits, kb_per_split, kb_total;
  const float* bias;
  const bf16* residual;
  int ld_res;
  const bf16* aux;
  int ld_aux;
  DropKey drop;
  const bf16* dot_with;
  int ld_dot;
  float* dot_out;
  int dot_rows;
  float* colsum_out;
};

struct GemmPlan {
  CUtensorMap tmA, tmB, tmC, tmC2;
  GemmArgs args;
  int variant;  // a_mn_major | b_mn_major << 1 | epilogue << 2
  int grid;
  int pair;  // 1 = CTA-pair kernel (cluster of 2)
};

// Returns 0 on success; on failure writes a message retrievable with bqa_last_error().
int gemm_plan(GemmPlan* plan, const GemmProblem& p, int num_sms);
int gemm_launch(const GemmPlan& plan, cudaStream_t stream);

// shared with attention: encode a tiled tensor map (bf16 / fp32, 128B swizzle)
int encode_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t cols, uint64_t rows,
                   uint64_t ld_elems, uint32_t box_cols, uint32_t box_rows);
int encode_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t d0, uint64_t d1, uint64_t d2,
                   uint64_t stride1_elems, uint64_t stride2_elems, uint32_t b0, uint32_t b1, uint32_t b2);

void set_error(const char* fmt, ...);

}  // namespace bqa
