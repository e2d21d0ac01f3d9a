// kernels.cuh — launch wrappers of every non-GEMM kernel in libbqa (host side).
// Each returns 0 on success; on failure the message is in bqa::last_error().
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "bqa_common.cuh"
#include "bqa_host.cuh"

namespace bqa {

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

// Launch with programmatic stream serialization: the grid may be scheduled while its predecessor in the stream is
// still draining (launch latency and the kernel's own prologue overlap the predecessor's tail).  The kernel must call
// pdl_wait() before it touches anything the predecessor may read or write, and pdl_trigger() only after that wait
// (so that completion stays transitive along a chain of such launches).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- rowwise.cu -------------------------------------------------------------------------
int launch_ln_fwd(const bf16* z, const float* gamma, const float* beta, bf16* y, int T, int H, float eps,
                  cudaStream_t st);
// ids_kept / tt_kept / mask_kept (each may be null): range-checked private copies of the inputs for the backward
int launch_embed_ln_fwd(const int64_t* ids, const int64_t* token_type, const uint8_t* key_mask, const float* word,
                        const float* pos, const float* type, const float* gamma, const float* beta, bf16* z0,
                        bf16* x0, int64_t* ids_kept, int64_t* tt_kept, uint8_t* mask_kept, int32_t* pos_kept, int T,
                        int S, int H, int V, int TV, int P, int pos_mode, int pad_id, float eps, DropKey drop,
                        int* err_flag, cudaStream_t st);
int launch_ln_bwd(const bf16* dy, const bf16* z, const float* gamma, bf16* dz, bf16* dz_masked, float* dgamma,
                  float* dbeta, float* dbias, int T, int H, float eps, DropKey drop_in, DropKey drop_out,
                  int num_sms, cudaStream_t st);
int launch_colsum(const bf16* X, int ld, float* out, int T, int N, int num_sms, cudaStream_t st);
int launch_prepare_inputs(const int32_t* ids32, int64_t* ids, int64_t* token_type, uint8_t* mask, int B, int S,
                          int sep_id, int type_coef, int* missing_sep, cudaStream_t st);
int launch_embed_bwd(const bf16* dz0, const int64_t* ids, const int64_t* token_type, const int32_t* pos_ids,
                     float* dword, float* dpos, float* dtype, int T, int S, int H, int V, int TV, int P, int pos_mode,
                     int pad_id, cudaStream_t st);

// ---- heads_loss.cu ----------------------------------------------------------------------
struct HeadParams {  // fp32 master weights of the pooler and the four heads (+ their gradients)
  const float *pool_w, *pool_b;    // [H,H], [H]
  const float *span_w, *span_b;    // [2,H], [2]
  const float *cls_w, *cls_b;      // [C,H], [C]
  const float *regs_w, *regs_b;    // [1,H], [1]
  const float *rege_w, *rege_b;    // [1,H], [1]
};
struct HeadGrads {
  float *pool_w, *pool_b, *span_w, *span_b, *cls_w, *cls_b, *regs_w, *regs_b, *rege_w, *rege_b;
};

int launch_span_head_fwd(const bf16* y, const float* span_w, const float* span_b, float* start, float* end, int T,
                         int H, cudaStream_t st);
// reg_s2 / reg_e2 (may be null): second copy of the two sigmoid outputs (kept for the backward)
int launch_pooler_heads_fwd(const bf16* y, const HeadParams& hp, float* pooled, float* cls, float* reg_s,
                            float* reg_e, float* reg_s2, float* reg_e2, int B, int S, int H, int C, DropKey cls_drop,
                            cudaStream_t st);

enum ClsLossKind : int { CLS_LOSS_SMOOTH = 0, CLS_LOSS_CE = 1, CLS_LOSS_FOCAL = 2 };
struct LossConfig {
  float w_start, w_end, w_start_reg, w_end_reg, w_cls;
  int cls_kind;
  float smooth_alpha;
  const float* cls_weight;  // [C] fp32 or null (CrossEntropyLoss(weight=...))
  float grad_scale;         // 1 / batch_split (trainer.py:198)
  float focal_alpha, focal_gamma;  // FocalLossWithLogits (loss.py:57-71)
};
// losses_out[6] = {start_class, end_class, start_reg, end_reg, cls, total}; all NaN when *err_flag != 0 (the
// forward saw an out-of-range token / type id, where the reference raises IndexError)
// d_reg_s / d_reg_e are gradients w.r.t. the sigmoid OUTPUTS (what autograd would hand back)
int launch_loss_fwd_bwd(const float* start, const float* end, const float* cls, const float* reg_s,
                        const float* reg_e, const int64_t* y_start, const int64_t* y_end, const int64_t* y_cls,
                        const float* y_reg_s, const float* y_reg_e, float* d_start, float* d_end, float* d_cls,
                        float* d_reg_s, float* d_reg_e, float* row_scratch, float* losses_out, int B, int S,
                        int C, const LossConfig& lc, int with_grad, cudaStream_t st, const int* err_flag = nullptr);
int launch_heads_bwd(const bf16* y, const float* pooled, const float* reg_s, const float* reg_e,
                     const float* d_start, const float* d_end, const float* d_cls, const float* d_reg_s,
                     const float* d_reg_e, const HeadParams& hp, const HeadGrads& hg, float* dpre_scratch,
                     float* dx0_scratch, bf16* dy, int B, int S, int H, int C, DropKey cls_drop, int num_sms,
                     cudaStream_t st);
int launch_decode(const float* start, const float* end, const float* cls, int64_t* start_idx, int64_t* end_idx,
                  int64_t* cls_idx, float* start_max, float* end_max, float* cls_prob, float* score, int B, int S,
                  int C, cudaStream_t st);

// ---- optimizer.cu -----------------------------------------------------------------------
// partial[blocks] <- per-block sum of squares over g[0..n) ; accumulate into *sumsq (fp32 atomics
// are avoided: a second tiny kernel adds the partials in a fixed order).
int launch_grad_sumsq(const float* g, size_t n, double* sumsq_accum, float* partial_scratch, int num_sms,
                      cudaStream_t st);
struct AdamWChunk {  // one contiguous run of elements of one tensor (same layout as bqa_adamw_chunk)
  uint64_t begin, end;
  float weight_decay;
  int32_t group;
};
struct LrTable {
  float lr[8];
};
int launch_clip_adamw(float* p, float* g, float* m, float* v, bf16* p_bf16, const AdamWChunk* chunks_dev,
                      int n_chunks, const double* sumsq, float max_grad_norm, LrTable lrs, float beta1,
                      float beta2, float eps, int zero_grad, cudaStream_t st);
int sumsq_scratch_floats(int num_sms);
int launch_clip_adamod(float* p, float* g, float* m, float* v, float* s, bf16* p_bf16, const AdamWChunk* chunks_dev,
                       int n_chunks, const double* sumsq, float max_grad_norm, LrTable lrs, float beta1,
                       float beta2, float beta3, float eps, float bias_mul, int zero_grad, cudaStream_t st);
int launch_f32_to_bf16(const float* src, bf16* dst, size_t n, int num_sms, cudaStream_t st);

// ---- attention_sm100.cu ------------------------------------------------------------------
struct AttnPlan;  // opaque here; see attention_sm100.cuh

}  // namespace bqa
