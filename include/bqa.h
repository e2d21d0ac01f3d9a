/* bqa.h — C ABI of libbqa, the sm_100a implementation of the BERT question-answering
 * fine-tuning step of neuro-inc/ml-recipe-distributed-pytorch.
 *
 * The reference exposes no FFI: its seam is the Python surface of
 *   modules/model/model/model.py:13-73     BertForQuestionAnswering.forward
 *   modules/model/model/loss.py:74-106     WeightedLoss.__call__
 *   modules/model/trainer/trainer.py:197-225,266-300   backward / clip / optimizer step
 *   modules/model/inference/predictor.py:114-120       span decoding
 * and every FLOP behind that surface is executed by third-party libraries.
 * This header is what a binding of that surface onto hand-written kernels links against.
 * Each entry point cites the reference lines it replaces.
 *
 * Conventions: every function returns 0 on success, non-zero on failure with a message in
 * bqa_last_error() (thread-local).  All `void*` tensor arguments are DEVICE pointers unless
 * the name starts with `host_`.  `stream` is a cudaStream_t passed as void*.  The library
 * never takes ownership of caller memory and has no CPU fallback: without a usable CUDA
 * device every compute entry point fails.
 */
#ifndef BQA_H_
#define BQA_H_

#include <stdint.h>

#if defined(__GNUC__)
#define BQA_API __attribute__((visibility("default")))
#else
#define BQA_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

BQA_API const char* bqa_last_error(void);
/* ABI version of this header; bumped on any incompatible change. */
BQA_API int bqa_abi_version(void);

/* ------------------------------------------------------------------------------------
 * Dense layer contraction on tcgen05 tensor cores (bf16 in, fp32 accumulate).
 * Replaces torch.nn.Linear forward/backward as used by the encoder:
 *   site-packages/transformers/models/bert/modeling_bert.py:179-181 (Q/K/V), :294-298
 *   (attention output), :339-342 (intermediate + GELU), :352-356 (output), and their
 *   autograd backward driven by modules/model/trainer/trainer.py:197-204.
 * ------------------------------------------------------------------------------------ */
enum bqa_epilogue {
  BQA_EPI_LINEAR = 0,   /* C = dropout(acc + bias) + residual (each optional), bf16 out      */
  BQA_EPI_GELU_FWD = 1, /* C = acc + bias, C2 = gelu_erf(C), bf16 out                        */
  BQA_EPI_GELU_BWD = 2, /* C = acc * gelu_erf'(aux), bf16 out                                */
  BQA_EPI_F32_ADD = 3   /* C += acc, fp32 out, split-K allowed (weight gradients)            */
};

typedef struct bqa_gemm_desc {
  int32_t M, N, K;
  const void* A; /* bf16; K-major: [M][lda]; MN-major: [K][lda]                              */
  int32_t lda;
  int32_t a_mn_major;
  const void* B; /* bf16; K-major: [N][ldb]; MN-major: [K][ldb]                              */
  int32_t ldb;
  int32_t b_mn_major;
  void* C; /* [M][ldc], bf16 (fp32 for BQA_EPI_F32_ADD)                                      */
  int32_t ldc;
  void* C2; /* second output of BQA_EPI_GELU_FWD, else NULL                                  */
  int32_t epilogue;
  const float* bias;    /* [N] fp32 or NULL                                                  */
  const void* residual; /* bf16 [M][ld_res] or NULL                                          */
  int32_t ld_res;
  const void* aux; /* bf16 [M][ld_aux], pre-activation for BQA_EPI_GELU_BWD                   */
  int32_t ld_aux;
  int32_t k_splits;   /* >= 1                                                                */
  float dropout_p;    /* 0 disables; applied to (acc + bias) before the residual             */
  uint64_t drop_seed; /* with drop_site selects the dropout stream                           */
  uint32_t drop_site;
} bqa_gemm_desc;

BQA_API int bqa_gemm_bf16(const bqa_gemm_desc* desc, void* stream);

#ifdef __cplusplus
}
#endif

#endif /* BQA_H_ */
