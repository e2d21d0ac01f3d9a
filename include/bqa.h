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
 * bqa_last_error() (thread-local).  All pointer arguments are DEVICE pointers unless the
 * parameter is a struct passed by the host or its name starts with `host_`.  `stream` is a
 * cudaStream_t passed as void*.  The library never takes ownership of caller memory, keeps
 * bound pointers only between bqa_bind() and bqa_destroy(), allocates no device memory of
 * its own, and has no CPU fallback: without a usable CUDA device every compute entry point
 * fails.  One handle is driven by one host thread at a time.
 */
#ifndef BQA_H_
#define BQA_H_

#include <stddef.h>
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
  BQA_EPI_GELU_FWD = 1, /* u = acc + bias: C = gelu_erf'(u), C2 = gelu_erf(u), bf16 out         */
  BQA_EPI_GELU_BWD = 2, /* C = acc * aux with aux = C of BQA_EPI_GELU_FWD (gelu'), bf16 out      */
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
  const void* aux; /* bf16 [M][ld_aux], gelu' values for BQA_EPI_GELU_BWD                     */
  int32_t ld_aux;
  int32_t k_splits;   /* >= 1                                                                */
  float dropout_p;    /* 0 disables; applied to (acc + bias) before the residual             */
  uint64_t drop_seed; /* with drop_site selects the dropout stream                           */
  uint32_t drop_site;
  /* Fused side outputs of the backward GEMMs (all optional, NULL disables; ABI version 3).
   * colsum_out (BQA_EPI_GELU_BWD): [N] fp32, += column sums of C (the bias gradient of the
   *   intermediate projection, modeling_bert.py:339-342, without a pass over C).
   * dot_with / dot_out (BQA_EPI_LINEAR, N a multiple of 64): per row m and 64-column block j,
   *   dot_out[((m / dot_rows) * (N / 64) + j) * dot_rows + m % dot_rows] = sum of C[m][64j..64j+63]
   *   * dot_with[m][64j..64j+63] (dot_with bf16 [M][ld_dot]) — the attention backward's
   *   D = rowsum(dO * O) per (batch, head, query) with M = B * S, dot_rows = S, written (not added). */
  float* colsum_out;
  const void* dot_with;
  int32_t ld_dot;
  float* dot_out;
  int32_t dot_rows;
} bqa_gemm_desc;

BQA_API int bqa_gemm_bf16(const bqa_gemm_desc* desc, void* stream);

/* ------------------------------------------------------------------------------------
 * Fused self-attention (head dim 64, S <= 512).  Replaces BertSelfAttention's
 * softmax(QK^T/8 + mask) -> dropout -> PV (modeling_bert.py:115-140,168-207) and its backward.
 *   qkv  [B*S][3H] bf16 (Q | K | V slabs), ctx [B*S][H] bf16, lse2 [B][A][S] fp32
 *   key_mask [B][S] uint8 (1 = attend) or NULL
 *   keep_bits [B][A][S][bqa_attention_keep_words(S)] uint32: dropout decisions written by the
 *   forward and read by the backward; opaque to the caller (per position r: the decisions of
 *   query r over the keys, then those of key r over the queries); may be NULL when dropout_p == 0
 *   backward: dctx [B*S][H] bf16 in, dqkv [B*S][3H] bf16 out, dsum [B][A][S] fp32 scratch;
 *   dqkv_bias_grad: [3H] fp32 or NULL, += column sums of dqkv (gradient of the fused Q/K/V bias)
 * ------------------------------------------------------------------------------------ */
BQA_API int bqa_attention_keep_words(int32_t S);
BQA_API int bqa_attention_forward(const void* qkv, void* ctx, float* lse2, uint32_t* keep_bits,
                                  const uint8_t* key_mask, int32_t B, int32_t S, int32_t A, float dropout_p,
                                  uint64_t drop_seed, uint32_t drop_site, void* stream);
BQA_API int bqa_attention_backward(const void* qkv, const void* ctx, const float* lse2, const uint32_t* keep_bits,
                                   const void* dctx, void* dqkv, float* dsum, const uint8_t* key_mask, int32_t B,
                                   int32_t S, int32_t A, float dropout_p, uint64_t drop_seed, uint32_t drop_site,
                                   float* dqkv_bias_grad, void* stream);

/* ------------------------------------------------------------------------------------
 * Row kernels (HBM-bound).  LayerNorm(eps) over H of bf16 rows with fp32 affine parameters
 * (modeling_bert.py:111,297,355) and its backward; column sums for bias gradients.
 * ------------------------------------------------------------------------------------ */
BQA_API int bqa_layernorm_forward(const void* z, const float* gamma, const float* beta, void* y, int32_t T,
                                  int32_t H, float eps, void* stream);
/* dz_masked/dbias may be NULL.  dgamma/dbeta/dbias are accumulated into (+=). */
BQA_API int bqa_layernorm_backward(const void* dy, const void* z, const float* gamma, void* dz, void* dz_masked,
                                   float* dgamma, float* dbeta, float* dbias, int32_t T, int32_t H, float eps,
                                   float drop_in_p, uint32_t drop_in_site, float drop_out_p,
                                   uint32_t drop_out_site, uint64_t drop_seed, void* stream);
BQA_API int bqa_colsum_bf16(const void* X, int32_t ld, float* out, int32_t T, int32_t N, void* stream);

/* ------------------------------------------------------------------------------------
 * Optimizer: global-norm clip + HF AdamW(correct_bias=False) over flat fp32 buffers.
 * Replaces torch.nn.utils.clip_grad_norm_ (trainer.py:221-225) and transformers.AdamW.step
 * as constructed in modules/init.py:125-138 (two param groups, decoupled decay after the
 * update, eps 1e-6, no bias correction) plus optimizer.zero_grad() (trainer.py:287).
 * ------------------------------------------------------------------------------------ */
typedef struct bqa_adamw_chunk { /* one run of elements of one parameter tensor */
  uint64_t begin, end;           /* element range in the flat buffers, multiples of 4      */
  float weight_decay;
  int32_t group; /* index into the per-call lr[] array                                     */
} bqa_adamw_chunk;

/* *sumsq (device double) += sum(g[0..n)^2).  scratch: >= bqa_sumsq_scratch_floats() floats. */
BQA_API int bqa_sumsq_scratch_floats(void);
BQA_API int bqa_grad_sumsq(const float* g, uint64_t n, double* sumsq, float* scratch, void* stream);
/* No counterpart in the reference.  Same result as bqa_grad_sumsq (*sumsq += sum(g[0..n)^2)) for
 * callers that have no scratch buffer to give and may not create one: processes that pre-reserve all
 * device memory in their own pool and forbid cudaMalloc afterwards, plugins without access to the
 * host application's allocator, latency-critical paths that cannot take cudaMalloc's device-wide
 * synchronisation.  The per-block partial sums live in a fixed 8 KiB array inside the library
 * (resident on every GPU that loads it until the process exits); the call allocates nothing.
 *   - n <= 2^31 elements, otherwise the call fails with a message (use bqa_grad_sumsq instead).
 *   - The first call on each device looks up the array's device address; make that first call before
 *     starting any stream capture on that device.
 *   - All calls on a device share the array and it keeps whatever the last call wrote; each call
 *     overwrites every entry it later reads, so no clearing is needed, but two calls that can run
 *     concurrently on the GPU (different streams) race on it — serialise them. */
BQA_API int bqa_grad_sumsq_noalloc(const float* g, uint64_t n, double* sumsq, void* stream);
/* chunks: DEVICE array.  host_lr: HOST array of n_groups (<= 8) learning rates.
 * max_grad_norm <= 0 disables clipping.  p_bf16 may be NULL.  zero_grad != 0 clears g. */
BQA_API int bqa_clip_adamw(float* p, float* g, float* m, float* v, void* p_bf16, const bqa_adamw_chunk* chunks,
                           int32_t n_chunks, const double* sumsq, float max_grad_norm, const float* host_lr,
                           int32_t n_groups, float beta1, float beta2, float eps, int32_t zero_grad,
                           void* stream);
/* AdaMod (modules/model/trainer/optim.py:42-100, params.optimizer == 'adamod') in the same fused form:
 * clip, Adam moments with bias correction (step = 1-based step count of this update), decoupled
 * weight decay applied before the update, the moving average `s` (exp_avg_lr, beta3) of the
 * per-element learning rates and the momental bound min(eta, s). */
BQA_API int bqa_clip_adamod(float* p, float* g, float* m, float* v, float* s, void* p_bf16,
                            const bqa_adamw_chunk* chunks, int32_t n_chunks, const double* sumsq,
                            float max_grad_norm, const float* host_lr, int32_t n_groups, float beta1, float beta2,
                            float beta3, float eps, int64_t step, int32_t zero_grad, void* stream);
BQA_API int bqa_f32_to_bf16(const float* src, void* dst, uint64_t n, void* stream);

/* ------------------------------------------------------------------------------------
 * Model inputs from compact token ids, on the device.  Replaces the host loops of collate_fun
 * (modules/model/dataset/split_dataset.py:480-499) and two of the three host-to-device copies of
 * Trainer._to_device (trainer.py:239-249): the host ships int32 ids only.
 *   input_ids = ids, attention_mask = ids > 0 (uint8), token_type_ids = 0 up to and including the
 *   first [SEP] of a row and 1 after it, padding included (all 0 when type_coef == 0: RoBERTa).
 *   *missing_sep (may be NULL) is set to 1 if a row has no [SEP] and type_coef != 0; the reference
 *   raises ValueError there (list.index), and so does the host-side collate of this package.
 * ------------------------------------------------------------------------------------ */
BQA_API int bqa_prepare_inputs(const int32_t* ids32, int32_t B, int32_t S, int32_t sep_token_id, int32_t type_coef,
                               int64_t* input_ids, int64_t* token_type_ids, uint8_t* attention_mask,
                               int32_t* missing_sep, void* stream);

/* ------------------------------------------------------------------------------------
 * Span decoding (modules/model/inference/predictor.py:114-120): per row, first-index arg max
 * and max of the start and end logits, arg max / max probability of softmax(cls), and
 * score = max_start + max_end - (start[0] + end[0]).
 * ------------------------------------------------------------------------------------ */
BQA_API int bqa_decode(const float* start, const float* end, const float* cls, int64_t* start_idx,
                       int64_t* end_idx, int64_t* cls_idx, float* start_max, float* end_max, float* cls_prob,
                       float* score, int32_t B, int32_t S, int32_t C, void* stream);

/* ------------------------------------------------------------------------------------
 * The model step.  Replaces BertForQuestionAnswering.forward (model.py:43-73 and, behind
 * it, transformers BertModel.forward), WeightedLoss.__call__ (loss.py:78-100) and
 * loss.backward() (trainer.py:197-204).
 * ------------------------------------------------------------------------------------ */
typedef struct bqa_config {
  int32_t num_layers, hidden, heads, intermediate;
  int32_t vocab, max_pos, type_vocab, num_labels;
  float ln_eps, hidden_dropout, attn_dropout;
  /* Embedding conventions of the encoder family (ABI version 4).  pad_token_id: row of the word table that is
   * nn.Embedding's padding_idx and receives no gradient (BERT 0, RoBERTa 1).  position_mode 0 (BERT,
   * modeling_bert.py:72-112): position id = index in the sequence.  position_mode 1 (RoBERTa, the reference's
   * second `--model` choice, model.py:9-10 -> RobertaModel; modeling_roberta.py create_position_ids_from_input_ids):
   * position id = pad_token_id + (number of non-padding tokens up to and including this one) for non-padding
   * tokens, pad_token_id for padding tokens; that row of the position table is a padding_idx too. */
  int32_t pad_token_id, position_mode;
} bqa_config;

/* Element offsets of every parameter tensor inside the flat buffers (fp32 master, fp32
 * gradient, bf16 shadow all share it).  Tensors are 128-element aligned; Q/K/V weights
 * (and biases) of a layer are adjacent so that they form one [3H,H] matrix.  Order:
 * embeddings, layer 0 .. L-1, pooler, heads — gradients become final in reverse order. */
typedef struct bqa_layout {
  int64_t word_emb, pos_emb, type_emb, emb_ln_g, emb_ln_b;
  int64_t layer0, layer_stride;
  int64_t qkv_w, qkv_b, ao_w, ao_b, ln1_g, ln1_b, i_w, i_b, o_w, o_b, ln2_g, ln2_b; /* within a layer */
  int64_t pool_w, pool_b, span_w, span_b, cls_w, cls_b, regs_w, regs_b, rege_w, rege_b;
  int64_t total; /* elements, multiple of 128 */
} bqa_layout;

typedef struct bqa_handle bqa_handle;

BQA_API int bqa_param_layout(const bqa_config* cfg, bqa_layout* out);
BQA_API int bqa_workspace_bytes(const bqa_config* cfg, int32_t B, int32_t S, int32_t training, uint64_t* out);
BQA_API int bqa_create(const bqa_config* cfg, bqa_handle** out);
BQA_API int bqa_destroy(bqa_handle* h);
/* params/grads fp32 [total], params_bf16 bf16 [total]; grads may be NULL for inference.
 * workspace: >= bqa_workspace_bytes for the largest (B, S, training) that will be used. */
BQA_API int bqa_bind(bqa_handle* h, float* params, float* grads, void* params_bf16, void* workspace,
                     uint64_t workspace_bytes);

/* input_ids/token_type_ids int64 [B][S] (token_type_ids may be NULL = zeros), attention_mask
 * uint8/bool [B][S] or NULL (= all ones), as produced by collate_fun
 * (modules/model/dataset/split_dataset.py:480-520).  Outputs fp32: start/end [B][S],
 * cls [B][num_labels], reg_start/reg_end [B] (after the sigmoid).  training != 0 applies
 * dropout (streams derived from seed and step) and keeps activations for bqa_backward.
 * training is a bit set: 1 = training forward; | 2 = no dropout inside the encoder; | 4 = no
 * dropout in front of the classifier (the reference's fine-tuning mode, init.py:86-119 +
 * trainer.py:226-233, puts only the selected sub-modules into train mode). */
BQA_API int bqa_forward(bqa_handle* h, const int64_t* input_ids, const int64_t* token_type_ids,
                        const uint8_t* attention_mask, int32_t B, int32_t S, int32_t training, uint64_t seed,
                        uint64_t step, float* start, float* end, float* cls, float* reg_start, float* reg_end,
                        void* stream);

enum bqa_cls_loss { BQA_CLS_SMOOTH = 0, BQA_CLS_CE = 1, BQA_CLS_FOCAL = 2 };
typedef struct bqa_loss_config {
  float w_start, w_end, w_start_reg, w_end_reg, w_cls; /* init.py:19-38                       */
  int32_t cls_loss;                                    /* params.loss: 'smooth' | 'ce' | 'focal' */
  float smooth_alpha;
  const float* cls_weight; /* DEVICE [num_labels] fp32 or NULL (loss='ce' label weights)       */
  float loss_scale;        /* 1 / batch_split (trainer.py:198); scales gradients only          */
  float focal_alpha, focal_gamma; /* FocalLossWithLogits(alpha, gamma) (loss.py:57-71), loss='focal'    */
} bqa_loss_config;

/* Loss of the last bqa_forward outputs against labels (int64 / fp32 [B]).  losses: DEVICE
 * fp32[6] = {start_class, end_class, start_reg, end_reg, cls, total} in the key order of
 * init_losses.  with_grad != 0 also stores d(total * loss_scale)/d(outputs) for bqa_backward. */
BQA_API int bqa_loss(bqa_handle* h, const float* start, const float* end, const float* cls,
                     const float* reg_start, const float* reg_end, const int64_t* y_start, const int64_t* y_end,
                     const int64_t* y_cls, const float* y_reg_start, const float* y_reg_end,
                     const bqa_loss_config* lc, int32_t with_grad, float* losses, void* stream);
/* *host_status = 1 if the last bqa_forward met a token id outside [0, vocab) or a token type outside
 * [0, type_vocab) (torch's nn.Embedding raises IndexError there, modeling_bert.py:72-112; the kernels clamp
 * the id to 0, set this flag and bqa_loss then reports NaN for all six scalars), else 0.  Synchronises
 * `stream`.  host_status is a HOST pointer. */
BQA_API int bqa_input_status(bqa_handle* h, int32_t* host_status, void* stream);
/* Handle-free form of bqa_loss for callers that own every buffer (the autograd-facing
 * WeightedLoss): d_* receive d(total * loss_scale)/d(output) when with_grad != 0 (d_reg_* are
 * w.r.t. the sigmoid outputs); row_scratch: >= 5 * B floats. */
BQA_API int bqa_qa_loss(const float* start, const float* end, const float* cls, const float* reg_start,
                        const float* reg_end, const int64_t* y_start, const int64_t* y_end, const int64_t* y_cls,
                        const float* y_reg_start, const float* y_reg_end, float* d_start, float* d_end,
                        float* d_cls, float* d_reg_start, float* d_reg_end, float* row_scratch, float* losses,
                        int32_t B, int32_t S, int32_t C, const bqa_loss_config* lc, int32_t with_grad,
                        void* stream);
/* Alternative to bqa_loss(with_grad): supply d(loss)/d(outputs) computed elsewhere (autograd). */
BQA_API int bqa_set_output_grads(bqa_handle* h, const float* d_start, const float* d_end, const float* d_cls,
                                 const float* d_reg_start, const float* d_reg_end, void* stream);
/* Backward phases: 0 = heads + pooler, 1..L = encoder layers L-1..0, L+1 = embeddings.
 * Runs phases [first, last]; gradients are ACCUMULATED into the bound gradient buffer.  After
 * phase k every parameter at or beyond the flat offset bqa_phase_grad_begin(k) is final. */
BQA_API int bqa_backward(bqa_handle* h, int32_t first_phase, int32_t last_phase, void* stream);
BQA_API int bqa_num_backward_phases(const bqa_handle* h);
BQA_API int64_t bqa_phase_grad_begin(const bqa_handle* h, int32_t phase);
/* Per-launch device timing of the step's kernels (CUDA events on the launch stream around every
 * GEMM / attention / row-kernel launch between begin and end).  end() synchronises and fills HOST
 * arrays of BQA_PROFILE_KINDS entries (milliseconds, algorithmic work, launches): 0 GEMM forward,
 * 1 dgrad, 2 wgrad, 3 attention forward, 4 attention backward — work = matmul FLOPs; 5 LayerNorm
 * forward, 6 LayerNorm backward, 7 embeddings + LayerNorm forward, 8 embedding backward — work =
 * algorithmic HBM bytes.  Measurement aid for bench.py. */
#define BQA_PROFILE_KINDS 9
BQA_API int bqa_profile_begin(bqa_handle* h);
BQA_API int bqa_profile_end(bqa_handle* h, double* host_ms, double* host_flops, int64_t* host_count);
/* Number of kernels launched by this handle since creation (bench.py's gpu_launches). */
BQA_API int64_t bqa_launch_count(const bqa_handle* h);

#ifdef __cplusplus
}
#endif

#endif /* BQA_H_ */
