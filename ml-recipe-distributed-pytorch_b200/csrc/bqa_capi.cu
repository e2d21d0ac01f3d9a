// bqa_capi.cu — extern "C" surface of libbqa (declared in include/bqa.h).
#include "../../include/bqa.h"

#include <math.h>

#include "bqa_host.cuh"
#include "gemm_sm100.cuh"

using namespace bqa;

extern "C" {

const char* bqa_last_error(void) { return bqa::last_error(); }

int bqa_abi_version(void) { return 4; }

int bqa_gemm_bf16(const bqa_gemm_desc* d, void* stream) {
  if (d == nullptr) {
    set_error("bqa_gemm_bf16: null descriptor");
    return 1;
  }
  int num_sms = 0;
  if (device_sm_count(&num_sms)) return 1;
  GemmProblem p{};
  p.M = d->M;
  p.N = d->N;
  p.K = d->K;
  p.A = d->A;
  p.lda = d->lda;
  p.a_mn_major = d->a_mn_major;
  p.B = d->B;
  p.ldb = d->ldb;
  p.b_mn_major = d->b_mn_major;
  p.C = d->C;
  p.ldc = d->ldc;
  p.C2 = d->C2;
  p.epilogue = d->epilogue;
  p.bias = d->bias;
  p.residual = d->residual;
  p.ld_res = d->ld_res;
  p.aux = d->aux;
  p.ld_aux = d->ld_aux;
  p.k_splits = d->k_splits;
  p.drop = make_drop_key(d->drop_seed, d->drop_site, d->dropout_p);
  p.colsum_out = d->colsum_out;
  p.dot_with = d->dot_with;
  p.ld_dot = d->ld_dot;
  p.dot_out = d->dot_out;
  p.dot_rows = d->dot_rows;
  GemmPlan plan;
  if (gemm_plan(&plan, p, num_sms)) return 1;
  return gemm_launch(plan, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// standalone operator entry points (unit-testable pieces of the step)
// ---------------------------------------------------------------------------------------
#include "attention_sm100.cuh"
#include "kernels.cuh"

extern "C" {

int bqa_attention_keep_words(int32_t S) { return attn_keep_words(S); }

int bqa_attention_forward(const void* qkv, void* ctx, float* lse2, uint32_t* keep_bits, const uint8_t* key_mask,
                          int32_t B, int32_t S, int32_t A, float dropout_p, uint64_t drop_seed, uint32_t drop_site,
                          void* stream) {
  AttnPlan plan;
  AttnShape s{B, S, A, A * 64};
  if (attn_plan(&plan, s, qkv, ctx, lse2, nullptr, nullptr, nullptr, keep_bits)) return 1;
  return attn_forward(plan, key_mask, make_drop_key(drop_seed, drop_site, dropout_p),
                      reinterpret_cast<cudaStream_t>(stream));
}

int bqa_attention_backward(const void* qkv, const void* ctx, const float* lse2, const uint32_t* keep_bits,
                           const void* dctx, void* dqkv, float* dsum, const uint8_t* key_mask, int32_t B, int32_t S,
                           int32_t A, float dropout_p, uint64_t drop_seed, uint32_t drop_site, float* dqkv_bias_grad,
                           void* stream) {
  AttnPlan plan;
  AttnShape s{B, S, A, A * 64};
  if (attn_plan(&plan, s, qkv, const_cast<void*>(ctx), const_cast<float*>(lse2), dctx, dsum, dqkv,
                const_cast<uint32_t*>(keep_bits)))
    return 1;
  return attn_backward(plan, key_mask, make_drop_key(drop_seed, drop_site, dropout_p),
                       reinterpret_cast<cudaStream_t>(stream), /*dsum_ready=*/false, dqkv_bias_grad);
}

int bqa_layernorm_forward(const void* z, const float* gamma, const float* beta, void* y, int32_t T, int32_t H,
                          float eps, void* stream) {
  return launch_ln_fwd(reinterpret_cast<const bf16*>(z), gamma, beta, reinterpret_cast<bf16*>(y), T, H, eps,
                       reinterpret_cast<cudaStream_t>(stream));
}

int bqa_layernorm_backward(const void* dy, const void* z, const float* gamma, void* dz, void* dz_masked,
                           float* dgamma, float* dbeta, float* dbias, int32_t T, int32_t H, float eps,
                           float drop_in_p, uint32_t drop_in_site, float drop_out_p, uint32_t drop_out_site,
                           uint64_t drop_seed, void* stream) {
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  return launch_ln_bwd(reinterpret_cast<const bf16*>(dy), reinterpret_cast<const bf16*>(z), gamma,
                       reinterpret_cast<bf16*>(dz), reinterpret_cast<bf16*>(dz_masked), dgamma, dbeta, dbias, T, H,
                       eps, make_drop_key(drop_seed, drop_in_site, drop_in_p),
                       make_drop_key(drop_seed, drop_out_site, drop_out_p), sms,
                       reinterpret_cast<cudaStream_t>(stream));
}

int bqa_colsum_bf16(const void* X, int32_t ld, float* out, int32_t T, int32_t N, void* stream) {
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  return launch_colsum(reinterpret_cast<const bf16*>(X), ld, out, T, N, sms, reinterpret_cast<cudaStream_t>(stream));
}

int bqa_sumsq_scratch_floats(void) { return sumsq_scratch_floats(256); }

int bqa_grad_sumsq(const float* g, uint64_t n, double* sumsq, float* scratch, void* stream) {
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  if (sms > 256) sms = 256;
  return launch_grad_sumsq(g, n, sumsq, scratch, sms, reinterpret_cast<cudaStream_t>(stream));
}

int bqa_clip_adamw(float* p, float* g, float* m, float* v, void* p_bf16, const bqa_adamw_chunk* chunks,
                   int32_t n_chunks, const double* sumsq, float max_grad_norm, const float* host_lr,
                   int32_t n_groups, float beta1, float beta2, float eps, int32_t zero_grad, void* stream) {
  static_assert(sizeof(bqa_adamw_chunk) == sizeof(AdamWChunk), "chunk layouts must match");
  if (n_groups < 1 || n_groups > 8 || host_lr == nullptr) {
    set_error("bqa_clip_adamw: need 1..8 parameter groups");
    return 1;
  }
  if (max_grad_norm > 0.0f && sumsq == nullptr) {
    set_error("bqa_clip_adamw: clipping needs the sum of squares");
    return 1;
  }
  LrTable t{};
  for (int i = 0; i < n_groups; ++i) t.lr[i] = host_lr[i];
  return launch_clip_adamw(p, g, m, v, reinterpret_cast<bf16*>(p_bf16), reinterpret_cast<const AdamWChunk*>(chunks),
                           n_chunks, sumsq, max_grad_norm, t, beta1, beta2, eps, zero_grad,
                           reinterpret_cast<cudaStream_t>(stream));
}

int bqa_clip_adamod(float* p, float* g, float* m, float* v, float* s, void* p_bf16, const bqa_adamw_chunk* chunks,
                    int32_t n_chunks, const double* sumsq, float max_grad_norm, const float* host_lr,
                    int32_t n_groups, float beta1, float beta2, float beta3, float eps, int64_t step,
                    int32_t zero_grad, void* stream) {
  if (n_groups < 1 || n_groups > 8 || host_lr == nullptr) {
    set_error("bqa_clip_adamod: need 1..8 parameter groups");
    return 1;
  }
  if (max_grad_norm > 0.0f && sumsq == nullptr) {
    set_error("bqa_clip_adamod: clipping needs the sum of squares");
    return 1;
  }
  if (step < 1 || s == nullptr) {
    set_error("bqa_clip_adamod: step must be >= 1 and the exp_avg_lr buffer non-null");
    return 1;
  }
  LrTable t{};
  for (int i = 0; i < n_groups; ++i) t.lr[i] = host_lr[i];
  // step_size = lr * sqrt(1 - beta2^t) / (1 - beta1^t)   (optim.py:83-85), in double on the host
  const double bc1 = 1.0 - pow(static_cast<double>(beta1), static_cast<double>(step));
  const double bc2 = 1.0 - pow(static_cast<double>(beta2), static_cast<double>(step));
  const float bias_mul = static_cast<float>(sqrt(bc2) / bc1);
  return launch_clip_adamod(p, g, m, v, s, reinterpret_cast<bf16*>(p_bf16),
                            reinterpret_cast<const AdamWChunk*>(chunks), n_chunks, sumsq, max_grad_norm, t, beta1,
                            beta2, beta3, eps, bias_mul, zero_grad, reinterpret_cast<cudaStream_t>(stream));
}

int bqa_f32_to_bf16(const float* src, void* dst, uint64_t n, void* stream) {
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  return launch_f32_to_bf16(src, reinterpret_cast<bf16*>(dst), n, sms, reinterpret_cast<cudaStream_t>(stream));
}

int bqa_decode(const float* start, const float* end, const float* cls, int64_t* start_idx, int64_t* end_idx,
               int64_t* cls_idx, float* start_max, float* end_max, float* cls_prob, float* score, int32_t B,
               int32_t S, int32_t C, void* stream) {
  if (B < 1 || S < 1 || C < 1) {
    set_error("bqa_decode: empty input (B=%d S=%d C=%d)", B, S, C);
    return 1;
  }
  return launch_decode(start, end, cls, start_idx, end_idx, cls_idx, start_max, end_max, cls_prob, score, B, S, C,
                       reinterpret_cast<cudaStream_t>(stream));
}

int bqa_prepare_inputs(const int32_t* ids32, int32_t B, int32_t S, int32_t sep_token_id, int32_t type_coef,
                       int64_t* input_ids, int64_t* token_type_ids, uint8_t* attention_mask, int32_t* missing_sep,
                       void* stream) {
  if (B < 1 || S < 1) {
    set_error("bqa_prepare_inputs: empty input (B=%d S=%d)", B, S);
    return 1;
  }
  return launch_prepare_inputs(ids32, input_ids, token_type_ids, attention_mask, B, S, sep_token_id, type_coef,
                               missing_sep, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
