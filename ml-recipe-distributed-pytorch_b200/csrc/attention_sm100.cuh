// attention_sm100.cuh — host interface of the fused self-attention kernels (attention_sm100.cu).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "bqa_common.cuh"

namespace bqa {

struct AttnShape {
  int B, S, A, H;  // batch, sequence length (<= 512), heads, hidden (= A * 64)
};

struct AttnPlan {
  CUtensorMap tm_qkv;   // [B][S][3H] bf16, box {64,128,1}
  CUtensorMap tm_kv96;  // qkv with box {64,96,1}: K / V chunks of the forward kernel
  CUtensorMap tm_dctx;  // [B][S][H]  bf16, box {64,128,1}
  CUtensorMap tm_q64;   // qkv with box {64,64,1}: 64-row Q chunks of the dK/dV kernel
  CUtensorMap tm_do64;  // dctx with box {64,64,1}
  CUtensorMap tm_dqkv32;  // dqkv with box {64,32,1}: dK/dV stores, one per TMEM lane quarter
  CUtensorMap tm_ctx;   // [B][S][H]  bf16 (forward output, TMA store)
  CUtensorMap tm_dqkv;  // [B][S][3H] bf16 (backward output, TMA store)
  AttnShape shape;
  int nchunk;     // ceil(S / 128)
  int tmem_cols;  // forward allocation (power of two holding the scores and, if it fits, the output)
  const bf16* qkv;
  bf16* ctx;
  float* lse2;   // [B][A][S] log2-domain log-sum-exp of the scaled, masked scores
  const bf16* dctx;
  float* dsum;   // [B][A][S] rowsum(dO * O)
  bf16* dqkv;    // [B*S][3H]
  uint32_t* keep;  // [B][A][S][attn_keep_words(S)] dropout keep bits (forward writes, backward reads) or null
};

// qkv/ctx/dctx/dqkv may be null when only the forward (or only a plan for it) is needed.
int attn_plan(AttnPlan* plan, const AttnShape& s, const void* qkv, void* ctx, float* lse2, const void* dctx,
              float* dsum, void* dqkv, uint32_t* keep);
// 32-bit words of keep bits per (batch, head, position) row for sequence length S: the decisions of
// query r over all keys, then those of key r over all queries (attention_sm100.cu: attn_keep_bits_kernel)
inline int attn_keep_words(int S) { return 8 * ((S + 127) / 128); }
// Writes the dropout keep bits of one attention call (plan.keep).  The decisions depend only on the
// key, so this can run ahead of / concurrently with whatever produces qkv.
int attn_make_keep_bits(const AttnPlan& plan, DropKey drop, cudaStream_t st);
// key_mask: [B][S] uint8 (1 = attend) or null.  drop.thresh == 0 disables dropout; otherwise plan.keep
// must already hold this call's keep bits (bits_ready) or they are generated first on `st`.
int attn_forward(const AttnPlan& plan, const uint8_t* key_mask, DropKey drop, cudaStream_t st,
                 bool bits_ready = false);
// dsum_ready: plan.dsum already holds D = rowsum(dO * O) (written by the epilogue of the GEMM that produced
// dctx); otherwise a small pre-pass computes it.
// dqkv_bias_grad: fp32 [3H] or null; the column sums of dqkv (gradient of the fused QKV bias) are accumulated
// into it by the kernels' drains, so that no separate pass over dqkv is needed.
int attn_backward(const AttnPlan& plan, const uint8_t* key_mask, DropKey drop, cudaStream_t st,
                  bool dsum_ready = false, float* dqkv_bias_grad = nullptr);

}  // namespace bqa
