// engine.cu — the BERT-QA step as a sequence of launches on one stream: parameter layout,
// activation workspace, forward, loss, phased backward.  Implements the handle part of
// include/bqa.h.
//
// Reference data flow being replaced (per micro-batch):
//   Trainer._train                        modules/model/trainer/trainer.py:275-290
//   -> BertForQuestionAnswering.forward   modules/model/model/model.py:43-73
//      -> BertModel.forward               site-packages/transformers/models/bert/modeling_bert.py:628
//   -> WeightedLoss.__call__              modules/model/model/loss.py:78-100
//   -> loss.backward()                    trainer.py:197-204
//
// HBM layout.  Parameters: one flat fp32 master buffer, one fp32 gradient buffer and one bf16
// shadow (GEMM operands) sharing the offsets of bqa_layout.  Activations: [T = B*S][width]
// bf16 row-major slabs carved from the caller's workspace; per layer the backward needs
//   qkv [T,3H] | ctx [T,H] | z1 [T,H] | a [T,H] | u [T,I] | g [T,I] | z2 [T,H] | x_{l+1} [T,H]
// plus lse2 [B,A,S] fp32; in inference two layer slots are ping-ponged instead of L.
#include "../../include/bqa.h"

#include <new>
#include <vector>

#include "attention_sm100.cuh"
#include "bqa_host.cuh"
#include "gemm_sm100.cuh"
#include "kernels.cuh"

using namespace bqa;

namespace {

constexpr int64_t ALIGN_ELEMS = 128;
inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

int compute_layout(const bqa_config& c, bqa_layout* L) {
  if (c.num_layers < 1 || c.hidden < 64 || c.hidden % 64 != 0 || c.heads * 64 != c.hidden || c.hidden > 1024 ||
      c.intermediate % 64 != 0 || c.intermediate < 64 || c.vocab < 1 || c.max_pos < 1 || c.type_vocab < 1 ||
      c.type_vocab > 2 || c.num_labels < 1 || c.num_labels > 32 || c.pad_token_id < 0 || c.pad_token_id >= c.vocab ||
      (c.position_mode != 0 && c.position_mode != 1)) {
    set_error("config: unsupported shape (L=%d H=%d A=%d I=%d V=%d P=%d TV=%d C=%d): need H = 64*A <= 1024, "
              "H and I multiples of 64, type_vocab <= 2",
              c.num_layers, c.hidden, c.heads, c.intermediate, c.vocab, c.max_pos, c.type_vocab, c.num_labels);
    return 1;
  }
  const int64_t H = c.hidden, I = c.intermediate;
  int64_t o = 0;
  auto take = [&](int64_t n) {
    const int64_t at = o;
    o = align_up(o + n, ALIGN_ELEMS);
    return at;
  };
  L->word_emb = take(static_cast<int64_t>(c.vocab) * H);
  L->pos_emb = take(static_cast<int64_t>(c.max_pos) * H);
  L->type_emb = take(static_cast<int64_t>(c.type_vocab) * H);
  L->emb_ln_g = take(H);
  L->emb_ln_b = take(H);
  L->layer0 = o;
  const int64_t base = o;
  L->qkv_w = take(3 * H * H) - base;
  L->qkv_b = take(3 * H) - base;
  L->ao_w = take(H * H) - base;
  L->ao_b = take(H) - base;
  L->ln1_g = take(H) - base;
  L->ln1_b = take(H) - base;
  L->i_w = take(I * H) - base;
  L->i_b = take(I) - base;
  L->o_w = take(H * I) - base;
  L->o_b = take(H) - base;
  L->ln2_g = take(H) - base;
  L->ln2_b = take(H) - base;
  L->layer_stride = o - base;
  o = base + L->layer_stride * c.num_layers;
  L->pool_w = take(H * H);
  L->pool_b = take(H);
  L->span_w = take(2 * H);
  L->span_b = take(2);
  L->cls_w = take(static_cast<int64_t>(c.num_labels) * H);
  L->cls_b = take(c.num_labels);
  L->regs_w = take(H);
  L->regs_b = take(1);
  L->rege_w = take(H);
  L->rege_b = take(1);
  L->total = o;
  return 0;
}

// dropout call sites of one step (1 + 3L + 1)
inline uint32_t site_embed() { return 0; }
inline uint32_t site_attn(int l) { return 1 + 3 * l; }
inline uint32_t site_ao(int l) { return 2 + 3 * l; }
inline uint32_t site_ffn(int l) { return 3 + 3 * l; }
inline uint32_t site_cls(int L) { return 1 + 3 * L; }

struct Bump {
  uint8_t* base;
  uint64_t off = 0;
  template <typename T>
  T* take(uint64_t count) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off = (off + count * sizeof(T) + 255) & ~static_cast<uint64_t>(255);
    return p;
  }
};

struct LayerActs {
  bf16 *qkv, *ctx, *z1, *a, *u, *g, *z2, *x_out;  // u holds gelu'(pre-activation), g the activation
  float* lse2;
  uint32_t* keep;  // attention-dropout keep bits
};

struct Workspace {
  bf16 *z0, *x0;
  std::vector<LayerActs> layer;  // L slots in training, 2 in inference
  float *pooled, *reg_s, *reg_e;  // pooler output and the two sigmoid outputs, kept for the backward
  float *d_start, *d_end, *d_cls, *d_reg_s, *d_reg_e;
  float *loss_rows, *dpre, *dx0;
  uint8_t* mask;
  int64_t *ids, *tt;
  int32_t* pos_ids;  // position ids of the forward (RoBERTa counts non-padding tokens)
  // backward temporaries
  bf16 *dy0, *dy1, *dres, *dgm, *du, *dctx, *dqkv;
  float* dsum;
  int* err_flag;
  uint64_t bytes;
};

void carve(Workspace* w, uint8_t* base, const bqa_config& c, int B, int S, bool training) {
  Bump b{base};
  const uint64_t T = static_cast<uint64_t>(B) * S, H = c.hidden, I = c.intermediate;
  w->z0 = b.take<bf16>(T * H);
  w->x0 = b.take<bf16>(T * H);
  const int slots = training ? c.num_layers : (c.num_layers < 2 ? c.num_layers : 2);
  w->layer.resize(slots);
  for (int i = 0; i < slots; ++i) {
    LayerActs& a = w->layer[i];
    a.qkv = b.take<bf16>(T * 3 * H);
    a.ctx = b.take<bf16>(T * H);
    a.z1 = b.take<bf16>(T * H);
    a.a = b.take<bf16>(T * H);
    a.u = b.take<bf16>(T * I);
    a.g = b.take<bf16>(T * I);
    a.z2 = b.take<bf16>(T * H);
    a.x_out = b.take<bf16>(T * H);
    a.lse2 = b.take<float>(static_cast<uint64_t>(B) * c.heads * S);
    a.keep = (training && c.attn_dropout > 0.0f)
                 ? b.take<uint32_t>(static_cast<uint64_t>(B) * c.heads * S * attn_keep_words(S))
                 : nullptr;
  }
  w->pooled = b.take<float>(static_cast<uint64_t>(B) * H);
  w->reg_s = b.take<float>(B);
  w->reg_e = b.take<float>(B);
  w->mask = b.take<uint8_t>(T);
  w->ids = b.take<int64_t>(T);
  w->tt = b.take<int64_t>(T);
  w->pos_ids = b.take<int32_t>(T);
  w->err_flag = b.take<int>(1);
  w->loss_rows = b.take<float>(static_cast<uint64_t>(B) * 5);
  if (training) {
    w->d_start = b.take<float>(T);
    w->d_end = b.take<float>(T);
    w->d_cls = b.take<float>(static_cast<uint64_t>(B) * c.num_labels);
    w->d_reg_s = b.take<float>(B);
    w->d_reg_e = b.take<float>(B);
    w->dpre = b.take<float>(static_cast<uint64_t>(B) * H);
    w->dx0 = b.take<float>(static_cast<uint64_t>(B) * H);
    w->dy0 = b.take<bf16>(T * H);
    w->dy1 = b.take<bf16>(T * H);
    w->dres = b.take<bf16>(T * H);
    w->dgm = b.take<bf16>(T * H);
    w->du = b.take<bf16>(T * I);
    w->dctx = b.take<bf16>(T * H);
    w->dqkv = b.take<bf16>(T * 3 * H);
    w->dsum = b.take<float>(static_cast<uint64_t>(B) * c.heads * S);
  } else {
    w->d_start = w->d_end = w->d_cls = w->d_reg_s = w->d_reg_e = w->dpre = w->dx0 = w->dsum = nullptr;
    w->dy0 = w->dy1 = w->dres = w->dgm = w->du = w->dctx = w->dqkv = nullptr;
  }
  w->bytes = b.off;
}

struct LayerPlans {
  GemmPlan qkv, ao, ffn1, ffn2;                   // forward
  GemmPlan d_ffn2, d_ffn1, d_ao, d_qkv;           // dgrad
  GemmPlan w_ffn2, w_ffn1, w_ao, w_qkv;           // wgrad
  AttnPlan attn;
};

struct Plan {
  int B = 0, S = 0;
  bool training = false;
  Workspace ws;
  std::vector<LayerPlans> layer;
};

}  // namespace

struct bqa_handle {
  bqa_config cfg;
  bqa_layout lay;
  int num_sms = 0;
  float* params = nullptr;
  float* grads = nullptr;
  bf16* params_bf16 = nullptr;
  uint8_t* workspace = nullptr;
  uint64_t workspace_bytes = 0;
  Plan plan;
  bool plan_valid = false;
  // state of the last forward
  uint64_t seed = 0, step = 0;
  bool last_training = false;
  bool enc_dropout_off = false, cls_dropout_off = false;  // bqa_forward training bits 2 / 4
  // attention keep bits generated ahead of time (at the end of a backward, for the next step): valid for
  // this (seed, step, B, S) if bits_ahead
  bool bits_ahead = false;
  uint64_t bits_seed = 0, bits_step = 0;
  int bits_B = 0, bits_S = 0;
  uint32_t cls_site = 0;
  bool has_mask = false;
  int64_t launches = 0;
  // optional per-launch timing of the tensor-core kernels (bench.py's roofline leg)
  struct Timed {
    cudaEvent_t a, b;
    double flops;
    // 0 = GEMM forward, 1 = dgrad, 2 = wgrad, 3 = attention forward, 4 = attention backward (flops = matmul FLOPs);
    // 5 = LayerNorm forward, 6 = LayerNorm backward, 7 = embeddings + LayerNorm forward, 8 = embedding backward
    // (flops = algorithmic HBM bytes)
    int kind;
  };
  // side stream on which the attention-dropout keep bits of all layers are generated while the main
  // stream runs the embeddings and the first GEMMs (the decisions depend only on the step's keys)
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr;
  cudaEvent_t ev_tail = nullptr;  // end of the ahead-of-time keep-bit generation on the side stream
  bool side_pending = false;      // that work has not been ordered before the main stream yet
  std::vector<cudaEvent_t> ev_bits;
  bool profiling = false;
  std::vector<Timed> timed;
  std::vector<cudaEvent_t> event_pool;
  cudaEvent_t get_event() {
    if (!event_pool.empty()) {
      cudaEvent_t e = event_pool.back();
      event_pool.pop_back();
      return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
  }
  void time_begin(int kind, double flops, cudaStream_t st) {
    if (!profiling) return;
    Timed t{get_event(), get_event(), flops, kind};
    cudaEventRecord(t.a, st);
    timed.push_back(t);
  }
  void time_end(cudaStream_t st) {
    if (!profiling) return;
    cudaEventRecord(timed.back().b, st);
  }

  const float* P(int64_t off) const { return params + off; }
  float* G(int64_t off) const { return grads + off; }
  const bf16* W(int64_t off) const { return params_bf16 + off; }
  int64_t layer_off(int l) const { return lay.layer0 + lay.layer_stride * l; }
  DropKey key_at(uint64_t seed_, uint64_t step_, uint32_t site, float p) const {
    if (p <= 0.0f) return make_drop_key(0, 0, 0.0f);
    return make_drop_key(seed_ ^ splitmix64(step_ + 0x51ed270b9f3a6c1dull), site, p);
  }
  DropKey key(uint32_t site, float p) const {
    if (!last_training || p <= 0.0f) return make_drop_key(0, 0, 0.0f);
    if (site == cls_site ? cls_dropout_off : enc_dropout_off) return make_drop_key(0, 0, 0.0f);
    return make_drop_key(seed ^ splitmix64(step + 0x51ed270b9f3a6c1dull), site, p);
  }
};

namespace {

int build_plan(bqa_handle* h, int B, int S, bool training) {
  h->bits_ahead = false;  // the keep-bit buffers move with the plan
  const bqa_config& c = h->cfg;
  Plan& p = h->plan;
  h->plan_valid = false;
  p.B = B;
  p.S = S;
  p.training = training;
  carve(&p.ws, h->workspace, c, B, S, training);
  if (p.ws.bytes > h->workspace_bytes) {
    set_error("workspace too small: need %llu bytes for B=%d S=%d training=%d, bound %llu",
              (unsigned long long)p.ws.bytes, B, S, (int)training, (unsigned long long)h->workspace_bytes);
    return 1;
  }
  const int T = B * S, H = c.hidden, I = c.intermediate;
  p.layer.assign(c.num_layers, LayerPlans{});
  const int slots = static_cast<int>(p.ws.layer.size());
  for (int l = 0; l < c.num_layers; ++l) {
    LayerPlans& lp = p.layer[l];
    const LayerActs& act = p.ws.layer[l % slots];
    const bf16* x_in = l == 0 ? p.ws.x0 : p.ws.layer[(l - 1) % slots].x_out;
    const int64_t lo = h->layer_off(l);
    const bqa_layout& Y = h->lay;
    auto fwd = [&](GemmPlan* out, const bf16* A, int K, const bf16* Wt, int N, void* C, void* C2, int epi,
                   const float* bias, const bf16* residual) {
      GemmProblem g{};
      g.M = T; g.N = N; g.K = K;
      g.A = A; g.lda = K; g.a_mn_major = 0;
      g.B = Wt; g.ldb = K; g.b_mn_major = 0;
      g.C = C; g.ldc = N; g.C2 = C2;
      g.epilogue = epi; g.bias = bias;
      g.residual = residual; g.ld_res = N;
      g.k_splits = 1;
      g.drop = make_drop_key(0, 0, 0.0f);
      return gemm_plan(out, g, h->num_sms);
    };
    if (fwd(&lp.qkv, x_in, H, h->W(lo + Y.qkv_w), 3 * H, act.qkv, nullptr, EPI_LINEAR, h->P(lo + Y.qkv_b), nullptr))
      return 1;
    if (fwd(&lp.ao, act.ctx, H, h->W(lo + Y.ao_w), H, act.z1, nullptr, EPI_LINEAR, h->P(lo + Y.ao_b), x_in)) return 1;
    if (fwd(&lp.ffn1, act.a, H, h->W(lo + Y.i_w), I, act.u, act.g, EPI_GELU_FWD, h->P(lo + Y.i_b), nullptr)) return 1;
    if (fwd(&lp.ffn2, act.g, I, h->W(lo + Y.o_w), H, act.z2, nullptr, EPI_LINEAR, h->P(lo + Y.o_b), act.a)) return 1;
    AttnShape as{B, S, c.heads, H};
    if (attn_plan(&lp.attn, as, act.qkv, act.ctx, act.lse2, training ? p.ws.dctx : nullptr, p.ws.dsum,
                  p.ws.dqkv, act.keep))
      return 1;
    if (!training) continue;
    // dgrad: dX[T,Kin] = dY[T,Nout] . W[Nout,Kin]   (A K-major, W is the MN-major operand)
    auto dgrad = [&](GemmPlan* out, const bf16* dY, int Nout, const bf16* Wt, int Kin, void* C, int epi,
                     const bf16* residual, const bf16* aux) {
      GemmProblem g{};
      g.M = T; g.N = Kin; g.K = Nout;
      g.A = dY; g.lda = Nout; g.a_mn_major = 0;
      g.B = Wt; g.ldb = Kin; g.b_mn_major = 1;
      g.C = C; g.ldc = Kin;
      g.epilogue = epi;
      g.residual = residual; g.ld_res = Kin;
      g.aux = aux; g.ld_aux = Kin;
      g.k_splits = 1;
      g.drop = make_drop_key(0, 0, 0.0f);
      return gemm_plan(out, g, h->num_sms);
    };
    // wgrad: dW[Nout,Kin] += dY[T,Nout]^T . X[T,Kin]   (both MN-major, fp32 reduce-add, split-K)
    auto wgrad = [&](GemmPlan* out, const bf16* dY, int Nout, const bf16* X, int Kin, float* dW) {
      GemmProblem g{};
      g.M = Nout; g.N = Kin; g.K = T;
      g.A = dY; g.lda = Nout; g.a_mn_major = 1;
      g.B = X; g.ldb = Kin; g.b_mn_major = 1;
      g.C = dW; g.ldc = Kin;
      g.epilogue = EPI_F32_ADD;
      g.k_splits = 0;  // automatic: whole waves of tiles, see gemm_plan
      g.drop = make_drop_key(0, 0, 0.0f);
      return gemm_plan(out, g, h->num_sms);
    };
    Workspace& w = p.ws;
    bf16* dy_out = (l & 1) ? w.dy1 : w.dy0;  // gradient w.r.t. this layer's input
    // without hidden dropout the masked gradient IS the residual gradient: read it in place
    const bf16* dgm = c.hidden_dropout > 0.0f ? w.dgm : w.dres;
    {
      // du = (dgm . W2) * gelu'(u); the epilogue also accumulates the bias gradient of the intermediate
      // projection (column sums of du) instead of a separate pass over du
      GemmProblem g{};
      g.M = T; g.N = I; g.K = H;
      g.A = dgm; g.lda = H; g.a_mn_major = 0;
      g.B = h->W(lo + Y.o_w); g.ldb = I; g.b_mn_major = 1;
      g.C = w.du; g.ldc = I;
      g.epilogue = EPI_GELU_BWD;
      g.aux = act.u; g.ld_aux = I;
      g.k_splits = 1;
      g.drop = make_drop_key(0, 0, 0.0f);
      g.colsum_out = h->G(lo + Y.i_b);
      if (gemm_plan(&lp.d_ffn2, g, h->num_sms)) return 1;
    }
    if (wgrad(&lp.w_ffn2, dgm, H, act.g, I, h->G(lo + Y.o_w))) return 1;
    if (dgrad(&lp.d_ffn1, w.du, I, h->W(lo + Y.i_w), H, w.dctx /* da, reuses dctx slab */, EPI_LINEAR, w.dres,
              nullptr))
      return 1;
    if (wgrad(&lp.w_ffn1, w.du, I, act.a, H, h->G(lo + Y.i_w))) return 1;
    {
      // dctx = dgm . Wo; its epilogue also emits D = rowsum(dctx * ctx) per (batch, head, query), which the
      // attention backward needs (no separate pass over dctx and ctx)
      GemmProblem g{};
      g.M = T; g.N = H; g.K = H;
      g.A = dgm; g.lda = H; g.a_mn_major = 0;
      g.B = h->W(lo + Y.ao_w); g.ldb = H; g.b_mn_major = 1;
      g.C = w.dctx; g.ldc = H;
      g.epilogue = EPI_LINEAR;
      g.k_splits = 1;
      g.drop = make_drop_key(0, 0, 0.0f);
      g.dot_with = act.ctx; g.ld_dot = H; g.dot_out = w.dsum; g.dot_rows = S;
      if (gemm_plan(&lp.d_ao, g, h->num_sms)) return 1;
    }
    if (wgrad(&lp.w_ao, dgm, H, act.ctx, H, h->G(lo + Y.ao_w))) return 1;
    if (dgrad(&lp.d_qkv, w.dqkv, 3 * H, h->W(lo + Y.qkv_w), H, dy_out, EPI_LINEAR, w.dres, nullptr)) return 1;
    if (wgrad(&lp.w_qkv, w.dqkv, 3 * H, x_in, H, h->G(lo + Y.qkv_w))) return 1;
  }
  h->plan_valid = true;
  return 0;
}

int launch_gemm(bqa_handle* h, GemmPlan& plan, DropKey drop, cudaStream_t st) {
  plan.args.drop = drop;
  h->launches += 1;
  const int kind = (plan.variant >> 2) == EPI_F32_ADD ? 2 : ((plan.variant & 2) ? 1 : 0);
  h->time_begin(kind, 2.0 * plan.args.M * plan.args.N * plan.args.K, st);
  const int rc = gemm_launch(plan, st);
  h->time_end(st);
  return rc;
}

// matmul FLOPs of one attention call: QK^T and PV forward (4 S^2 d per head), 7 products backward
double attn_flops(const bqa_handle* h, int B, int S, bool backward) {
  const double per = 2.0 * B * h->cfg.heads * static_cast<double>(S) * S * 64.0;
  return backward ? 7.0 * per : 2.0 * per;
}

HeadParams head_params(const bqa_handle* h) {
  const bqa_layout& Y = h->lay;
  HeadParams hp;
  hp.pool_w = h->P(Y.pool_w); hp.pool_b = h->P(Y.pool_b);
  hp.span_w = h->P(Y.span_w); hp.span_b = h->P(Y.span_b);
  hp.cls_w = h->P(Y.cls_w); hp.cls_b = h->P(Y.cls_b);
  hp.regs_w = h->P(Y.regs_w); hp.regs_b = h->P(Y.regs_b);
  hp.rege_w = h->P(Y.rege_w); hp.rege_b = h->P(Y.rege_b);
  return hp;
}

HeadGrads head_grads(const bqa_handle* h) {
  const bqa_layout& Y = h->lay;
  HeadGrads hg;
  hg.pool_w = h->G(Y.pool_w); hg.pool_b = h->G(Y.pool_b);
  hg.span_w = h->G(Y.span_w); hg.span_b = h->G(Y.span_b);
  hg.cls_w = h->G(Y.cls_w); hg.cls_b = h->G(Y.cls_b);
  hg.regs_w = h->G(Y.regs_w); hg.regs_b = h->G(Y.regs_b);
  hg.rege_w = h->G(Y.rege_w); hg.rege_b = h->G(Y.rege_b);
  return hg;
}

}  // namespace

extern "C" {

int bqa_param_layout(const bqa_config* cfg, bqa_layout* out) {
  if (cfg == nullptr || out == nullptr) {
    set_error("bqa_param_layout: null argument");
    return 1;
  }
  return compute_layout(*cfg, out);
}

int bqa_workspace_bytes(const bqa_config* cfg, int32_t B, int32_t S, int32_t training, uint64_t* out) {
  bqa_layout L;
  if (cfg == nullptr || out == nullptr || compute_layout(*cfg, &L)) return 1;
  if (B < 1 || S < 1 || S > 512 || S > cfg->max_pos) {
    set_error("bqa_workspace_bytes: B=%d S=%d out of range (S <= min(512, max_pos=%d))", B, S, cfg->max_pos);
    return 1;
  }
  Workspace w;
  carve(&w, nullptr, *cfg, B, S, training != 0);
  *out = w.bytes;
  return 0;
}

int bqa_create(const bqa_config* cfg, bqa_handle** out) {
  if (cfg == nullptr || out == nullptr) {
    set_error("bqa_create: null argument");
    return 1;
  }
  bqa_layout L;
  if (compute_layout(*cfg, &L)) return 1;
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  int dev = 0, major = 0;
  BQA_CUDA_CHECK(cudaGetDevice(&dev));
  BQA_CUDA_CHECK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) {
    set_error("bqa_create: device compute capability %d.x is not sm_100 (this library has no other code path)", major);
    return 1;
  }
  bqa_handle* h = new (std::nothrow) bqa_handle();
  if (h == nullptr) {
    set_error("bqa_create: out of host memory");
    return 1;
  }
  h->cfg = *cfg;
  h->lay = L;
  h->num_sms = sms;
  *out = h;
  return 0;
}

int bqa_destroy(bqa_handle* h) {
  if (h != nullptr) {
    for (auto& t : h->timed) {
      cudaEventDestroy(t.a);
      cudaEventDestroy(t.b);
    }
    for (auto e : h->event_pool) cudaEventDestroy(e);
    for (auto e : h->ev_bits) cudaEventDestroy(e);
    if (h->ev_fork != nullptr) cudaEventDestroy(h->ev_fork);
    if (h->ev_tail != nullptr) cudaEventDestroy(h->ev_tail);
    if (h->side != nullptr) cudaStreamDestroy(h->side);
  }
  delete h;
  return 0;
}

int bqa_profile_begin(bqa_handle* h) {
  if (h == nullptr) {
    set_error("bqa_profile_begin: null handle");
    return 1;
  }
  for (auto& t : h->timed) {
    h->event_pool.push_back(t.a);
    h->event_pool.push_back(t.b);
  }
  h->timed.clear();
  h->profiling = true;
  return 0;
}

int bqa_profile_end(bqa_handle* h, double* host_ms, double* host_flops, int64_t* host_count) {
  if (h == nullptr || host_ms == nullptr || host_flops == nullptr || host_count == nullptr) {
    set_error("bqa_profile_end: null argument");
    return 1;
  }
  h->profiling = false;
  for (int k = 0; k < BQA_PROFILE_KINDS; ++k) {
    host_ms[k] = 0.0;
    host_flops[k] = 0.0;
    host_count[k] = 0;
  }
  for (auto& t : h->timed) {
    BQA_CUDA_CHECK(cudaEventSynchronize(t.b));
    float ms = 0.0f;
    BQA_CUDA_CHECK(cudaEventElapsedTime(&ms, t.a, t.b));
    host_ms[t.kind] += ms;
    host_flops[t.kind] += t.flops;
    host_count[t.kind] += 1;
  }
  return 0;
}

int bqa_bind(bqa_handle* h, float* params, float* grads, void* params_bf16, void* workspace,
             uint64_t workspace_bytes) {
  if (h == nullptr || params == nullptr || params_bf16 == nullptr || workspace == nullptr) {
    set_error("bqa_bind: null argument");
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) |
       reinterpret_cast<uintptr_t>(params_bf16) | reinterpret_cast<uintptr_t>(workspace)) & 255u) {
    set_error("bqa_bind: buffers must be 256-byte aligned");
    return 1;
  }
  h->params = params;
  h->grads = grads;
  h->params_bf16 = reinterpret_cast<bf16*>(params_bf16);
  h->workspace = reinterpret_cast<uint8_t*>(workspace);
  h->workspace_bytes = workspace_bytes;
  h->plan_valid = false;
  return 0;
}

int bqa_forward(bqa_handle* h, const int64_t* input_ids, const int64_t* token_type_ids,
                const uint8_t* attention_mask, int32_t B, int32_t S, int32_t training, uint64_t seed,
                uint64_t step, float* start, float* end, float* cls, float* reg_start, float* reg_end,
                void* stream) {
  if (h == nullptr || h->params == nullptr) {
    set_error("bqa_forward: handle not bound");
    return 1;
  }
  if (B < 1 || S < 1 || S > 512 || S > h->cfg.max_pos) {
    set_error("bqa_forward: B=%d S=%d out of range (S <= min(512, max_pos=%d))", B, S, h->cfg.max_pos);
    return 1;
  }
  if (training && h->grads == nullptr) {
    set_error("bqa_forward: training forward needs a bound gradient buffer");
    return 1;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool tr = training != 0;
  if (!h->plan_valid || h->plan.B != B || h->plan.S != S || h->plan.training != tr) {
    if (build_plan(h, B, S, tr)) return 1;
  }
  h->seed = seed;
  h->step = step;
  h->last_training = tr;
  h->enc_dropout_off = (training & 2) != 0;
  h->cls_dropout_off = (training & 4) != 0;
  h->cls_site = site_cls(h->cfg.num_layers);
  const bqa_config& c = h->cfg;
  const bqa_layout& Y = h->lay;
  Workspace& w = h->plan.ws;
  const int T = B * S, H = c.hidden;
  const size_t tb = static_cast<size_t>(T);

  // The private copies of the inputs that the backward needs (embedding scatter, key mask) are written by the
  // embedding kernel itself; only the error flag is cleared here.
  const uint8_t* mask = attention_mask != nullptr ? w.mask : nullptr;
  h->has_mask = mask != nullptr;

  const bool attn_drop = tr && c.attn_dropout > 0.0f && !h->enc_dropout_off;
  const bool ahead = attn_drop && h->bits_ahead && h->bits_seed == seed && h->bits_step == step && h->bits_B == B &&
                     h->bits_S == S;
  if (h->side_pending && !ahead) {
    // keep bits generated ahead of time that this forward will not consume (other shape, plan rebuilt, eval,
    // dropout off): those side-stream kernels may still be writing into the workspace that the kernels below
    // are about to reuse — order them first
    BQA_CUDA_CHECK(cudaStreamWaitEvent(st, h->ev_tail, 0));
  }
  h->side_pending = false;  // consumed bits are ordered by the per-layer events
  BQA_CUDA_CHECK(cudaMemsetAsync(w.err_flag, 0, sizeof(int), st));
  if (attn_drop) {
    if (h->side == nullptr) {
      BQA_CUDA_CHECK(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
      BQA_CUDA_CHECK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
      BQA_CUDA_CHECK(cudaEventCreateWithFlags(&h->ev_tail, cudaEventDisableTiming));
      h->ev_bits.resize(c.num_layers);
      for (auto& e : h->ev_bits) BQA_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (!ahead) {
      BQA_CUDA_CHECK(cudaEventRecord(h->ev_fork, st));  // everything that read last step's bits is behind us
      BQA_CUDA_CHECK(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
      for (int l = 0; l < c.num_layers; ++l) {
        if (attn_make_keep_bits(h->plan.layer[l].attn, h->key(site_attn(l), c.attn_dropout), h->side)) return 1;
        BQA_CUDA_CHECK(cudaEventRecord(h->ev_bits[l], h->side));
      }
      h->launches += c.num_layers;
    }
  }
  h->bits_ahead = false;
  // algorithmic bytes: three fp32 table rows in, two bf16 rows out per token
  h->time_begin(7, static_cast<double>(T) * H * (3 * 4 + 2 * 2), st);
  if (launch_embed_ln_fwd(input_ids, token_type_ids, attention_mask, h->P(Y.word_emb), h->P(Y.pos_emb),
                          h->P(Y.type_emb), h->P(Y.emb_ln_g), h->P(Y.emb_ln_b), w.z0, w.x0, w.ids, w.tt, w.mask,
                          w.pos_ids, T, S, H, c.vocab, c.type_vocab, c.max_pos, c.position_mode, c.pad_token_id, c.ln_eps,
                          h->key(site_embed(), c.hidden_dropout), w.err_flag, st))
    return 1;
  h->time_end(st);
  h->launches += 1;
  const int slots = static_cast<int>(w.layer.size());
  const DropKey nodrop = make_drop_key(0, 0, 0.0f);
  for (int l = 0; l < c.num_layers; ++l) {
    LayerPlans& lp = h->plan.layer[l];
    const LayerActs& act = w.layer[l % slots];
    const int64_t lo = h->layer_off(l);
    if (launch_gemm(h, lp.qkv, nodrop, st)) return 1;
    if (attn_drop) BQA_CUDA_CHECK(cudaStreamWaitEvent(st, h->ev_bits[l], 0));
    h->time_begin(3, attn_flops(h, B, S, false), st);
    if (attn_forward(lp.attn, mask, h->key(site_attn(l), c.attn_dropout), st, /*bits_ready=*/true)) return 1;
    h->time_end(st);
    if (launch_gemm(h, lp.ao, h->key(site_ao(l), c.hidden_dropout), st)) return 1;
    const double ln_fwd_bytes = static_cast<double>(T) * H * 2 * 2;  // one bf16 row in, one out
    h->time_begin(5, ln_fwd_bytes, st);
    if (launch_ln_fwd(act.z1, h->P(lo + Y.ln1_g), h->P(lo + Y.ln1_b), act.a, T, H, c.ln_eps, st)) return 1;
    h->time_end(st);
    if (launch_gemm(h, lp.ffn1, nodrop, st)) return 1;
    if (launch_gemm(h, lp.ffn2, h->key(site_ffn(l), c.hidden_dropout), st)) return 1;
    h->time_begin(5, ln_fwd_bytes, st);
    if (launch_ln_fwd(act.z2, h->P(lo + Y.ln2_g), h->P(lo + Y.ln2_b), act.x_out, T, H, c.ln_eps, st)) return 1;
    h->time_end(st);
    h->launches += 3;
  }
  const bf16* y = w.layer[(c.num_layers - 1) % slots].x_out;
  const HeadParams hp = head_params(h);
  // the outputs go straight into the caller's buffers; the backward keeps its own copy of the sigmoid outputs
  if (launch_span_head_fwd(y, hp.span_w, hp.span_b, start, end, T, H, st)) return 1;
  if (launch_pooler_heads_fwd(y, hp, w.pooled, cls, reg_start, reg_end, w.reg_s, w.reg_e, B, S, H, c.num_labels,
                              h->key(site_cls(c.num_layers), c.hidden_dropout), st))
    return 1;
  h->launches += 3;
  return 0;
}

static int to_loss_config(const bqa_loss_config* lc, LossConfig* k) {
  if (lc == nullptr) {
    set_error("loss: null configuration");
    return 1;
  }
  if (lc->cls_loss != BQA_CLS_SMOOTH && lc->cls_loss != BQA_CLS_CE && lc->cls_loss != BQA_CLS_FOCAL) {
    set_error("loss: unsupported class loss %d", lc->cls_loss);
    return 1;
  }
  k->w_start = lc->w_start; k->w_end = lc->w_end;
  k->w_start_reg = lc->w_start_reg; k->w_end_reg = lc->w_end_reg; k->w_cls = lc->w_cls;
  k->cls_kind = lc->cls_loss == BQA_CLS_CE ? CLS_LOSS_CE : (lc->cls_loss == BQA_CLS_FOCAL ? CLS_LOSS_FOCAL : CLS_LOSS_SMOOTH);
  k->focal_alpha = lc->focal_alpha;
  k->focal_gamma = lc->focal_gamma;
  k->smooth_alpha = lc->smooth_alpha;
  k->cls_weight = lc->cls_weight;
  k->grad_scale = lc->loss_scale;
  return 0;
}

int bqa_qa_loss(const float* start, const float* end, const float* cls, const float* reg_start,
                const float* reg_end, const int64_t* y_start, const int64_t* y_end, const int64_t* y_cls,
                const float* y_reg_start, const float* y_reg_end, float* d_start, float* d_end, float* d_cls,
                float* d_reg_start, float* d_reg_end, float* row_scratch, float* losses, int32_t B, int32_t S,
                int32_t C, const bqa_loss_config* lc, int32_t with_grad, void* stream) {
  LossConfig k;
  if (to_loss_config(lc, &k)) return 1;
  if (B < 1 || S < 1 || C < 1 || C > 32) {
    set_error("bqa_qa_loss: bad shape B=%d S=%d C=%d", B, S, C);
    return 1;
  }
  if (with_grad && (d_start == nullptr || d_end == nullptr || d_cls == nullptr || d_reg_start == nullptr ||
                    d_reg_end == nullptr)) {
    set_error("bqa_qa_loss: gradient buffers missing");
    return 1;
  }
  return launch_loss_fwd_bwd(start, end, cls, reg_start, reg_end, y_start, y_end, y_cls, y_reg_start, y_reg_end,
                             d_start, d_end, d_cls, d_reg_start, d_reg_end, row_scratch, losses, B, S, C, k,
                             with_grad, reinterpret_cast<cudaStream_t>(stream));
}

int bqa_loss(bqa_handle* h, const float* start, const float* end, const float* cls, const float* reg_start,
             const float* reg_end, const int64_t* y_start, const int64_t* y_end, const int64_t* y_cls,
             const float* y_reg_start, const float* y_reg_end, const bqa_loss_config* lc, int32_t with_grad,
             float* losses, void* stream) {
  if (h == nullptr || !h->plan_valid || lc == nullptr) {
    set_error("bqa_loss: no forward has been run on this handle");
    return 1;
  }
  if (with_grad && !h->plan.training) {
    set_error("bqa_loss: gradients requested but the last forward was not a training forward");
    return 1;
  }
  Workspace& w = h->plan.ws;
  LossConfig k;
  if (to_loss_config(lc, &k)) return 1;
  h->launches += 2;
  return launch_loss_fwd_bwd(start, end, cls, reg_start, reg_end, y_start, y_end, y_cls, y_reg_start, y_reg_end,
                             w.d_start, w.d_end, w.d_cls, w.d_reg_s, w.d_reg_e, w.loss_rows, losses, h->plan.B,
                             h->plan.S, h->cfg.num_labels, k, with_grad, reinterpret_cast<cudaStream_t>(stream),
                             w.err_flag);
}

int bqa_input_status(bqa_handle* h, int32_t* host_status, void* stream) {
  if (h == nullptr || host_status == nullptr || !h->plan_valid) {
    set_error("bqa_input_status: no forward has been run on this handle");
    return 1;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int flag = 0;
  BQA_CUDA_CHECK(cudaMemcpyAsync(&flag, h->plan.ws.err_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
  BQA_CUDA_CHECK(cudaStreamSynchronize(st));
  *host_status = flag;
  return 0;
}

int bqa_set_output_grads(bqa_handle* h, const float* d_start, const float* d_end, const float* d_cls,
                         const float* d_reg_start, const float* d_reg_end, void* stream) {
  if (h == nullptr || !h->plan_valid || !h->plan.training) {
    set_error("bqa_set_output_grads: the last forward was not a training forward");
    return 1;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Workspace& w = h->plan.ws;
  const size_t T = static_cast<size_t>(h->plan.B) * h->plan.S, B = h->plan.B, C = h->cfg.num_labels;
  auto put = [&](float* dst, const float* src, size_t n) -> int {
    if (src != nullptr)
      BQA_CUDA_CHECK(cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st));
    else
      BQA_CUDA_CHECK(cudaMemsetAsync(dst, 0, n * sizeof(float), st));
    return 0;
  };
  if (put(w.d_start, d_start, T) || put(w.d_end, d_end, T) || put(w.d_cls, d_cls, B * C) ||
      put(w.d_reg_s, d_reg_start, B) || put(w.d_reg_e, d_reg_end, B))
    return 1;
  return 0;
}

int bqa_num_backward_phases(const bqa_handle* h) { return h == nullptr ? 0 : h->cfg.num_layers + 2; }

int64_t bqa_phase_grad_begin(const bqa_handle* h, int32_t phase) {
  if (h == nullptr) return -1;
  const int L = h->cfg.num_layers;
  if (phase <= 0) return h->lay.pool_w;
  if (phase <= L) return h->layer_off(L - phase);
  return 0;
}

int64_t bqa_launch_count(const bqa_handle* h) { return h == nullptr ? 0 : h->launches; }

int bqa_backward(bqa_handle* h, int32_t first_phase, int32_t last_phase, void* stream) {
  if (h == nullptr || !h->plan_valid || !h->plan.training || !h->last_training) {
    set_error("bqa_backward: the last forward was not a training forward");
    return 1;
  }
  const bqa_config& c = h->cfg;
  const bqa_layout& Y = h->lay;
  const int L = c.num_layers;
  if (first_phase < 0 || last_phase > L + 1 || first_phase > last_phase) {
    set_error("bqa_backward: phase range [%d, %d] outside [0, %d]", first_phase, last_phase, L + 1);
    return 1;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  Workspace& w = h->plan.ws;
  const int B = h->plan.B, S = h->plan.S, T = B * S, H = c.hidden, I = c.intermediate;
  const DropKey nodrop = make_drop_key(0, 0, 0.0f);
  const uint8_t* mask = h->has_mask ? w.mask : nullptr;

  for (int phase = first_phase; phase <= last_phase; ++phase) {
    if (phase == 0) {
      // gradient of the last hidden states comes out of the heads in slot dy[(L) & 1]
      bf16* dy = (L & 1) ? w.dy1 : w.dy0;
      const bf16* y = w.layer[L - 1].x_out;
      if (launch_heads_bwd(y, w.pooled, w.reg_s, w.reg_e, w.d_start, w.d_end, w.d_cls, w.d_reg_s, w.d_reg_e,
                           head_params(h), head_grads(h), w.dpre, w.dx0, dy, B, S, H, c.num_labels,
                           h->key(site_cls(L), c.hidden_dropout), h->num_sms, st))
        return 1;
      h->launches += 4;
    } else if (phase <= L) {
      const int l = L - phase;
      LayerPlans& lp = h->plan.layer[l];
      const LayerActs& act = w.layer[l];
      const int64_t lo = h->layer_off(l);
      const bf16* dy_in = ((l + 1) & 1) ? w.dy1 : w.dy0;  // gradient w.r.t. this layer's output
      const bool hd = c.hidden_dropout > 0.0f;
      // LN2 backward: dres = d z2 (residual branch), dgm = d (ffn2 GEMM output) = dres * keep-mask
      const double ln_bwd_bytes = static_cast<double>(T) * H * 2 * (hd ? 4 : 3);  // dy, z in; dz (+ masked dz) out
      h->time_begin(6, ln_bwd_bytes, st);
      if (launch_ln_bwd(dy_in, act.z2, h->P(lo + Y.ln2_g), w.dres, hd ? w.dgm : w.dres, h->G(lo + Y.ln2_g),
                        h->G(lo + Y.ln2_b), h->G(lo + Y.o_b), T, H, c.ln_eps, nodrop,
                        h->key(site_ffn(l), c.hidden_dropout), h->num_sms, st))
        return 1;
      h->time_end(st);
      if (launch_gemm(h, lp.w_ffn2, nodrop, st)) return 1;
      if (launch_gemm(h, lp.d_ffn2, nodrop, st)) return 1;  // du = (dgm . W2) * gelu'(u)
      if (launch_gemm(h, lp.w_ffn1, nodrop, st)) return 1;
      if (launch_gemm(h, lp.d_ffn1, nodrop, st)) return 1;  // da = du . W1 + dres   (into the dctx slab)
      // LN1 backward on da
      h->time_begin(6, ln_bwd_bytes, st);
      if (launch_ln_bwd(w.dctx, act.z1, h->P(lo + Y.ln1_g), w.dres, hd ? w.dgm : w.dres, h->G(lo + Y.ln1_g),
                        h->G(lo + Y.ln1_b), h->G(lo + Y.ao_b), T, H, c.ln_eps, nodrop,
                        h->key(site_ao(l), c.hidden_dropout), h->num_sms, st))
        return 1;
      h->time_end(st);
      if (launch_gemm(h, lp.w_ao, nodrop, st)) return 1;
      if (launch_gemm(h, lp.d_ao, nodrop, st)) return 1;  // dctx = dgm . Wo
      h->time_begin(4, attn_flops(h, B, S, true), st);
      if (attn_backward(lp.attn, mask, h->key(site_attn(l), c.attn_dropout), st, /*dsum_ready=*/true,
                        /*dqkv_bias_grad=*/h->G(lo + Y.qkv_b)))
        return 1;
      h->time_end(st);
      if (launch_gemm(h, lp.w_qkv, nodrop, st)) return 1;
      if (launch_gemm(h, lp.d_qkv, nodrop, st)) return 1;  // dx = dqkv . Wqkv + dres
      h->launches += 4;  // 2 LayerNorm backward, 2 attention backward (GEMMs count themselves)
    } else {
      // embeddings: x0 = dropout(LN(z0)); dz0 scattered into the three tables
      const bf16* dy_in = w.dy0;  // layer 0 wrote its input gradient into dy[0 & 1]
      h->time_begin(6, static_cast<double>(T) * H * 2 * 3, st);
      if (launch_ln_bwd(dy_in, w.z0, h->P(Y.emb_ln_g), w.dres, nullptr, h->G(Y.emb_ln_g), h->G(Y.emb_ln_b), nullptr,
                        T, H, c.ln_eps, h->key(site_embed(), c.hidden_dropout), nodrop, h->num_sms, st))
        return 1;
      h->time_end(st);
      // one bf16 row in per token for each of the two kernels; one fp32 table row read-modify-written per token
      h->time_begin(8, static_cast<double>(T) * H * (2 * 2 + 2 * 4), st);
      if (launch_embed_bwd(w.dres, w.ids, w.tt, w.pos_ids, h->G(Y.word_emb), h->G(Y.pos_emb), h->G(Y.type_emb), T, S, H,
                           c.vocab, c.type_vocab, c.max_pos, c.position_mode, c.pad_token_id, st))
        return 1;
      h->time_end(st);
      h->launches += 2;
    }
  }
  // The attention-dropout decisions depend only on (seed, step, site): once layer 0's attention backward
  // has read this step's bits, the next step's are generated on the side stream, where they overlap the
  // rest of this step (embedding backward, gradient exchange, clip + AdamW: HBM-bound work that leaves the
  // issue slots idle) instead of the next forward's GEMMs.  bqa_forward uses them if it is called with the
  // predicted (seed, step + 1, B, S) and generates its own otherwise.
  if (last_phase >= L && first_phase <= L && c.attn_dropout > 0.0f && !h->enc_dropout_off && h->side != nullptr) {
    BQA_CUDA_CHECK(cudaEventRecord(h->ev_fork, st));
    BQA_CUDA_CHECK(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
    for (int l = 0; l < L; ++l) {
      if (attn_make_keep_bits(h->plan.layer[l].attn, h->key_at(h->seed, h->step + 1, site_attn(l), c.attn_dropout),
                              h->side))
        return 1;
      BQA_CUDA_CHECK(cudaEventRecord(h->ev_bits[l], h->side));
    }
    BQA_CUDA_CHECK(cudaEventRecord(h->ev_tail, h->side));
    h->side_pending = true;
    h->launches += L;
    h->bits_ahead = true;
    h->bits_seed = h->seed;
    h->bits_step = h->step + 1;
    h->bits_B = B;
    h->bits_S = S;
  }
  return 0;
}

}  // extern "C"
