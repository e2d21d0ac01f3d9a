// heads_loss.cu — pooler, the four QA heads, the weighted multi-task loss (forward and the
// gradients it sends into the heads), the heads' backward and validate.py-style span decoding.
//
// Reference:
//   BertPooler.forward                site-packages/transformers/models/bert/modeling_bert.py:462-468
//   heads                             modules/model/model/model.py:30-41,54-65
//   WeightedLoss / losses             modules/model/model/loss.py:5-38,74-100, modules/init.py:18-40
//   decode                            modules/model/inference/predictor.py:114-120
// All of this is < 0.01 % of the step's FLOPs; the kernels are latency/HBM-bound warp-shuffle
// reductions in fp32 on fp32 master weights (logits stay fp32 so that argmax ties resolve as
// in the reference: first maximal index).
#include "kernels.cuh"

namespace bqa {

namespace {

__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = 0.0f;
  for (int w = 0; w < nw; ++w) s += red[w];  // fixed order: deterministic
  return s;
}

__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = red[0];
  for (int w = 1; w < nw; ++w) s = fmaxf(s, red[w]);
  return s;
}

// dot of a bf16 row with an fp32 vector, warp-cooperative; H % 8 == 0
__device__ __forceinline__ float warp_dot_bf16_f32(const bf16* __restrict__ x, const float* __restrict__ w, int H,
                                                   int lane) {
  float acc = 0.0f;
  for (int c = lane * 8; c < H; c += 256) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + c));
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + c));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + c + 4));
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    acc += a.x * w0.x + a.y * w0.y + b.x * w0.z + b.y * w0.w + cc.x * w1.x + cc.y * w1.y + d.x * w1.z + d.y * w1.w;
  }
  return warp_sum(acc);
}

// ---------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
span_head_fwd_kernel(const bf16* __restrict__ y, const float* __restrict__ w, const float* __restrict__ b,
                     float* __restrict__ start, float* __restrict__ end, int T, int H) {
  const int lane = threadIdx.x & 31;
  const int t = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (t >= T) return;
  const bf16* row = y + static_cast<size_t>(t) * H;
  const float s = warp_dot_bf16_f32(row, w, H, lane);
  const float e = warp_dot_bf16_f32(row, w + H, H, lane);
  if (lane == 0) {
    start[t] = s + b[0];
    end[t] = e + b[1];
  }
}

// pooled[b][j] = tanh(y[b*S][:] . Wp[j][:] + bp[j]);  grid (ceil(H/32), B), 8 warps x 4 rows
__global__ void __launch_bounds__(256)
pooler_fwd_kernel(const bf16* __restrict__ y, const float* __restrict__ Wp, const float* __restrict__ bp,
                  float* __restrict__ pooled, int S, int H) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const bf16* x = y + static_cast<size_t>(b) * S * H;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int j = blockIdx.x * 32 + warp * 4 + r;
    if (j < H) {
      const float v = warp_dot_bf16_f32(x, Wp + static_cast<size_t>(j) * H, H, lane);
      if (lane == 0) pooled[static_cast<size_t>(b) * H + j] = tanhf(v + bp[j]);
    }
  }
}

// cls[b][c] = dropout(pooled[b]) . Wc[c] + bc[c]; reg = sigmoid(pooled[b] . w + b);  grid B
__global__ void __launch_bounds__(256)
heads_fwd_kernel(const float* __restrict__ pooled, HeadParams hp, float* __restrict__ cls, float* __restrict__ reg_s,
                 float* __restrict__ reg_e, float* __restrict__ reg_s2, float* __restrict__ reg_e2, int H, int C,
                 DropKey drop) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int b = blockIdx.x;
  const float* p = pooled + static_cast<size_t>(b) * H;
  for (int o = warp; o < C + 2; o += nw) {
    const float* w = o < C ? hp.cls_w + static_cast<size_t>(o) * H : (o == C ? hp.regs_w : hp.rege_w);
    float acc = 0.0f;
    for (int j = lane; j < H; j += 32) {
      float pv = p[j];
      if (o < C && drop.thresh != 0u) pv = drop_keep(drop, static_cast<uint32_t>(b) * H + j) ? pv * drop.scale : 0.0f;
      acc += pv * w[j];
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      if (o < C) {
        cls[static_cast<size_t>(b) * C + o] = acc + hp.cls_b[o];
      } else {
        const float z = acc + (o == C ? hp.regs_b[0] : hp.rege_b[0]);
        const float sg = 1.0f / (1.0f + __expf(-z));
        (o == C ? reg_s : reg_e)[b] = sg;
        float* second = o == C ? reg_s2 : reg_e2;
        if (second != nullptr) second[b] = sg;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// loss: one block per batch row writes its five contributions; a 1-block kernel adds them
// in a fixed order (deterministic) and forms the weighted total.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
loss_rows_kernel(const float* __restrict__ start, const float* __restrict__ end, const float* __restrict__ cls,
                 const float* __restrict__ reg_s, const float* __restrict__ reg_e, const int64_t* __restrict__ y_start,
                 const int64_t* __restrict__ y_end, const int64_t* __restrict__ y_cls,
                 const float* __restrict__ y_reg_s, const float* __restrict__ y_reg_e, float* __restrict__ d_start,
                 float* __restrict__ d_end, float* __restrict__ d_cls, float* __restrict__ d_reg_s,
                 float* __restrict__ d_reg_e, float* __restrict__ row_scratch, int B, int S, int C, LossConfig lc,
                 int with_grad) {
  __shared__ float red[8];
  __shared__ float s_cnt[3];
  const int b = blockIdx.x;
  // grid (B, 3): blockIdx.y = 0 / 1: start / end cross entropy of row b; 2: the regressions and the class loss —
  // three short dependent chains side by side instead of one long one (the kernel is pure latency)
  const int part = blockIdx.y;
  // denominator of the part's mean-reduced classification loss
  if (threadIdx.x < 32) {
    float cnt = 0.0f;
    for (int i = threadIdx.x; i < B; i += 32) {
      if (part == 0) cnt += (y_start[i] != -1) ? 1.0f : 0.0f;
      if (part == 1) cnt += (y_end[i] != -1) ? 1.0f : 0.0f;
      if (part == 2) {
        const long long yc = y_cls[i];
        if (lc.cls_kind == CLS_LOSS_FOCAL) {
          if (yc != -1) cnt += 1.0f;  // NLLLoss(ignore_index=-1)
        } else if (yc != -100) {
          cnt += (lc.cls_kind == CLS_LOSS_CE && lc.cls_weight != nullptr && yc >= 0 && yc < C) ? lc.cls_weight[yc]
                                                                                                 : 1.0f;
        }
      }
    }
    cnt = warp_sum(cnt);
    if (threadIdx.x == 0) s_cnt[part] = cnt;
  }
  __syncthreads();

  // ---- start / end : CrossEntropyLoss(ignore_index=-1) over the S positions (init.py:34-35)
  if (part < 2) {
    const int which = part;
    const float* logit = (which == 0 ? start : end) + static_cast<size_t>(b) * S;
    float* dl = (which == 0 ? d_start : d_end) + static_cast<size_t>(b) * S;
    const long long y = (which == 0 ? y_start : y_end)[b];
    const float w = which == 0 ? lc.w_start : lc.w_end;
    const float nvalid = s_cnt[which];
    float mx = -INFINITY;
    for (int s = threadIdx.x; s < S; s += blockDim.x) mx = fmaxf(mx, logit[s]);
    mx = block_max(mx, red);
    float se = 0.0f;
    for (int s = threadIdx.x; s < S; s += blockDim.x) se += __expf(logit[s] - mx);
    se = block_sum(se, red);
    const float lse = mx + __logf(se);
    const bool valid = (y != -1);
    if (threadIdx.x == 0) {
      float l = 0.0f;
      if (valid && y >= 0 && y < S) l = (lse - logit[y]) / nvalid;
      if (nvalid == 0.0f) l = __int_as_float(0x7fc00000);  // torch: mean over zero targets is NaN
      row_scratch[b * 5 + which] = l;
    }
    if (with_grad) {
      const float gs = valid ? w * lc.grad_scale / nvalid : 0.0f;
      for (int s = threadIdx.x; s < S; s += blockDim.x) {
        const float p = __expf(logit[s] - lse);
        dl[s] = gs * (p - ((valid && s == y) ? 1.0f : 0.0f));
      }
    }
    return;
  }

  // ---- regressions : MSELoss on the sigmoid outputs (init.py:36-37)
  if (threadIdx.x == 0) {
    for (int which = 0; which < 2; ++which) {
      const float sg = (which == 0 ? reg_s : reg_e)[b];
      const float t = (which == 0 ? y_reg_s : y_reg_e)[b];
      const float w = which == 0 ? lc.w_start_reg : lc.w_end_reg;
      const float d = sg - t;
      row_scratch[b * 5 + 2 + which] = d * d / static_cast<float>(B);
      if (with_grad) (which == 0 ? d_reg_s : d_reg_e)[b] = w * lc.grad_scale * 2.0f * d / static_cast<float>(B);
    }
    // ---- answer-type class loss (init.py:22-31)
    const float* z = cls + static_cast<size_t>(b) * C;
    float mx = -INFINITY;
    for (int c = 0; c < C; ++c) mx = fmaxf(mx, z[c]);
    float se = 0.0f;
    for (int c = 0; c < C; ++c) se += __expf(z[c] - mx);
    const float lse = mx + __logf(se);
    const long long y = y_cls[b];
    float l = 0.0f;
    if (lc.cls_kind == CLS_LOSS_SMOOTH && lc.smooth_alpha > 0.0f) {
      // LabelSmoothingLossWithLogits (loss.py:5-38): KLDivLoss(batchmean) against
      // fill = alpha / (C - 1) everywhere, 1 - alpha at the label
      const float fill = lc.smooth_alpha / static_cast<float>(C - 1);
      const float conf = 1.0f - lc.smooth_alpha;
      float tsum = 0.0f;
      for (int c = 0; c < C; ++c) {
        const float t = (c == y) ? conf : fill;
        tsum += t;
        if (t > 0.0f) l += t * (__logf(t) - (z[c] - lse));
      }
      l /= static_cast<float>(B);
      if (with_grad)
        for (int c = 0; c < C; ++c) {
          const float t = (c == y) ? conf : fill;
          d_cls[static_cast<size_t>(b) * C + c] =
              lc.w_cls * lc.grad_scale * (__expf(z[c] - lse) * tsum - t) / static_cast<float>(B);
        }
    } else if (lc.cls_kind == CLS_LOSS_FOCAL) {
      // FocalLossWithLogits (loss.py:57-71): NLLLoss(ignore_index=-1, mean) of alpha (1 - p)^gamma log p
      //   f(p_y) = -alpha (1 - p_y)^gamma log p_y ;  d f / d z_c = f'(p_y) p_y (delta_cy - p_c)
      const bool valid = (y != -1) && y >= 0 && y < C;
      const float den = s_cnt[2];
      float fprime_p = 0.0f;
      if (valid) {
        const float logp = z[y] - lse;
        const float py = __expf(logp);
        const float om = fmaxf(1.0f - py, 0.0f);
        const float pw = __powf(om, lc.focal_gamma);             // (1 - p)^gamma
        const float pw1 = lc.focal_gamma == 0.0f ? 0.0f : lc.focal_gamma * __powf(om, lc.focal_gamma - 1.0f);
        l = -lc.focal_alpha * pw * logp / den;
        // f'(p) p = alpha (gamma (1 - p)^(gamma - 1) p log p - (1 - p)^gamma)
        fprime_p = lc.focal_alpha * (pw1 * py * logp - pw);
      }
      if (den == 0.0f) l = __int_as_float(0x7fc00000);
      if (with_grad)
        for (int c = 0; c < C; ++c)
          d_cls[static_cast<size_t>(b) * C + c] =
              valid ? lc.w_cls * lc.grad_scale * fprime_p * ((c == y ? 1.0f : 0.0f) - __expf(z[c] - lse)) / den : 0.0f;
    } else {
      // nn.CrossEntropyLoss(weight) / NLLLoss(ignore_index=-100): weighted mean over valid rows
      const bool valid = (y != -100) && y >= 0 && y < C;
      const float wy = (valid && lc.cls_kind == CLS_LOSS_CE && lc.cls_weight != nullptr) ? lc.cls_weight[y] : 1.0f;
      const float den = s_cnt[2];
      if (valid) l = wy * (lse - z[y]) / den;
      if (den == 0.0f) l = __int_as_float(0x7fc00000);
      if (with_grad)
        for (int c = 0; c < C; ++c)
          d_cls[static_cast<size_t>(b) * C + c] =
              valid ? lc.w_cls * lc.grad_scale * wy * (__expf(z[c] - lse) - (c == y ? 1.0f : 0.0f)) / den : 0.0f;
    }
    row_scratch[b * 5 + 4] = l;
  }
}

__global__ void loss_finalize_kernel(const float* __restrict__ row_scratch, float* __restrict__ losses, int B,
                                     LossConfig lc, const int* __restrict__ err_flag) {
  if (threadIdx.x < 5) {
    float s = 0.0f;
    for (int b = 0; b < B; ++b) s += row_scratch[b * 5 + threadIdx.x];
    losses[threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // same accumulation order as WeightedLoss.__call__ (loss.py:83-96): dict order of init_losses
    float total = 0.0f;
    total += lc.w_start * losses[0];
    total += lc.w_end * losses[1];
    total += lc.w_start_reg * losses[2];
    total += lc.w_end_reg * losses[3];
    total += lc.w_cls * losses[4];
    losses[5] = total;
    if (err_flag != nullptr && *err_flag != 0) {
#pragma unroll
      for (int i = 0; i < 6; ++i) losses[i] = __int_as_float(0x7fc00000);
    }
  }
}

// ---------------------------------------------------------------------------------------
// backward of the heads
// ---------------------------------------------------------------------------------------
// grid B: dpre[b][j] = (dcls . Wc[:,j] * keep + drs * wrs[j] + dre * wre[j]) * (1 - pooled^2)
// and the head weight gradients (B-way atomics on C*H + 2H addresses)
__global__ void __launch_bounds__(256)
pooled_bwd_kernel(const float* __restrict__ pooled, const float* __restrict__ reg_s,
                  const float* __restrict__ reg_e, const float* __restrict__ d_cls, const float* __restrict__ d_rs,
                  const float* __restrict__ d_re, HeadParams hp, HeadGrads hg, float* __restrict__ dpre, int H,
                  int C, DropKey drop) {
  const int b = blockIdx.x;
  // through the sigmoid: d/dz = d/dsigma * sigma * (1 - sigma)
  const float drs = d_rs[b] * reg_s[b] * (1.0f - reg_s[b]);
  const float dre = d_re[b] * reg_e[b] * (1.0f - reg_e[b]);
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    const float pv = pooled[static_cast<size_t>(b) * H + j];
    float keep = 1.0f;
    if (drop.thresh != 0u) keep = drop_keep(drop, static_cast<uint32_t>(b) * H + j) ? drop.scale : 0.0f;
    float dp = drs * hp.regs_w[j] + dre * hp.rege_w[j];
    for (int c = 0; c < C; ++c) {
      const float dc = d_cls[static_cast<size_t>(b) * C + c];
      dp += dc * hp.cls_w[static_cast<size_t>(c) * H + j] * keep;
      atomicAdd(hg.cls_w + static_cast<size_t>(c) * H + j, dc * pv * keep);
    }
    atomicAdd(hg.regs_w + j, drs * pv);
    atomicAdd(hg.rege_w + j, dre * pv);
    dpre[static_cast<size_t>(b) * H + j] = dp * (1.0f - pv * pv);
  }
  if (threadIdx.x < C) atomicAdd(hg.cls_b + threadIdx.x, d_cls[static_cast<size_t>(b) * C + threadIdx.x]);
  if (threadIdx.x == 0) {
    atomicAdd(hg.regs_b, drs);
    atomicAdd(hg.rege_b, dre);
  }
}

// dx0[b][i] += sum_{j in slice} dpre[b][j] * Wp[j][i] for every batch row b at once, so that Wp is read once.
// grid (ceil(H/256), ceil(H/PJ)); dx0 is zeroed by the caller; batch rows are taken PB at a time.
constexpr int PJ = 32;  // j rows of Wp per block
constexpr int PB = 32;  // batch rows per pass
__global__ void __launch_bounds__(256)
pooler_dx_kernel(const float* __restrict__ dpre, const float* __restrict__ Wp, float* __restrict__ dx0, int B,
                 int H) {
  __shared__ float s_d[PB][PJ];
  const int i = blockIdx.x * 256 + threadIdx.x;
  const int j0 = blockIdx.y * PJ;
  float w[PJ];
#pragma unroll
  for (int j = 0; j < PJ; ++j) w[j] = (i < H && j0 + j < H) ? __ldg(Wp + static_cast<size_t>(j0 + j) * H + i) : 0.0f;
  for (int b0 = 0; b0 < B; b0 += PB) {
    __syncthreads();
    for (int e = threadIdx.x; e < PB * PJ; e += 256) {
      const int b = b0 + e / PJ, j = j0 + e % PJ;
      s_d[e / PJ][e % PJ] = (b < B && j < H) ? dpre[static_cast<size_t>(b) * H + j] : 0.0f;
    }
    __syncthreads();
    if (i < H) {
      for (int b = 0; b < PB && b0 + b < B; ++b) {
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < PJ; ++j) acc += s_d[b][j] * w[j];
        atomicAdd(dx0 + static_cast<size_t>(b0 + b) * H + i, acc);
      }
    }
  }
}

// dWp[j][i] += sum_b dpre[b][j] * x0[b][i], dbp[j] += sum_b dpre[b][j], x0[b] = [CLS] row of batch b.
// grid (ceil(H/256), ceil(H/PJ)): the block keeps its 256 columns of the [CLS] rows in registers-by-batch
// through shared memory and produces PJ rows of dWp.
__global__ void __launch_bounds__(256)
pooler_wgrad_kernel(const float* __restrict__ dpre, const bf16* __restrict__ y, float* __restrict__ dWp,
                    float* __restrict__ dbp, int B, int S, int H) {
  __shared__ __align__(16) float s_d[PB][PJ];
  const int i = blockIdx.x * 256 + threadIdx.x;
  const int j0 = blockIdx.y * PJ;
  float acc[PJ];
#pragma unroll
  for (int j = 0; j < PJ; ++j) acc[j] = 0.0f;
  float bsum = 0.0f;  // thread j < PJ of the first column block: bias gradient of row j0 + j
  for (int b0 = 0; b0 < B; b0 += PB) {
    __syncthreads();
    for (int e = threadIdx.x; e < PB * PJ; e += 256) {
      const int b = b0 + e / PJ, j = j0 + e % PJ;
      s_d[e / PJ][e % PJ] = (b < B && j < H) ? dpre[static_cast<size_t>(b) * H + j] : 0.0f;
    }
    __syncthreads();
    // all [CLS] elements of the pass first: the loads are independent, one latency instead of PB
    float xs[PB];
#pragma unroll
    for (int b = 0; b < PB; ++b)
      xs[b] = (i < H && b0 + b < B) ? __bfloat162float(y[static_cast<size_t>(b0 + b) * S * H + i]) : 0.0f;
#pragma unroll
    for (int b = 0; b < PB; ++b) {  // rows beyond B are zeros in s_d and xs
#pragma unroll
      for (int j = 0; j < PJ; j += 4) {  // 16-byte broadcast reads: the loop is bound by shared-memory instructions
        const float4 d4 = *reinterpret_cast<const float4*>(&s_d[b][j]);
        acc[j] += d4.x * xs[b];
        acc[j + 1] += d4.y * xs[b];
        acc[j + 2] += d4.z * xs[b];
        acc[j + 3] += d4.w * xs[b];
      }
      if (blockIdx.x == 0 && threadIdx.x < PJ) bsum += s_d[b][threadIdx.x];
    }
  }
  if (i < H) {
#pragma unroll
    for (int j = 0; j < PJ; ++j)
      if (j0 + j < H) dWp[static_cast<size_t>(j0 + j) * H + i] += acc[j];  // single writer per element
  }
  if (blockIdx.x == 0 && threadIdx.x < PJ && j0 + threadIdx.x < H) dbp[j0 + threadIdx.x] += bsum;
}

// dy[t][:] = ds[t] * Ws[0][:] + de[t] * Ws[1][:] (+ dx0[b][:] on the [CLS] row)   -> bf16
// dWs[k][:] += sum_t d_k[t] * y[t][:], dbs[k] += sum_t d_k[t]     (persistent grid, register partials)
constexpr int SPAN_WARPS = 8;
__global__ void __launch_bounds__(SPAN_WARPS * 32)
span_bwd_kernel(const bf16* __restrict__ y, const float* __restrict__ d_start, const float* __restrict__ d_end,
                const float* __restrict__ Ws, const float* __restrict__ dx0, bf16* __restrict__ dy,
                float* __restrict__ dWs, float* __restrict__ dbs, int T, int S, int H) {
  // One warp per token row, rows round-robin over a persistent grid of one block per SM (the weight-gradient
  // partials end in 2 H atomics per BLOCK: few blocks keep that cheap).  The next row's activations are loaded
  // before the current row is processed.
  __shared__ float red[2 * 1024 + 2];  // block partials, combined with shared-memory atomics
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int MV = 4;
  for (int c = threadIdx.x; c < 2 * 1024 + 2; c += blockDim.x) red[c] = 0.0f;
  __syncthreads();
  float acc0[MV][8], acc1[MV][8];
#pragma unroll
  for (int j = 0; j < MV; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc0[j][i] = acc1[j][i] = 0.0f;
  float sb0 = 0.0f, sb1 = 0.0f;
  const int stride = gridDim.x * SPAN_WARPS;
  int t = blockIdx.x * SPAN_WARPS + warp;
  uint4 nxt[MV];
#pragma unroll
  for (int j = 0; j < MV; ++j) {
    const int c = (j * 32 + lane) * 8;
    nxt[j] = (t < T && c < H) ? __ldg(reinterpret_cast<const uint4*>(y + static_cast<size_t>(t) * H + c)) : make_uint4(0, 0, 0, 0);
  }
  for (; t < T; t += stride) {
    uint4 cur[MV];
#pragma unroll
    for (int j = 0; j < MV; ++j) cur[j] = nxt[j];
    const int tn = t + stride;
    if (tn < T) {
#pragma unroll
      for (int j = 0; j < MV; ++j) {
        const int c = (j * 32 + lane) * 8;
        if (c < H) nxt[j] = __ldg(reinterpret_cast<const uint4*>(y + static_cast<size_t>(tn) * H + c));
      }
    }
    const float ds = d_start[t], de = d_end[t];
    sb0 += ds;
    sb1 += de;
    const bool is_cls = (t % S) == 0;
    const int b = t / S;
#pragma unroll
    for (int j = 0; j < MV; ++j) {
      const int c = (j * 32 + lane) * 8;
      if (c < H) {
        const uint4 u = cur[j];
        float yv[8];
        const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
        yv[0] = a.x; yv[1] = a.y; yv[2] = bb.x; yv[3] = bb.y; yv[4] = cc.x; yv[5] = cc.y; yv[6] = d.x; yv[7] = d.y;
        const float4 w00 = __ldg(reinterpret_cast<const float4*>(Ws + c)), w01 = __ldg(reinterpret_cast<const float4*>(Ws + c + 4));
        const float4 w10 = __ldg(reinterpret_cast<const float4*>(Ws + H + c)), w11 = __ldg(reinterpret_cast<const float4*>(Ws + H + c + 4));
        const float ws0[8] = {w00.x, w00.y, w00.z, w00.w, w01.x, w01.y, w01.z, w01.w};
        const float ws1[8] = {w10.x, w10.y, w10.z, w10.w, w11.x, w11.y, w11.z, w11.w};
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          acc0[j][i] += ds * yv[i];
          acc1[j][i] += de * yv[i];
          o[i] = ds * ws0[i] + de * ws1[i];
        }
        if (is_cls) {
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] += dx0[static_cast<size_t>(b) * H + c + i];
        }
        uint4 w;
        w.x = pack_bf16x2(o[0], o[1]);
        w.y = pack_bf16x2(o[2], o[3]);
        w.z = pack_bf16x2(o[4], o[5]);
        w.w = pack_bf16x2(o[6], o[7]);
        *reinterpret_cast<uint4*>(dy + static_cast<size_t>(t) * H + c) = w;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < MV; ++j) {
    if ((j * 32 + lane) * 8 < H) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        atomicAdd(&red[(j * 32 + lane) * 8 + i], acc0[j][i]);
        atomicAdd(&red[1024 + (j * 32 + lane) * 8 + i], acc1[j][i]);
      }
    }
  }
  sb0 = warp_sum(sb0);
  sb1 = warp_sum(sb1);
  if (lane == 0) {
    atomicAdd(&red[2048], sb0);
    atomicAdd(&red[2049], sb1);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    atomicAdd(dWs + c, red[c]);
    atomicAdd(dWs + H + c, red[1024 + c]);
  }
  if (threadIdx.x < 2) atomicAdd(dbs + threadIdx.x, red[2048 + threadIdx.x]);
}

// ---------------------------------------------------------------------------------------
// decode (predictor.py:114-120): torch.max(dim=-1) -> first maximal index on ties
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void argmax_combine(float& v, int& i, float ov, int oi) {
  if (ov > v || (ov == v && oi < i)) {
    v = ov;
    i = oi;
  }
}

__global__ void __launch_bounds__(256)
decode_kernel(const float* __restrict__ start, const float* __restrict__ end, const float* __restrict__ cls,
              int64_t* __restrict__ start_idx, int64_t* __restrict__ end_idx, int64_t* __restrict__ cls_idx,
              float* __restrict__ start_max, float* __restrict__ end_max, float* __restrict__ cls_prob,
              float* __restrict__ score, int S, int C) {
  __shared__ float s_v[8];
  __shared__ int s_i[8];
  __shared__ float s_best[2];
  const int b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int which = 0; which < 2; ++which) {
    const float* x = (which == 0 ? start : end) + static_cast<size_t>(b) * S;
    float v = -INFINITY;
    int idx = 0x7fffffff;
    for (int s = threadIdx.x; s < S; s += blockDim.x) argmax_combine(v, idx, x[s], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      argmax_combine(v, idx, ov, oi);
    }
    __syncthreads();
    if (lane == 0) {
      s_v[warp] = v;
      s_i[warp] = idx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < 8; ++w) argmax_combine(v, idx, s_v[w], s_i[w]);
      if (idx == 0x7fffffff) idx = 0;
      (which == 0 ? start_idx : end_idx)[b] = idx;
      (which == 0 ? start_max : end_max)[b] = v;
      s_best[which] = v;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // same association as predictor.py:120: (max_s + max_e) - (s[0] + e[0])
    score[b] = (s_best[0] + s_best[1]) - (start[static_cast<size_t>(b) * S] + end[static_cast<size_t>(b) * S]);
    const float* z = cls + static_cast<size_t>(b) * C;
    float mx = z[0];
    int mi = 0;
    for (int c = 1; c < C; ++c)
      if (z[c] > mx) {
        mx = z[c];
        mi = c;
      }
    float se = 0.0f;
    for (int c = 0; c < C; ++c) se += expf(z[c] - mx);
    cls_idx[b] = mi;
    cls_prob[b] = 1.0f / se;  // softmax value at the arg max
  }
}

}  // namespace

int launch_span_head_fwd(const bf16* y, const float* span_w, const float* span_b, float* start, float* end, int T,
                         int H, cudaStream_t st) {
  span_head_fwd_kernel<<<(T + 3) / 4, 128, 0, st>>>(y, span_w, span_b, start, end, T, H);
  return check_launch("span_head_fwd");
}

int launch_pooler_heads_fwd(const bf16* y, const HeadParams& hp, float* pooled, float* cls, float* reg_s,
                            float* reg_e, float* reg_s2, float* reg_e2, int B, int S, int H, int C, DropKey cls_drop,
                            cudaStream_t st) {
  dim3 grid((H + 31) / 32, B);
  pooler_fwd_kernel<<<grid, 256, 0, st>>>(y, hp.pool_w, hp.pool_b, pooled, S, H);
  if (check_launch("pooler_fwd")) return 1;
  heads_fwd_kernel<<<B, 256, 0, st>>>(pooled, hp, cls, reg_s, reg_e, reg_s2, reg_e2, H, C, cls_drop);
  return check_launch("heads_fwd");
}

int launch_loss_fwd_bwd(const float* start, const float* end, const float* cls, const float* reg_s,
                        const float* reg_e, const int64_t* y_start, const int64_t* y_end, const int64_t* y_cls,
                        const float* y_reg_s, const float* y_reg_e, float* d_start, float* d_end, float* d_cls,
                        float* d_reg_s, float* d_reg_e, float* row_scratch, float* losses_out, int B, int S,
                        int C, const LossConfig& lc, int with_grad, cudaStream_t st, const int* err_flag) {
  loss_rows_kernel<<<dim3(B, 3), 256, 0, st>>>(start, end, cls, reg_s, reg_e, y_start, y_end, y_cls, y_reg_s, y_reg_e,
                                      d_start, d_end, d_cls, d_reg_s, d_reg_e, row_scratch, B, S, C, lc,
                                      with_grad);
  if (check_launch("loss_rows")) return 1;
  loss_finalize_kernel<<<1, 32, 0, st>>>(row_scratch, losses_out, B, lc, err_flag);
  return check_launch("loss_finalize");
}

int launch_heads_bwd(const bf16* y, const float* pooled, const float* reg_s, const float* reg_e,
                     const float* d_start, const float* d_end, const float* d_cls, const float* d_reg_s,
                     const float* d_reg_e, const HeadParams& hp, const HeadGrads& hg, float* dpre_scratch,
                     float* dx0_scratch, bf16* dy, int B, int S, int H, int C, DropKey cls_drop, int num_sms,
                     cudaStream_t st) {
  if (H > 1024 || H % 8 != 0) {
    set_error("heads backward: H=%d must be a multiple of 8 and <= 1024", H);
    return 1;
  }
  pooled_bwd_kernel<<<B, 256, 0, st>>>(pooled, reg_s, reg_e, d_cls, d_reg_s, d_reg_e, hp, hg, dpre_scratch, H, C,
                                       cls_drop);
  if (check_launch("pooled_bwd")) return 1;
  dim3 g2((H + 255) / 256, (H + PJ - 1) / PJ);
  BQA_CUDA_CHECK(cudaMemsetAsync(dx0_scratch, 0, sizeof(float) * B * H, st));
  pooler_dx_kernel<<<g2, 256, 0, st>>>(dpre_scratch, hp.pool_w, dx0_scratch, B, H);
  if (check_launch("pooler_dx")) return 1;
  pooler_wgrad_kernel<<<g2, 256, 0, st>>>(dpre_scratch, y, hg.pool_w, hg.pool_b, B, S, H);
  if (check_launch("pooler_wgrad")) return 1;
  const int T = B * S;
  int grid = num_sms;
  if (grid > (T + SPAN_WARPS - 1) / SPAN_WARPS) grid = (T + SPAN_WARPS - 1) / SPAN_WARPS;
  span_bwd_kernel<<<grid, SPAN_WARPS * 32, 0, st>>>(y, d_start, d_end, hp.span_w, dx0_scratch, dy, hg.span_w, hg.span_b, T, S,
                                        H);
  return check_launch("span_bwd");
}

int launch_decode(const float* start, const float* end, const float* cls, int64_t* start_idx, int64_t* end_idx,
                  int64_t* cls_idx, float* start_max, float* end_max, float* cls_prob, float* score, int B, int S,
                  int C, cudaStream_t st) {
  decode_kernel<<<B, 256, 0, st>>>(start, end, cls, start_idx, end_idx, cls_idx, start_max, end_max, cls_prob,
                                   score, S, C);
  return check_launch("decode");
}

}  // namespace bqa
