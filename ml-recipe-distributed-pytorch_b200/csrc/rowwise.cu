// rowwise.cu — HBM-bound row kernels of the encoder: embeddings + LayerNorm, LayerNorm
// forward/backward, bias-gradient column sums and the embedding-gradient scatter.
//
// Reference call sites (all executed by ATen in the reference):
//   BertEmbeddings.forward      site-packages/transformers/models/bert/modeling_bert.py:72-112
//   BertSelfOutput / BertOutput LayerNorm(dropout(dense(h)) + x)          ...:294-298,352-356
//   their autograd backward      reference modules/model/trainer/trainer.py:197-204
//
// Layout: activations are [T = B*S tokens][H] bf16 row-major.  One warp owns one row and
// keeps it in registers (H <= 1024: at most four 16-byte vectors per lane); statistics are
// fp32 with the biased two-pass variance torch uses.  Every pass touches each byte once.
#include "kernels.cuh"

namespace bqa {

namespace {

constexpr int MAXV = 4;  // 16-byte vectors per lane -> H <= 1024
constexpr int ROW_WARPS = 4;

struct RowVec {
  float v[MAXV][8];
};

__device__ __forceinline__ void load_row_bf16(const bf16* __restrict__ row, int H, int lane, RowVec& r) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(row + c));
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      r.v[j][0] = a.x; r.v[j][1] = a.y; r.v[j][2] = b.x; r.v[j][3] = b.y;
      r.v[j][4] = cc.x; r.v[j][5] = cc.y; r.v[j][6] = d.x; r.v[j][7] = d.y;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) r.v[j][i] = 0.0f;
    }
  }
}

__device__ __forceinline__ void store_row_bf16(bf16* __restrict__ row, int H, int lane, const RowVec& r) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
      uint4 u;
      u.x = pack_bf16x2(r.v[j][0], r.v[j][1]);
      u.y = pack_bf16x2(r.v[j][2], r.v[j][3]);
      u.z = pack_bf16x2(r.v[j][4], r.v[j][5]);
      u.w = pack_bf16x2(r.v[j][6], r.v[j][7]);
      *reinterpret_cast<uint4*>(row + c) = u;
    }
  }
}

__device__ __forceinline__ void load_vec_f32(const float* __restrict__ p, int H, int lane, RowVec& r) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(p + c));
      const float4 b = __ldg(reinterpret_cast<const float4*>(p + c + 4));
      r.v[j][0] = a.x; r.v[j][1] = a.y; r.v[j][2] = a.z; r.v[j][3] = a.w;
      r.v[j][4] = b.x; r.v[j][5] = b.y; r.v[j][6] = b.z; r.v[j][7] = b.w;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) r.v[j][i] = 0.0f;
    }
  }
}

struct RawRow {
  uint4 u[MAXV];
};

__device__ __forceinline__ void load_raw(const bf16* __restrict__ row, int H, int lane, RawRow& r) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    r.u[j] = c < H ? __ldg(reinterpret_cast<const uint4*>(row + c)) : make_uint4(0u, 0u, 0u, 0u);
  }
}

__device__ __forceinline__ void unpack_raw(const RawRow& raw, RowVec& r) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const float2 a = unpack_bf16x2(raw.u[j].x), b = unpack_bf16x2(raw.u[j].y);
    const float2 c = unpack_bf16x2(raw.u[j].z), d = unpack_bf16x2(raw.u[j].w);
    r.v[j][0] = a.x; r.v[j][1] = a.y; r.v[j][2] = b.x; r.v[j][3] = b.y;
    r.v[j][4] = c.x; r.v[j][5] = c.y; r.v[j][6] = d.x; r.v[j][7] = d.y;
  }
}

// round a register row to bf16 precision in place (what a later kernel will read back)
__device__ __forceinline__ void round_row_bf16(RowVec& r) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[j][i] = __bfloat162float(__float2bfloat16_rn(r.v[j][i]));
}

__device__ __forceinline__ void row_stats(const RowVec& r, int H, int lane, float eps, float& mean, float& rstd) {
  float s = 0.0f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) s += r.v[j][i];
  mean = warp_sum(s) / static_cast<float>(H);
  float q = 0.0f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = r.v[j][i] - mean;
        q += d * d;
      }
    }
  }
  rstd = rsqrtf(warp_sum(q) / static_cast<float>(H) + eps);
}

// y = LayerNorm(z) * gamma + beta, optional dropout on y (embeddings only)
__device__ __forceinline__ void ln_apply(RowVec& r, const RowVec& g, const RowVec& b, int H, int lane, float mean,
                                         float rstd, const DropKey& drop, uint32_t row_index) {
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float y = (r.v[j][i] - mean) * rstd * g.v[j][i] + b.v[j][i];
        if (drop.thresh != 0u) y = drop_keep(drop, row_index * H + c + i) ? y * drop.scale : 0.0f;
        r.v[j][i] = y;
      }
    }
  }
}

// NV = 16-byte vectors per lane (H <= 256 NV).  Persistent grid (every block resident, 8 per SM): rows are dealt
// round-robin to the warps, so every SM streams the same number of rows and there is no partial last wave; the
// loads of a warp's next row are issued before the current row is reduced.  gamma / beta are read at use
// (L1-resident) so that only the rows themselves live in registers.
constexpr int LN_FWD_BLOCKS_PER_SM = 8;
template <int NV>
__global__ void __launch_bounds__(ROW_WARPS * 32, LN_FWD_BLOCKS_PER_SM)
ln_fwd_kernel(const bf16* __restrict__ z, const float* __restrict__ gamma, const float* __restrict__ beta,
              bf16* __restrict__ y, int T, int H, float eps) {
  pdl_wait();     // launched with programmatic serialization: the producer of z has completed from here on
  pdl_trigger();  // the next kernel (a GEMM) may start its prologue now; it waits for this grid before reading
  const int lane = threadIdx.x & 31;
  const int row_stride = gridDim.x * ROW_WARPS;
  int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  const float invH = 1.0f / static_cast<float>(H);
  uint4 nraw[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = (j * 32 + lane) * 8;
    nraw[j] = c < H ? __ldg(reinterpret_cast<const uint4*>(z + static_cast<size_t>(row) * H + c)) : make_uint4(0, 0, 0, 0);
  }
  for (; row < T; row += row_stride) {
    float x[NV][8];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const uint4 u = nraw[j];
      const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z), a3 = unpack_bf16x2(u.w);
      x[j][0] = a0.x; x[j][1] = a0.y; x[j][2] = a1.x; x[j][3] = a1.y;
      x[j][4] = a2.x; x[j][5] = a2.y; x[j][6] = a3.x; x[j][7] = a3.y;
    }
    const int nrow = row + row_stride;
    if (nrow < T) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int c = (j * 32 + lane) * 8;
        if (c < H) nraw[j] = __ldg(reinterpret_cast<const uint4*>(z + static_cast<size_t>(nrow) * H + c));
      }
    }
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < NV; ++j)
#pragma unroll
      for (int i = 0; i < 8; ++i) s += x[j][i];
    const float mean = warp_sum(s) * invH;
    float qv = 0.0f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      if ((j * 32 + lane) * 8 < H) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float t = x[j][i] - mean;
          qv = fmaf(t, t, qv);
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(qv) * invH + eps);
    bf16* yr = y + static_cast<size_t>(row) * H;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (j * 32 + lane) * 8;
      if (c < H) {
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c + 4));
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c + 4));
        uint4 o;
        o.x = pack_bf16x2((x[j][0] - mean) * rstd * g0.x + b0.x, (x[j][1] - mean) * rstd * g0.y + b0.y);
        o.y = pack_bf16x2((x[j][2] - mean) * rstd * g0.z + b0.z, (x[j][3] - mean) * rstd * g0.w + b0.w);
        o.z = pack_bf16x2((x[j][4] - mean) * rstd * g1.x + b1.x, (x[j][5] - mean) * rstd * g1.y + b1.y);
        o.w = pack_bf16x2((x[j][6] - mean) * rstd * g1.z + b1.z, (x[j][7] - mean) * rstd * g1.w + b1.w);
        *reinterpret_cast<uint4*>(yr + c) = o;
      }
    }
  }
}

// z0 = word[ids] + pos[s] + type[tt]  (fp32 tables, fp32 sum, stored as bf16)
// x0 = dropout(LayerNorm(z0))
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_ln_fwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ token_type,
                    const uint8_t* __restrict__ key_mask, const float* __restrict__ word,
                    const float* __restrict__ pos, const float* __restrict__ type, const float* __restrict__ gamma,
                    const float* __restrict__ beta, bf16* __restrict__ z0, bf16* __restrict__ x0,
                    int64_t* __restrict__ ids_kept, int64_t* __restrict__ tt_kept, uint8_t* __restrict__ mask_kept,
                    int32_t* __restrict__ pos_kept, int T, int S, int H, int V, int TV, int P, int pos_mode, int pad_id,
                    float eps, DropKey drop, int* __restrict__ err_flag) {
  pdl_trigger();  // the next kernel (a GEMM) may start its prologue now; it waits for this grid before reading
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  long long id = ids[row];
  long long tt = token_type != nullptr ? token_type[row] : 0;
  if (id < 0 || id >= V || tt < 0 || tt >= TV) {
    // torch raises IndexError here; the flag turns the step's loss into NaN and bqa_input_status() reports it
    if (lane == 0) atomicExch(err_flag, 1);
    id = 0;
    tt = 0;
  }
  // the backward (embedding scatter, key mask) reads these private, range-checked copies of the inputs
  if (lane == 0) {
    if (ids_kept != nullptr) ids_kept[row] = id;
    if (tt_kept != nullptr) tt_kept[row] = tt;
    if (mask_kept != nullptr && key_mask != nullptr) mask_kept[row] = key_mask[row];
  }
  int s = row % S;
  if (pos_mode != 0) {
    // RoBERTa: pad_id + (non-padding tokens up to and including this one), pad_id for a padding token
    const int64_t* seq = ids + static_cast<size_t>(row - s);
    int cnt = 0;
    for (int j = lane; j <= s; j += 32) cnt += seq[j] != pad_id ? 1 : 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    s = ids[row] != pad_id ? cnt + pad_id : pad_id;
  }
  if (s >= P) {
    if (lane == 0) atomicExch(err_flag, 1);  // beyond the position table: IndexError in the reference
    s = 0;
  }
  if (lane == 0 && pos_kept != nullptr) pos_kept[row] = s;
  RowVec w, p, t, g, b;
  load_vec_f32(word + static_cast<size_t>(id) * H, H, lane, w);
  load_vec_f32(pos + static_cast<size_t>(s) * H, H, lane, p);
  load_vec_f32(type + static_cast<size_t>(tt) * H, H, lane, t);
#pragma unroll
  for (int j = 0; j < MAXV; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) w.v[j][i] = (w.v[j][i] + t.v[j][i]) + p.v[j][i];
  round_row_bf16(w);
  store_row_bf16(z0 + static_cast<size_t>(row) * H, H, lane, w);
  load_vec_f32(gamma, H, lane, g);
  load_vec_f32(beta, H, lane, b);
  float mean, rstd;
  row_stats(w, H, lane, eps, mean, rstd);
  ln_apply(w, g, b, H, lane, mean, rstd, drop, row);
  store_row_bf16(x0 + static_cast<size_t>(row) * H, H, lane, w);
}

// LayerNorm backward.  dz = rstd * (dxhat - mean(dxhat) - xhat * mean(dxhat * xhat)),
// dgamma += sum_rows dy * xhat, dbeta += sum_rows dy.
//   drop_in : dy is first multiplied by the keep mask of a dropout applied AFTER this LN (embeddings)
//   drop_out: additionally emits dz * keep-mask (gradient of the GEMM output that was dropped
//             BEFORE the residual add) and accumulates its column sums into dbias
// Rows are distributed round-robin over a persistent grid, one warp per row, NV = 16-byte vectors per lane
// (H <= 256 NV).  The kernel is bound by instruction issue, not by HBM, unless the per-element work is small:
// a lane always meets the same columns, so the three column partials (dgamma, dbeta, dbias) live in registers
// for the whole kernel (no shared-memory read-modify-write per row), xhat and the output come from fused
// multiply-adds with per-row constants, and the next row's loads are issued before the current row is reduced.
// Partials are combined once per block (shared memory) and once per grid (atomicAdd).
// FAST: the encoder layers' case fixed at compile time (H == NV * 256, no dropout after this LayerNorm, masked copy
// with dropout and its column sums wanted, dz_masked != dz): the uniform branches around every 8-element group cost 8 %
// of the kernel's instructions.
template <int NV, bool FAST>
__global__ void __launch_bounds__(ROW_WARPS * 32, NV <= 3 ? 3 : 2)
ln_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z, const float* __restrict__ gamma,
              bf16* __restrict__ dz, bf16* __restrict__ dz_masked, float* __restrict__ dgamma,
              float* __restrict__ dbeta, float* __restrict__ dbias, int T, int H, float eps, DropKey drop_in,
              DropKey drop_out) {
  pdl_wait();     // launched with programmatic serialization: the producers of dy / z have completed from here on
  pdl_trigger();  // the next kernel (a GEMM) may start its prologue now; it waits for this grid before reading
  extern __shared__ float ln_smem[];
  constexpr int W = NV * 256;           // padded row width
  float* s_gamma = ln_smem;             // [W]
  float* s_acc = ln_smem + W;           // [ROW_WARPS][W], reused for the three reductions at the end
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  for (int c = threadIdx.x; c < W; c += blockDim.x) s_gamma[c] = c < H ? gamma[c] : 0.0f;
  __syncthreads();
  const float invH = 1.0f / static_cast<float>(H);
  const bool masking = FAST || dz_masked != nullptr;
  const bool want_dbias = FAST || dbias != nullptr;

  float ag[NV][8], ab[NV][8], am[NV][8];
#pragma unroll
  for (int j = 0; j < NV; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) ag[j][i] = ab[j][i] = am[j][i] = 0.0f;

  const int row_stride = gridDim.x * ROW_WARPS;
  int row = blockIdx.x * ROW_WARPS + warp;
  uint4 nz[NV], nd[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int c = (j * 32 + lane) * 8;
    const bool ok = row < T && (FAST || c < H);
    nz[j] = ok ? __ldg(reinterpret_cast<const uint4*>(z + static_cast<size_t>(row) * H + c)) : make_uint4(0, 0, 0, 0);
    nd[j] = ok ? __ldg(reinterpret_cast<const uint4*>(dy + static_cast<size_t>(row) * H + c)) : make_uint4(0, 0, 0, 0);
  }
  for (; row < T; row += row_stride) {
    float x[NV][8], d[NV][8];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float2 a0 = unpack_bf16x2(nz[j].x), a1 = unpack_bf16x2(nz[j].y), a2 = unpack_bf16x2(nz[j].z),
                   a3 = unpack_bf16x2(nz[j].w);
      x[j][0] = a0.x; x[j][1] = a0.y; x[j][2] = a1.x; x[j][3] = a1.y;
      x[j][4] = a2.x; x[j][5] = a2.y; x[j][6] = a3.x; x[j][7] = a3.y;
      const float2 b0 = unpack_bf16x2(nd[j].x), b1 = unpack_bf16x2(nd[j].y), b2 = unpack_bf16x2(nd[j].z),
                   b3 = unpack_bf16x2(nd[j].w);
      d[j][0] = b0.x; d[j][1] = b0.y; d[j][2] = b1.x; d[j][3] = b1.y;
      d[j][4] = b2.x; d[j][5] = b2.y; d[j][6] = b3.x; d[j][7] = b3.y;
    }
    const int nrow = row + row_stride;
    if (nrow < T) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int c = (j * 32 + lane) * 8;
        if (FAST || c < H) {
          nz[j] = __ldg(reinterpret_cast<const uint4*>(z + static_cast<size_t>(nrow) * H + c));
          nd[j] = __ldg(reinterpret_cast<const uint4*>(dy + static_cast<size_t>(nrow) * H + c));
        }
      }
    }
    // The arithmetic below runs on packed fp32 pairs (FFMA2 / FMUL2 / FADD2: one issue slot for two elements; the
    // kernel is bound by instruction issue).  Each element sees the same operations as in scalar form.
    // row statistics (biased two-pass variance, as torch)
    uint64_t sp = f2_splat(0.0f);
#pragma unroll
    for (int j = 0; j < NV; ++j)
#pragma unroll
      for (int i = 0; i < 8; i += 2) sp = f2_add(sp, f2_pack(x[j][i], x[j][i + 1]));
    float s_lo, s_hi;
    f2_unpack(sp, s_lo, s_hi);
    const float mean = warp_sum(s_lo + s_hi) * invH;
    const uint64_t nmean2 = f2_splat(-mean);
    uint64_t qp = f2_splat(0.0f);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      if (FAST || (j * 32 + lane) * 8 < H) {
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          const uint64_t t2 = f2_add(f2_pack(x[j][i], x[j][i + 1]), nmean2);
          qp = f2_fma(t2, t2, qp);
        }
      }
    }
    float q_lo, q_hi;
    f2_unpack(qp, q_lo, q_hi);
    const float rstd = rsqrtf(warp_sum(q_lo + q_hi) * invH + eps);
    const float nmr = -mean * rstd;  // xhat = x * rstd + nmr
    const uint64_t rstd2 = f2_splat(rstd), nmr2 = f2_splat(nmr);
    uint64_t s1p = f2_splat(0.0f), s2p = f2_splat(0.0f);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (j * 32 + lane) * 8;
      if (FAST || c < H) {
        const float4 g0 = *reinterpret_cast<const float4*>(s_gamma + c), g1 = *reinterpret_cast<const float4*>(s_gamma + c + 4);
        const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        if (!FAST && drop_in.thresh != 0u) {
          const uint32_t base = static_cast<uint32_t>(row) * H + c;  // even
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            bool k0, k1;
            drop_keep2(drop_in, base + i, k0, k1);
            d[j][i] = k0 ? d[j][i] * drop_in.scale : 0.0f;
            d[j][i + 1] = k1 ? d[j][i + 1] * drop_in.scale : 0.0f;
          }
        }
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          const uint64_t dy2 = f2_pack(d[j][i], d[j][i + 1]);
          const uint64_t xhat2 = f2_fma(f2_pack(x[j][i], x[j][i + 1]), rstd2, nmr2);
          const uint64_t dxhat2 = f2_mul(dy2, f2_pack(gv[i], gv[i + 1]));
          f2_unpack(f2_fma(dy2, xhat2, f2_pack(ag[j][i], ag[j][i + 1])), ag[j][i], ag[j][i + 1]);
          f2_unpack(f2_add(f2_pack(ab[j][i], ab[j][i + 1]), dy2), ab[j][i], ab[j][i + 1]);
          s1p = f2_add(s1p, dxhat2);
          s2p = f2_fma(dxhat2, xhat2, s2p);
          f2_unpack(xhat2, x[j][i], x[j][i + 1]);
          f2_unpack(dxhat2, d[j][i], d[j][i + 1]);
        }
      }
    }
    float s1_lo, s1_hi, s2_lo, s2_hi;
    f2_unpack(s1p, s1_lo, s1_hi);
    f2_unpack(s2p, s2_lo, s2_hi);
    const float s1 = warp_sum(s1_lo + s1_hi) * invH;
    const float s2 = warp_sum(s2_lo + s2_hi) * invH;
    // dz = rstd * (dxhat - s1 - xhat * s2) = dxhat * rstd + (xhat * k2 + k1)
    const uint64_t k1p = f2_splat(-s1 * rstd), k2p = f2_splat(-s2 * rstd);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (j * 32 + lane) * 8;
      if (FAST || c < H) {
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
          float v0, v1;
          f2_unpack(f2_fma(f2_pack(d[j][i], d[j][i + 1]), rstd2, f2_fma(f2_pack(x[j][i], x[j][i + 1]), k2p, k1p)), v0, v1);
          pk[i >> 1] = pack_bf16x2(v0, v1);
        }
        *reinterpret_cast<uint4*>(dz + static_cast<size_t>(row) * H + c) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        if (masking) {
          // consumers on the GEMM side read the bf16-rounded value: mask and scale THAT value
          uint32_t pm[4];
          const uint32_t base = static_cast<uint32_t>(row) * H + c;  // even: H and c are multiples of 8
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const float2 rr = unpack_bf16x2(pk[i >> 1]);
            float m0 = rr.x, m1 = rr.y;
            if (FAST || drop_out.thresh != 0u) {
              bool k0, k1b;
              drop_keep2(drop_out, base + i, k0, k1b);
              f2_unpack(f2_mul(f2_pack(m0, m1), f2_pack(k0 ? drop_out.scale : 0.0f, k1b ? drop_out.scale : 0.0f)), m0, m1);
            }
            pm[i >> 1] = pack_bf16x2(m0, m1);
            if (want_dbias) {
              const float2 mr = unpack_bf16x2(pm[i >> 1]);
              f2_unpack(f2_add(f2_pack(am[j][i], am[j][i + 1]), f2_pack(mr.x, mr.y)), am[j][i], am[j][i + 1]);
            }
          }
          if (FAST || dz_masked != dz)
            *reinterpret_cast<uint4*>(dz_masked + static_cast<size_t>(row) * H + c) = make_uint4(pm[0], pm[1], pm[2], pm[3]);
        } else if (want_dbias) {
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const float2 rr = unpack_bf16x2(pk[i >> 1]);
            am[j][i] += rr.x;
            am[j][i + 1] += rr.y;
          }
        }
      }
    }
  }
  // block reduction of the three column partials, one after the other through the same buffer
  auto reduce_out = [&](float (&acc)[NV][8], float* out) {
    if (out == nullptr) return;  // uniform
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int c = (j * 32 + lane) * 8;
      *reinterpret_cast<float4*>(s_acc + warp * W + c) = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
      *reinterpret_cast<float4*>(s_acc + warp * W + c + 4) = make_float4(acc[j][4], acc[j][5], acc[j][6], acc[j][7]);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float t = 0.0f;
#pragma unroll
      for (int w = 0; w < ROW_WARPS; ++w) t += s_acc[w * W + c];
      atomicAdd(out + c, t);
    }
  };
  reduce_out(ag, dgamma);
  reduce_out(ab, dbeta);
  reduce_out(am, dbias);
}

// out[n] += sum_t X[t][n]  (bias gradients of the fused QKV and intermediate projections)
__global__ void __launch_bounds__(256)
colsum_kernel(const bf16* __restrict__ X, int ld, float* __restrict__ out, int T, int N) {
  pdl_trigger();  // the next kernel (a GEMM) may start its prologue now; it waits for this grid before reading
  __shared__ float red[8][32 * 8 + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + tx) * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (c < N) {
    const int step = gridDim.y * 8;
    int r = blockIdx.y * 8 + ty;
    for (; r + 3 * step < T; r += 4 * step) {  // four independent 16-byte loads in flight per thread
      uint4 u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) u[k] = __ldg(reinterpret_cast<const uint4*>(X + static_cast<size_t>(r + k * step) * ld + c));
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 a = unpack_bf16x2(u[k].x), b = unpack_bf16x2(u[k].y), cc = unpack_bf16x2(u[k].z), d = unpack_bf16x2(u[k].w);
        acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y;
        acc[4] += cc.x; acc[5] += cc.y; acc[6] += d.x; acc[7] += d.y;
      }
    }
    for (; r < T; r += step) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(X + static_cast<size_t>(r) * ld + c));
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y;
      acc[4] += cc.x; acc[5] += cc.y; acc[6] += d.x; acc[7] += d.y;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[ty][tx * 8 + i] = acc[i];
  __syncthreads();
  const int col = threadIdx.x;  // 256 columns per block
  float s = 0.0f;
#pragma unroll
  for (int w = 0; w < 8; ++w) s += red[w][col];
  const int gc = blockIdx.x * 256 + col;
  if (gc < N) atomicAdd(out + gc, s);
}

// Embedding gradient: dword[ids[t]] += dz0[t], dpos[t % S] += dz0[t], dtype[tt[t]] += dz0[t].
// The word table is nn.Embedding(..., padding_idx=pad_token_id) in BertEmbeddings
// (modeling_bert.py:59): the padding row receives no gradient.
//   word rows : one warp per token, fp32 atomics (rows are mostly distinct; [CLS]/[SEP] collide B-fold)
//   pos/type  : one block per position s sums its B rows in registers: dpos[s] has a single writer
//               (no atomics), the two type rows get one atomic per block and column instead of one
//               per token and column (T-way contention on two rows was 1.1 ms per step).
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_word_bwd_kernel(const bf16* __restrict__ dz0, const int64_t* __restrict__ ids, float* __restrict__ dword,
                      int T, int H, int V, int pad_id) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  const long long id = ids[row];
  if (id == pad_id || id < 0 || id >= V) return;  // out-of-range ids were reported by the forward
  RowVec d;
  load_row_bf16(dz0 + static_cast<size_t>(row) * H, H, lane, d);
  float* w = dword + static_cast<size_t>(id) * H;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
      // two 16-byte vector reductions instead of eight scalar ones
      red_add_f32x4(w + c, d.v[j][0], d.v[j][1], d.v[j][2], d.v[j][3]);
      red_add_f32x4(w + c + 4, d.v[j][4], d.v[j][5], d.v[j][6], d.v[j][7]);
    }
  }
}

// dpos[pos_ids[t]] += dz0[t] for position ids that are not the index in the sequence (RoBERTa); the padding
// index of the position table receives no gradient.  One warp per token, 16-byte vector reductions.
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_pos_scatter_bwd_kernel(const bf16* __restrict__ dz0, const int32_t* __restrict__ pos_ids, float* __restrict__ dpos,
                             int T, int H, int P, int pad_id) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  const int pid = pos_ids[row];
  if (pid == pad_id || pid < 0 || pid >= P) return;
  RowVec d;
  load_row_bf16(dz0 + static_cast<size_t>(row) * H, H, lane, d);
  float* w = dpos + static_cast<size_t>(pid) * H;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
      red_add_f32x4(w + c, d.v[j][0], d.v[j][1], d.v[j][2], d.v[j][3]);
      red_add_f32x4(w + c + 4, d.v[j][4], d.v[j][5], d.v[j][6], d.v[j][7]);
    }
  }
}

// block = one position s: 128 column-vector threads x PT_GROUPS batch groups, combined through shared memory
constexpr int PT_GROUPS = 4;
__global__ void __launch_bounds__(128 * PT_GROUPS)
embed_pos_type_bwd_kernel(const bf16* __restrict__ dz0, const int64_t* __restrict__ token_type,
                          float* __restrict__ dpos, float* __restrict__ dtype, int B, int S, int H, int TV) {
  __shared__ float s_acc[PT_GROUPS - 1][2][1024];
  __shared__ int s_any[PT_GROUPS - 1][2];
  const int s = blockIdx.x;
  const int grp = threadIdx.x >> 7;
  const int c = (threadIdx.x & 127) * 8;  // H <= 1024 -> one 16-byte vector per thread
  float acc[2][8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[0][i] = acc[1][i] = 0.0f;
  bool any[2] = {false, false};
  if (c < H) {
    for (int b = grp; b < B; b += PT_GROUPS) {
      const size_t row = static_cast<size_t>(b) * S + s;
      // two accumulators: type_vocab <= 2 (engine.cu: compute_layout); with one type row everything goes to row 0
      const int tt = (token_type != nullptr && TV > 1 && token_type[row] == 1) ? 1 : 0;
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(dz0 + row * H + c));
      const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      const float v[8] = {a.x, a.y, bb.x, bb.y, cc.x, cc.y, d.x, d.y};
      any[tt] = true;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (tt) acc[1][i] += v[i]; else acc[0][i] += v[i];
      }
    }
  }
  if (grp > 0 && c < H) {
#pragma unroll
    for (int t = 0; t < 2; ++t) {
#pragma unroll
      for (int i = 0; i < 8; ++i) s_acc[grp - 1][t][c + i] = acc[t][i];
      if (c == 0) s_any[grp - 1][t] = any[t];  // the token type of a row is the same for every column
    }
  }
  __syncthreads();
  if (grp != 0 || c >= H) return;
  for (int g = 0; g < PT_GROUPS - 1; ++g) {
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      any[t] = any[t] || s_any[g][t] != 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[t][i] += s_acc[g][t][c + i];
    }
  }
  if (dpos != nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dpos[static_cast<size_t>(s) * H + c + i] += acc[0][i] + acc[1][i];
  }
#pragma unroll
  for (int t = 0; t < 2; ++t)
    if (any[t]) {
      red_add_f32x4(dtype + static_cast<size_t>(t) * H + c, acc[t][0], acc[t][1], acc[t][2], acc[t][3]);
      red_add_f32x4(dtype + static_cast<size_t>(t) * H + c + 4, acc[t][4], acc[t][5], acc[t][6], acc[t][7]);
    }
}

// Model inputs from compact token ids (collate_fun, split_dataset.py:480-499, done on the device):
//   input_ids = ids (widened), attention_mask = ids > 0,
//   token_type_ids = 0 up to and including the first [SEP] of the row, 1 after it (and on the padding);
//   all zeros when type_coef == 0 (RoBERTa).  A row without [SEP] gets token type 0 everywhere and sets
//   *missing_sep (the host-side collate raises for it like the reference's list.index).
__global__ void __launch_bounds__(128)
prepare_inputs_kernel(const int32_t* __restrict__ ids32, int64_t* __restrict__ ids, int64_t* __restrict__ token_type,
                      uint8_t* __restrict__ mask, int S, int sep_id, int type_coef, int* __restrict__ missing_sep) {
  __shared__ int s_first[4];
  const int b = blockIdx.x;
  const int32_t* row = ids32 + static_cast<size_t>(b) * S;
  int first = S;
  for (int j = threadIdx.x; j < S; j += blockDim.x)
    if (row[j] == sep_id) {
      first = j;
      break;  // this thread's positions are increasing
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
  if ((threadIdx.x & 31) == 0) s_first[threadIdx.x >> 5] = first;
  __syncthreads();
  first = min(min(s_first[0], s_first[1]), min(s_first[2], s_first[3]));
  if (first == S && type_coef != 0 && threadIdx.x == 0 && missing_sep != nullptr) *missing_sep = 1;
  for (int j = threadIdx.x; j < S; j += blockDim.x) {
    const int32_t id = row[j];
    const size_t o = static_cast<size_t>(b) * S + j;
    ids[o] = id;
    mask[o] = id > 0 ? 1 : 0;
    token_type[o] = (type_coef != 0 && j > first) ? 1 : 0;
  }
}

}  // namespace

static inline int row_blocks(int T) { return (T + ROW_WARPS - 1) / ROW_WARPS; }

int launch_ln_fwd(const bf16* z, const float* gamma, const float* beta, bf16* y, int T, int H, float eps,
                  cudaStream_t st) {
  if (H % 8 != 0 || H > MAXV * 256) {
    set_error("layernorm: H=%d must be a multiple of 8 and <= %d", H, MAXV * 256);
    return 1;
  }
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  int grid = sms * LN_FWD_BLOCKS_PER_SM;
  if (grid > row_blocks(T)) grid = row_blocks(T);
  switch ((H + 255) / 256) {
    case 1: launch_pdl(ln_fwd_kernel<1>, grid, ROW_WARPS * 32, 0, st, z, gamma, beta, y, T, H, eps); break;
    case 2: launch_pdl(ln_fwd_kernel<2>, grid, ROW_WARPS * 32, 0, st, z, gamma, beta, y, T, H, eps); break;
    case 3: launch_pdl(ln_fwd_kernel<3>, grid, ROW_WARPS * 32, 0, st, z, gamma, beta, y, T, H, eps); break;
    default: launch_pdl(ln_fwd_kernel<4>, grid, ROW_WARPS * 32, 0, st, z, gamma, beta, y, T, H, eps); break;
  }
  return check_launch("ln_fwd");
}

int launch_embed_ln_fwd(const int64_t* ids, const int64_t* token_type, const uint8_t* key_mask, const float* word,
                        const float* pos, const float* type, const float* gamma, const float* beta, bf16* z0,
                        bf16* x0, int64_t* ids_kept, int64_t* tt_kept, uint8_t* mask_kept, int32_t* pos_kept, int T,
                        int S, int H, int V, int TV, int P, int pos_mode, int pad_id, float eps, DropKey drop,
                        int* err_flag, cudaStream_t st) {
  if (H % 8 != 0 || H > MAXV * 256) {
    set_error("embeddings: H=%d must be a multiple of 8 and <= %d", H, MAXV * 256);
    return 1;
  }
  embed_ln_fwd_kernel<<<row_blocks(T), ROW_WARPS * 32, 0, st>>>(ids, token_type, key_mask, word, pos, type, gamma,
                                                              beta, z0, x0, ids_kept, tt_kept, mask_kept, pos_kept, T,
                                                              S, H, V, TV, P, pos_mode, pad_id, eps, drop, err_flag);
  return check_launch("embed_ln_fwd");
}

int launch_ln_bwd(const bf16* dy, const bf16* z, const float* gamma, bf16* dz, bf16* dz_masked, float* dgamma,
                  float* dbeta, float* dbias, int T, int H, float eps, DropKey drop_in, DropKey drop_out,
                  int num_sms, cudaStream_t st) {
  if (H % 8 != 0 || H > MAXV * 256) {
    set_error("layernorm backward: H=%d must be a multiple of 8 and <= %d", H, MAXV * 256);
    return 1;
  }
  const int nv = (H + 255) / 256;
  const size_t smem = static_cast<size_t>(nv) * 256 * (1 + ROW_WARPS) * sizeof(float);
  int grid = num_sms * (nv <= 3 ? 3 : 2);  // every block resident: see __launch_bounds__ of the kernel
  const bool fast = H == nv * 256 && drop_in.thresh == 0u && drop_out.thresh != 0u && dz_masked != nullptr &&
                    dz_masked != dz && dbias != nullptr;
  if (grid > row_blocks(T)) grid = row_blocks(T);
#define BQA_LN_BWD(NV)                                                                                         \
  {                                                                                                            \
    static PerDeviceOnce once;                                                                                 \
    int once_dev; if (once.needs_setup(&once_dev)) {                                                                                  \
      BQA_CUDA_CHECK(cudaFuncSetAttribute(ln_bwd_kernel<NV, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          static_cast<int>(smem)));                                            \
      BQA_CUDA_CHECK(cudaFuncSetAttribute(ln_bwd_kernel<NV, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          static_cast<int>(smem)));                                            \
      once.mark_done(once_dev);                                                                                        \
    }                                                                                                          \
    if (fast)                                                                                                  \
      launch_pdl(ln_bwd_kernel<NV, true>, grid, ROW_WARPS * 32, smem, st, dy, z, gamma, dz, dz_masked, dgamma, dbeta, \
                 dbias, T, H, eps, drop_in, drop_out);                                                         \
    else                                                                                                       \
      launch_pdl(ln_bwd_kernel<NV, false>, grid, ROW_WARPS * 32, smem, st, dy, z, gamma, dz, dz_masked, dgamma, dbeta, \
                 dbias, T, H, eps, drop_in, drop_out);                                                         \
  }
  switch (nv) {
    case 1: BQA_LN_BWD(1) break;
    case 2: BQA_LN_BWD(2) break;
    case 3: BQA_LN_BWD(3) break;
    default: BQA_LN_BWD(4) break;
  }
#undef BQA_LN_BWD
  return check_launch("ln_bwd");
}

int launch_colsum(const bf16* X, int ld, float* out, int T, int N, int num_sms, cudaStream_t st) {
  if (N % 8 != 0 || ld % 8 != 0) {
    set_error("colsum: N=%d and ld=%d must be multiples of 8", N, ld);
    return 1;
  }
  dim3 grid((N + 255) / 256, 1);
  int gy = (num_sms * 4 + grid.x - 1) / grid.x;
  const int max_gy = (T + 7) / 8;
  grid.y = gy < max_gy ? gy : max_gy;
  if (grid.y < 1) grid.y = 1;
  colsum_kernel<<<grid, 256, 0, st>>>(X, ld, out, T, N);
  return check_launch("colsum");
}

int launch_embed_bwd(const bf16* dz0, const int64_t* ids, const int64_t* token_type, const int32_t* pos_ids,
                     float* dword, float* dpos, float* dtype, int T, int S, int H, int V, int TV, int P, int pos_mode,
                     int pad_id, cudaStream_t st) {
  if (H % 8 != 0 || H > 1024) {
    set_error("embedding backward: H=%d must be a multiple of 8 and <= 1024", H);
    return 1;
  }
  embed_word_bwd_kernel<<<row_blocks(T), ROW_WARPS * 32, 0, st>>>(dz0, ids, dword, T, H, V, pad_id);
  if (check_launch("embed_word_bwd")) return 1;
  // BERT: the position id is the index in the sequence, one block per position owns its row of the table;
  // RoBERTa: scatter by the position ids the forward kept, and only the type row comes out of the per-position kernel
  embed_pos_type_bwd_kernel<<<S, 128 * PT_GROUPS, 0, st>>>(dz0, token_type, pos_mode == 0 ? dpos : nullptr, dtype, T / S,
                                                          S, H, TV);
  if (check_launch("embed_pos_type_bwd")) return 1;
  if (pos_mode != 0) {
    if (pos_ids == nullptr) {
      set_error("embedding backward: position ids of the forward are missing");
      return 1;
    }
    embed_pos_scatter_bwd_kernel<<<row_blocks(T), ROW_WARPS * 32, 0, st>>>(dz0, pos_ids, dpos, T, H, P, pad_id);
    return check_launch("embed_pos_scatter_bwd");
  }
  return 0;
}

int launch_prepare_inputs(const int32_t* ids32, int64_t* ids, int64_t* token_type, uint8_t* mask, int B, int S,
                          int sep_id, int type_coef, int* missing_sep, cudaStream_t st) {
  prepare_inputs_kernel<<<B, 128, 0, st>>>(ids32, ids, token_type, mask, S, sep_id, type_coef, missing_sep);
  return check_launch("prepare_inputs");
}

}  // namespace bqa
