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

__global__ void __launch_bounds__(ROW_WARPS * 32)
ln_fwd_kernel(const bf16* __restrict__ z, const float* __restrict__ gamma, const float* __restrict__ beta,
              bf16* __restrict__ y, int T, int H, float eps) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  RowVec r, g, b;
  load_row_bf16(z + static_cast<size_t>(row) * H, H, lane, r);
  load_vec_f32(gamma, H, lane, g);
  load_vec_f32(beta, H, lane, b);
  float mean, rstd;
  row_stats(r, H, lane, eps, mean, rstd);
  DropKey none{0, 0, 0, 1.0f};
  ln_apply(r, g, b, H, lane, mean, rstd, none, row);
  store_row_bf16(y + static_cast<size_t>(row) * H, H, lane, r);
}

// z0 = word[ids] + pos[s] + type[tt]  (fp32 tables, fp32 sum, stored as bf16)
// x0 = dropout(LayerNorm(z0))
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_ln_fwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ token_type,
                    const float* __restrict__ word, const float* __restrict__ pos, const float* __restrict__ type,
                    const float* __restrict__ gamma, const float* __restrict__ beta, bf16* __restrict__ z0,
                    bf16* __restrict__ x0, int T, int S, int H, int V, int TV, float eps, DropKey drop,
                    int* __restrict__ err_flag) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  long long id = ids[row];
  long long tt = token_type != nullptr ? token_type[row] : 0;
  if (id < 0 || id >= V || tt < 0 || tt >= TV) {
    if (lane == 0) atomicExch(err_flag, 1);  // torch raises IndexError here; surfaced by the host wrapper
    id = 0;
    tt = 0;
  }
  const int s = row % S;
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
// Rows are distributed round-robin over a persistent grid; per-warp column partials live in
// registers and are reduced once per block (smem) then once per grid (atomicAdd).
__global__ void __launch_bounds__(ROW_WARPS * 32)
ln_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z, const float* __restrict__ gamma,
              bf16* __restrict__ dz, bf16* __restrict__ dz_masked, float* __restrict__ dgamma,
              float* __restrict__ dbeta, float* __restrict__ dbias, int T, int H, float eps, DropKey drop_in,
              DropKey drop_out) {
  __shared__ float red[ROW_WARPS][MAXV * 32 * 8 + 8];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  RowVec g;
  load_vec_f32(gamma, H, lane, g);
  RowVec acc_g, acc_b, acc_bias;
#pragma unroll
  for (int j = 0; j < MAXV; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc_g.v[j][i] = acc_b.v[j][i] = acc_bias.v[j][i] = 0.0f;

  for (int row = blockIdx.x * ROW_WARPS + warp; row < T; row += gridDim.x * ROW_WARPS) {
    RowVec x, d;
    load_row_bf16(z + static_cast<size_t>(row) * H, H, lane, x);
    load_row_bf16(dy + static_cast<size_t>(row) * H, H, lane, d);
    float mean, rstd;
    row_stats(x, H, lane, eps, mean, rstd);
    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
    for (int j = 0; j < MAXV; ++j) {
      const int c = (j * 32 + lane) * 8;
      if (c < H) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float dyv = d.v[j][i];
          if (drop_in.thresh != 0u)
            dyv = drop_keep(drop_in, static_cast<uint32_t>(row) * H + c + i) ? dyv * drop_in.scale : 0.0f;
          const float xhat = (x.v[j][i] - mean) * rstd;
          const float dxhat = dyv * g.v[j][i];
          acc_g.v[j][i] += dyv * xhat;
          acc_b.v[j][i] += dyv;
          s1 += dxhat;
          s2 += dxhat * xhat;
          x.v[j][i] = xhat;
          d.v[j][i] = dxhat;
        }
      }
    }
    s1 = warp_sum(s1) / static_cast<float>(H);
    s2 = warp_sum(s2) / static_cast<float>(H);
#pragma unroll
    for (int j = 0; j < MAXV; ++j)
#pragma unroll
      for (int i = 0; i < 8; ++i) d.v[j][i] = rstd * (d.v[j][i] - s1 - x.v[j][i] * s2);
    store_row_bf16(dz + static_cast<size_t>(row) * H, H, lane, d);
    if (dz_masked != nullptr) {
      // what the GEMM-side consumers (dgrad, wgrad, bias grad) read is the bf16-rounded value
      round_row_bf16(d);
#pragma unroll
      for (int j = 0; j < MAXV; ++j) {
        const int c = (j * 32 + lane) * 8;
        if (c < H) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float m = d.v[j][i];
            if (drop_out.thresh != 0u)
              m = drop_keep(drop_out, static_cast<uint32_t>(row) * H + c + i) ? m * drop_out.scale : 0.0f;
            d.v[j][i] = m;
          }
        }
      }
      if (dz_masked != dz) store_row_bf16(dz_masked + static_cast<size_t>(row) * H, H, lane, d);
    }
    if (dbias != nullptr) {
      round_row_bf16(d);
#pragma unroll
      for (int j = 0; j < MAXV; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc_bias.v[j][i] += d.v[j][i];
    }
  }

  // block reduction of the three column partials, one at a time through the same buffer
  for (int which = 0; which < 3; ++which) {
    float* out = which == 0 ? dgamma : (which == 1 ? dbeta : dbias);
    if (out == nullptr) continue;  // uniform across the block
    const RowVec& a = which == 0 ? acc_g : (which == 1 ? acc_b : acc_bias);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < MAXV; ++j)
#pragma unroll
      for (int i = 0; i < 8; ++i) red[warp][(j * 32 + lane) * 8 + i] = a.v[j][i];
    __syncthreads();
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
      float s = 0.0f;
#pragma unroll
      for (int w = 0; w < ROW_WARPS; ++w) s += red[w][c];
      atomicAdd(out + c, s);
    }
  }
}

// out[n] += sum_t X[t][n]  (bias gradients of the fused QKV and intermediate projections)
__global__ void __launch_bounds__(256)
colsum_kernel(const bf16* __restrict__ X, int ld, float* __restrict__ out, int T, int N) {
  __shared__ float red[8][32 * 8 + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = (blockIdx.x * 32 + tx) * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (c < N) {
    for (int r = blockIdx.y * 8 + ty; r < T; r += gridDim.y * 8) {
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
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_bwd_kernel(const bf16* __restrict__ dz0, const int64_t* __restrict__ ids, const int64_t* __restrict__ token_type,
                 float* __restrict__ dword, float* __restrict__ dpos, float* __restrict__ dtype, int T, int S,
                 int H, int pad_id) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROW_WARPS + (threadIdx.x >> 5);
  if (row >= T) return;
  RowVec d;
  load_row_bf16(dz0 + static_cast<size_t>(row) * H, H, lane, d);
  const long long id = ids[row];
  const long long tt = token_type != nullptr ? token_type[row] : 0;
  float* w = dword + static_cast<size_t>(id) * H;
  float* p = dpos + static_cast<size_t>(row % S) * H;
  float* t = dtype + static_cast<size_t>(tt) * H;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = (j * 32 + lane) * 8;
    if (c < H) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (id != pad_id) atomicAdd(w + c + i, d.v[j][i]);
        atomicAdd(p + c + i, d.v[j][i]);
        atomicAdd(t + c + i, d.v[j][i]);
      }
    }
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
  ln_fwd_kernel<<<row_blocks(T), ROW_WARPS * 32, 0, st>>>(z, gamma, beta, y, T, H, eps);
  return check_launch("ln_fwd");
}

int launch_embed_ln_fwd(const int64_t* ids, const int64_t* token_type, const float* word, const float* pos,
                        const float* type, const float* gamma, const float* beta, bf16* z0, bf16* x0, int T, int S,
                        int H, int V, int TV, float eps, DropKey drop, int* err_flag, cudaStream_t st) {
  if (H % 8 != 0 || H > MAXV * 256) {
    set_error("embeddings: H=%d must be a multiple of 8 and <= %d", H, MAXV * 256);
    return 1;
  }
  embed_ln_fwd_kernel<<<row_blocks(T), ROW_WARPS * 32, 0, st>>>(ids, token_type, word, pos, type, gamma, beta, z0,
                                                              x0, T, S, H, V, TV, eps, drop, err_flag);
  return check_launch("embed_ln_fwd");
}

int launch_ln_bwd(const bf16* dy, const bf16* z, const float* gamma, bf16* dz, bf16* dz_masked, float* dgamma,
                  float* dbeta, float* dbias, int T, int H, float eps, DropKey drop_in, DropKey drop_out,
                  int num_sms, cudaStream_t st) {
  if (H % 8 != 0 || H > MAXV * 256) {
    set_error("layernorm backward: H=%d must be a multiple of 8 and <= %d", H, MAXV * 256);
    return 1;
  }
  int grid = num_sms * 4;
  if (grid > row_blocks(T)) grid = row_blocks(T);
  ln_bwd_kernel<<<grid, ROW_WARPS * 32, 0, st>>>(dy, z, gamma, dz, dz_masked, dgamma, dbeta, dbias, T, H, eps,
                                                 drop_in, drop_out);
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

int launch_embed_bwd(const bf16* dz0, const int64_t* ids, const int64_t* token_type, float* dword, float* dpos,
                     float* dtype, int T, int S, int H, int pad_id, cudaStream_t st) {
  embed_bwd_kernel<<<row_blocks(T), ROW_WARPS * 32, 0, st>>>(dz0, ids, token_type, dword, dpos, dtype, T, S, H,
                                                             pad_id);
  return check_launch("embed_bwd");
}

}  // namespace bqa
