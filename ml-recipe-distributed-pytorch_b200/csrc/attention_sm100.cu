// attention_sm100.cu — fused multi-head self-attention for sm_100a (head dim 64, S <= 512).
//
// Reference: BertSelfAttention.forward + eager_attention_forward
// (site-packages/transformers/models/bert/modeling_bert.py:115-140,168-207):
//   P = softmax(Q K^T / sqrt(d) + mask) ; P = dropout(P) ; ctx = P V
// and its autograd backward.  The reference materialises the [B,A,S,S] score tensor in HBM
// several times per layer; here scores live only in tensor memory:
//
//   forward  (persistent; a CTA walks a contiguous range of (batch, head, 128-query tile) tiles):
//     TMA -> smem Q, K, V | tcgen05.mma S = Q K^T into TMEM (128 lanes x S columns, fp32)
//     | 512 softmax threads, four per query row: row max, exp2, row sum straight out of TMEM
//     | P (bf16, unscaled, dropped) -> 128B-swizzled smem | tcgen05.mma O += P V (V is the MN-major operand)
//     | O * (dropout scale / rowsum) -> ctx, log-sum-exp -> lse2.
//   backward = two kernels with the same structure, each recomputing its score tile:
//     dQ  kernel (persistent, CTA = 128-query tile): S, dP = dO V^T ; dS = P (dP - D) ; dQ += dS K
//     dKV kernel (two CTAs per SM, CTA = 128-key tile, 64-query chunks), transposed problem so that
//                rows are keys: S^T = K Q^T, dP^T = V dO^T ; dV += P^T dO ; dK += dS^T Q
//   with D = rowsum(dO * O) from a small HBM-bound pre-pass.
// Dropout decisions are the stateless hash of (batch, head, query, key), identical in all kernels.
#include "attention_sm100.cuh"

#include <type_traits>

#include "bqa_host.cuh"
#include "gemm_sm100.cuh"
#include "kernels.cuh"

namespace bqa {

namespace {

constexpr int CH = 16384;  // bytes of one 128 x 64 bf16 tile
constexpr int ROW_THREADS = 512;               // 16 warps: 4 threads per tile row (column quarters)
constexpr int ATT_THREADS = ROW_THREADS + 32;  // + 1 control warp (TMA + MMA issue, TMEM alloc)
constexpr int CTRL_WARP = ROW_THREADS / 32;
// dK/dV kernel: two CTAs per SM, 8 row warps (2 threads per tile row) + 1 control warp each
constexpr int HC = 8192;  // bytes of one 64 x 64 bf16 tile
constexpr int BW_ROW_THREADS = 256;
constexpr int BW_THREADS = BW_ROW_THREADS + 32;
constexpr int BW_CTRL_WARP = BW_ROW_THREADS / 32;

struct AttnArgs {
  int B, S, A, H, nchunk, nchunk64, nchunk96, tmem_cols;
  float scale, scale_log2;
  const uint8_t* key_mask;
  bf16* ctx;
  float* lse2;
  float* dsum;
  bf16* dqkv;
  uint32_t* keep;  // [B][A][S][attn_keep_words(S)] dropout keep bits, or null
  float* dbias;    // [3H] fp32 or null: column sums of dqkv (gradient of the fused QKV bias), accumulated
  DropKey drop;
};

#ifdef BQA_ATTN_TRACE
__device__ unsigned long long g_attn_trace[4096 * 16];
#define ATRACE(slot)                                                                                        \
  do {                                                                                                      \
    const unsigned bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                     \
    if (bl < 4096) g_attn_trace[bl * 16 + (slot)] = clock64();                                              \
  } while (0)
// accumulated phase durations of one thread over the CTA's life: ATACC(k) adds the time since the
// previous ATACC to bucket k, ATACC_WRITE stores buckets 1..8 next to the ATRACE slots 0 and 9
#define ATACC_DECL long long tacc_[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}; long long tprev_ = clock64()
#define ATACC(k)                                                                                             \
  do {                                                                                                      \
    const long long now_ = clock64();                                                                       \
    tacc_[k] += now_ - tprev_;                                                                              \
    tprev_ = now_;                                                                                          \
  } while (0)
#define ATACC_WRITE()                                                                                        \
  do {                                                                                                      \
    const unsigned bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                     \
    if (bl < 4096)                                                                                          \
      for (int k_ = 1; k_ < 9; ++k_) g_attn_trace[bl * 16 + k_] = tacc_[k_];                                \
  } while (0)
// SM id and wall-clock stamps of a CTA's life (slots 10-12): which CTAs were co-resident
#define ATRACE_WHERE(slot)                                                                                  \
  do {                                                                                                      \
    const unsigned bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                     \
    unsigned smid_;                                                                                         \
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid_));                                                      \
    if (bl < 4096) {                                                                                        \
      g_attn_trace[bl * 16 + 10] = smid_;                                                                   \
      g_attn_trace[bl * 16 + (slot)] = global_timer_ns();                                                   \
    }                                                                                                       \
  } while (0)
// MMA-warp buckets (slots 13..15), same accumulate-since-last-call scheme
#define MTACC_DECL long long macc_[3] = {0, 0, 0}; long long mprev_ = clock64()
#define MTACC(k)                                                                                             \
  do {                                                                                                      \
    const long long now_ = clock64();                                                                       \
    macc_[k] += now_ - mprev_;                                                                              \
    mprev_ = now_;                                                                                          \
  } while (0)
#define MTACC_WRITE()                                                                                        \
  do {                                                                                                      \
    const unsigned bl = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                     \
    if (bl < 4096 && (threadIdx.x & 31) == 0)                                                               \
      for (int k_ = 0; k_ < 3; ++k_) g_attn_trace[bl * 16 + 13 + k_] = macc_[k_];                           \
  } while (0)
#else
#define MTACC_DECL do {} while (0)
#define MTACC(k) do {} while (0)
#define MTACC_WRITE() do {} while (0)
#define ATRACE_WHERE(slot) do {} while (0)
#define ATRACE(slot) do {} while (0)
#define ATACC_DECL do {} while (0)
#define ATACC(k) do {} while (0)
#define ATACC_WRITE() do {} while (0)
#endif

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// D[128 x N] (+)= A[128 x 64] * B[N x 64]^T, both operands K-major tiles of 128-byte rows
__device__ __forceinline__ void mma_k64(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc) {
#pragma unroll
  for (int k = 0; k < 4; ++k)
    umma_bf16(tmem_d, umma_desc_kmajor(a_addr + k * 32), umma_desc_kmajor(b_addr + k * 32), idesc, k > 0 ? 1u : 0u);
}

// D[128 x 64] (+)= A[128 x 128] * B[128 x 64]; A = two K-major atoms (CH apart), B = MN-major tile
__device__ __forceinline__ void mma_k128_mn(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc,
                                            bool accumulate_first) {
#pragma unroll
  for (int k = 0; k < 8; ++k)
    umma_bf16(tmem_d, umma_desc_kmajor(a_addr + (k >> 2) * CH + (k & 3) * 32),
              umma_desc_mnmajor(b_addr + k * 2048, CH), idesc, (accumulate_first || k > 0) ? 1u : 0u);
}

// D[128 x 64] (+)= A[128 x 64] * B[64 x 64]; A = one K-major atom column, B = MN-major tile of 64 rows
__device__ __forceinline__ void mma_k64_mn(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc,
                                           bool accumulate_first) {
#pragma unroll
  for (int k = 0; k < 4; ++k)
    umma_bf16(tmem_d, umma_desc_kmajor(a_addr + k * 32), umma_desc_mnmajor(b_addr + k * 2048, CH), idesc,
              (accumulate_first || k > 0) ? 1u : 0u);
}

// 8 consecutive bf16 of row r, 16-byte chunk c16 (0..15) of a [128 x 128] K-major operand tile
__device__ __forceinline__ void st_tile16(uint8_t* tile, int r, int c16, const float* f) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]);
  o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]);
  o.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(tile + (c16 >> 3) * CH + r * 128 + (((c16 & 7) ^ (r & 7)) << 4)) = o;
}

// same, from already packed bf16 pairs
__device__ __forceinline__ void st_tile16p(uint8_t* tile, int r, int c16, uint4 o) {
  *reinterpret_cast<uint4*>(tile + (c16 >> 3) * CH + r * 128 + (((c16 & 7) ^ (r & 7)) << 4)) = o;
}

__device__ __forceinline__ void st_row_bf16_16(bf16* dst, const float* f) {
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    uint4 o;
    o.x = pack_bf16x2(f[8 * q], f[8 * q + 1]);
    o.y = pack_bf16x2(f[8 * q + 2], f[8 * q + 3]);
    o.z = pack_bf16x2(f[8 * q + 4], f[8 * q + 5]);
    o.w = pack_bf16x2(f[8 * q + 6], f[8 * q + 7]);
    *reinterpret_cast<uint4*>(dst + 8 * q) = o;
  }
}

// 16 consecutive bf16 (columns 16 qd .. 16 qd + 15) of row r of a [128 x 64] output tile laid out for a
// SWIZZLE_128B TMA store
__device__ __forceinline__ void st_out16(uint8_t* tile, int r, int qd, const float* f) {
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    uint4 o;
    o.x = pack_bf16x2(f[8 * q], f[8 * q + 1]);
    o.y = pack_bf16x2(f[8 * q + 2], f[8 * q + 3]);
    o.z = pack_bf16x2(f[8 * q + 4], f[8 * q + 5]);
    o.w = pack_bf16x2(f[8 * q + 6], f[8 * q + 7]);
    *reinterpret_cast<uint4*>(tile + r * 128 + ((((qd * 2 + q)) ^ (r & 7)) << 4)) = o;
  }
}

// sum over `rows` consecutive rows (from row0) of column `col` (0..63) of a [rows x 64] bf16 tile staged in the
// SWIZZLE_128B layout st_out16 / st_tile16 write
__device__ __forceinline__ float staged_colsum(const uint8_t* tile, int row0, int rows, int col) {
  float s = 0.0f;
  for (int rr = row0; rr < row0 + rows; ++rr) {
    const uint16_t raw = *reinterpret_cast<const uint16_t*>(tile + rr * 128 + ((((col >> 3) ^ (rr & 7))) << 4) + (col & 7) * 2);
    s += __uint_as_float(static_cast<uint32_t>(raw) << 16);
  }
  return s;
}

__device__ __forceinline__ uint8_t* align1024(uint8_t* p) { return p + ((1024u - (smem_u32(p) & 1023u)) & 1023u); }

// additive key bias (0 / -inf) of batch row b and, per 32 keys, the bit mask of attended keys.
// Called by the ROW_THREADS (= 512 >= Sp) softmax threads, whole warps.
__device__ __forceinline__ void fill_key_bias(float* kbias, uint32_t* kvalid, const uint8_t* key_mask, int b, int S,
                                              int Sp, int tid) {
  const bool ok = tid < S && (key_mask == nullptr || key_mask[static_cast<size_t>(b) * S + tid] != 0);
  if (tid < Sp) kbias[tid] = ok ? 0.0f : -INFINITY;
  const uint32_t w = __ballot_sync(0xffffffffu, ok);
  if ((tid & 31) == 0) kvalid[tid >> 5] = w;
}

// =========================================================================================
// dropout keep bits of one attention call, in two layouts per (batch, head) row r (W = 4 * nchunk words):
//   keep[(bh * S + r) * 2 W + w]      bit k : (query r, key 32 w + k)      forward, dQ kernel
//   keep[(bh * S + r) * 2 W + W + w]  bit q : (query 32 w + q, key r)      dK/dV kernel (rows are keys)
// One warp per 32 queries x 32 keys block: lane = query; the key-major words are the 32 x 32 bit
// transpose of the query-major ones (5 shuffle steps).
// =========================================================================================
__global__ void __launch_bounds__(256)
attn_keep_bits_kernel(uint32_t* __restrict__ keep, int BH, int S, int W, DropKey dk) {
  const int lane = threadIdx.x & 31;
  const int qblocks = (S + 31) >> 5;
  const long long items = static_cast<long long>(BH) * qblocks * W;
  const long long stride = static_cast<long long>(gridDim.x) * (blockDim.x >> 5);
  for (long long it = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); it < items;
       it += stride) {
    const int w = static_cast<int>(it % W);
    const int qb = static_cast<int>((it / W) % qblocks);
    const int bh = static_cast<int>(it / (static_cast<long long>(W) * qblocks));
    const int q = qb * 32 + lane, key0 = w * 32;
    uint32_t bits = 0u;
    if (q < S && key0 < S) {
      bits = keep_bits32(dk, (static_cast<uint32_t>(bh) * S + q) * static_cast<uint32_t>(S) + key0);
      if (key0 + 32 > S) bits &= (1u << (S - key0)) - 1u;  // keys beyond S belong to the next row's indices
    }
    if (q < S) keep[(static_cast<size_t>(bh) * S + q) * (2 * W) + w] = bits;
    // 32 x 32 bit transpose across the warp: afterwards bit i of lane k = bit k of lane i
    uint32_t x = bits;
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
      const uint32_t m = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u
                                                                                                   : 0x55555555u;
      const uint32_t o = __shfl_xor_sync(0xffffffffu, x, j);
      x = (lane & j) ? (((o >> j) & m) | (x & ~m)) : ((x & m) | ((o & m) << j));
    }
    const int key = key0 + lane;
    if (key < S) keep[(static_cast<size_t>(bh) * S + key) * (2 * W) + W + qb] = x;
  }
}

// =========================================================================================
// forward
// =========================================================================================
// Two CTAs per SM, each: 4 softmax warps (one thread per query row of a 128-query tile), one MMA-issuing warp, one
// TMA warp.  A CTA walks a contiguous range of (batch, head, 128-query tile) tiles; per tile the keys stream
// through in chunks of KC = 96 (K and V tiles in a 3-stage TMA ring), online softmax:
//   S_g = Q K_g^T            tcgen05.mma into one of TWO score buffers in TMEM (96 fp32 columns each)
//   softmax threads          read the 96 scores of their row, update the running maximum, exponentiate, and
//                            write P_g as bf16 pairs IN PLACE over the first 48 columns of the same buffer
//   O += P_g V_g             tcgen05.mma with the A operand (P) read from TMEM, V the MN-major smem operand
// so scores and probabilities never touch shared memory, and the score MMA of chunk g+1 (other buffer) runs
// while the threads work on chunk g.  TMEM per CTA: 2 x 96 score columns + 64 output columns = 256.
// The running maximum is lazy (the output accumulator is rescaled only when the maximum grows by more than
// 2^8, decided per warp), which keeps the result exact: the final normalisation uses the same maximum.
// Barrier phases are derived from the per-CTA chunk / tile counters, never from local toggles.
constexpr int KC = 96;                 // keys per chunk
constexpr int KCB = KC * 128;          // bytes of one K or V chunk tile (96 rows x 128 B)
constexpr int FW_SOFT = 128;           // softmax threads
constexpr int FW_THREADS = FW_SOFT + 64;
constexpr int FW_MMA_WARP = 4, FW_TMA_WARP = 5;
constexpr int FW_RING = 3;             // K / V ring depth
constexpr uint32_t FW_COL_O = 192;     // output accumulator columns [192, 256)
constexpr float FW_RESCALE_LOG2 = 8.0f;

template <bool KB>
__global__ void __launch_bounds__(FW_THREADS, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_kv96,
                const __grid_constant__ CUtensorMap tm_ctx, const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sQ = smem;                    // 2 x CH; a finished tile's buffer stages its output tile
  uint8_t* sK = sQ + 2 * CH;             // FW_RING x KCB
  uint8_t* sV = sK + FW_RING * KCB;      // FW_RING x KCB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + FW_RING * KCB);
  uint64_t* q_full = bars + 0;    // [2]
  uint64_t* q_free = bars + 2;    // [2] last score MMA of the tile done + output store has read the buffer
  uint64_t* k_full = bars + 4;    // [3]
  uint64_t* k_free = bars + 7;    // [3]
  uint64_t* v_full = bars + 10;   // [3]
  uint64_t* v_free = bars + 13;   // [3]
  uint64_t* s_full = bars + 16;   // [2]
  uint64_t* p_full = bars + 18;   // [2]
  uint64_t* pv_done = bars + 20;  // [2]
  uint64_t* o_full = bars + 22;
  uint64_t* o_free = bars + 23;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
  uint32_t* kvalid = tmem_slot + 2;  // [18] bit k of word w: key 32 w + k is attended
  // keep words of the softmax threads' rows: [2 buffers][3 words][FW_SOFT threads], filled by cp.async
  uint32_t* kw_stage = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(bars) + 512);

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx();
  const int S = a.S, NQ = a.nchunk, NC = a.nchunk96;
  const int tiles = a.B * a.A * NQ;
  const int t0 = static_cast<int>(static_cast<long long>(blockIdx.x) * tiles / gridDim.x);
  const int t1 = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * tiles / gridDim.x);
  if (tid == 0) ATRACE(0);

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_free[i], 2);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], FW_SOFT);
      mbar_init(&pv_done[i], 1);
    }
    for (int i = 0; i < FW_RING; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_free[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_free[i], 1);
    }
    mbar_init(o_full, 1);
    mbar_init(o_free, FW_SOFT);
    fence_mbar_init();
  }
  if (warp == FW_MMA_WARP) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  if (tid == FW_TMA_WARP * 32) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_kv96);
    tma_prefetch_desc(&tm_ctx);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();     // launched with programmatic serialization: barrier set-up and TMEM allocation above overlapped the
                  // previous kernel's tail; its results are complete from here on
  pdl_trigger();  // the next kernel may start its prologue; it waits for this grid before reading

  if (warp == FW_TMA_WARP) {
    // ---------------- TMA producer: Q per tile (double-buffered), K / V chunks through the ring ----------------
    const bool leader = elect_one();
    uint32_t g = 0, slot = 0, round = 0;  // chunk counter, g % FW_RING, g / FW_RING
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int bh = t / NQ, qt = t - bh * NQ, b = bh / a.A, h = bh - b * a.A;
      const int qb = it & 1;
      if (it >= 2) mbar_wait(&q_free[qb], ((it >> 1) - 1) & 1);
      if (leader) {
        mbar_arrive_expect_tx(&q_full[qb], CH);
        tma_load_3d(sQ + qb * CH, &tm_qkv, h * 64, qt * 128, b, &q_full[qb]);
      }
      __syncwarp();
      for (int j = 0; j < NC; ++j, ++g) {
        if (round > 0) mbar_wait(&k_free[slot], (round - 1) & 1);
        if (leader) {
          mbar_arrive_expect_tx(&k_full[slot], KCB);
          tma_load_3d(sK + slot * KCB, &tm_kv96, a.H + h * 64, j * KC, b, &k_full[slot]);
        }
        __syncwarp();
        if (round > 0) mbar_wait(&v_free[slot], (round - 1) & 1);
        if (leader) {
          mbar_arrive_expect_tx(&v_full[slot], KCB);
          tma_load_3d(sV + slot * KCB, &tm_kv96, 2 * a.H + h * 64, j * KC, b, &v_full[slot]);
        }
        __syncwarp();
        if (++slot == FW_RING) {
          slot = 0;
          ++round;
        }
      }
    }
  } else if (warp == FW_MMA_WARP) {
    // ---------------- MMA issuer ----------------
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, KC, 0, 0);
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);
    // the last chunk of a row of keys may be partial: its score MMA computes only the columns that hold keys
    // (rounded up to the MMA's granularity of 16) and its P V MMA only those k-steps
    const int last_keys = S - (NC - 1) * KC;                  // 1 .. KC
    const int last_ksteps = (last_keys + 15) >> 4;            // k-steps of 16 keys
    const uint32_t idesc_s_last = umma_idesc_bf16(128, last_ksteps * 16, 0, 0);
    const bool leader = elect_one();
    // O += P V for chunk gp (the jp-th chunk of this CTA's itp-th tile, ring slot sp of round rp)
    auto issue_pv = [&](uint32_t gp, int itp, int jp, uint32_t sp, uint32_t rp) {
      const uint32_t bp = gp & 1u;
      mbar_wait(&p_full[bp], (gp >> 1) & 1u);
      mbar_wait(&v_full[sp], rp & 1u);
      if (jp == 0 && itp > 0) mbar_wait(o_free, (itp - 1) & 1);  // the previous tile's output has been read
      tc_fence_after();
      if (leader) {
        const uint32_t v_addr = smem_u32(sV + sp * KCB);
        const int ksteps = jp == NC - 1 ? last_ksteps : KC / 16;
#pragma unroll
        for (int k = 0; k < KC / 16; ++k)
          if (k < ksteps)
            umma_bf16_ts(tmem + FW_COL_O, tmem + bp * KC + k * 8, umma_desc_mnmajor(v_addr + k * 2048, KCB), idesc_o,
                         (jp > 0 || k > 0) ? 1u : 0u);
        umma_commit(&pv_done[bp]);
        umma_commit(&v_free[sp]);
        if (jp == NC - 1) umma_commit(o_full);
      }
      __syncwarp();
    };
    uint32_t g = 0, slot = 0, round = 0;
    int p_it = 0, p_j = 0;        // tile / chunk position of chunk g - 1
    uint32_t p_slot = 0, p_round = 0;
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int qb = it & 1;
      for (int j = 0; j < NC; ++j, ++g) {
        const uint32_t bsel = g & 1u;
        // a new tile's first score MMA must not delay the previous tile's last P V
        if (j == 0 && g > 0) issue_pv(g - 1, p_it, p_j, p_slot, p_round);
        if (j == 0) mbar_wait(&q_full[qb], (it >> 1) & 1);
        if (g >= 2) mbar_wait(&pv_done[bsel], ((g >> 1) - 1u) & 1u);  // P of chunk g - 2 consumed: buffer free
        mbar_wait(&k_full[slot], round & 1u);
        tc_fence_after();
        if (leader) {
          const uint32_t q_addr = smem_u32(sQ + qb * CH), k_addr = smem_u32(sK + slot * KCB);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem + bsel * KC, umma_desc_kmajor(q_addr + k * 32), umma_desc_kmajor(k_addr + k * 32),
                      j == NC - 1 ? idesc_s_last : idesc_s, k > 0 ? 1u : 0u);
          umma_commit(&s_full[bsel]);
          umma_commit(&k_free[slot]);
          if (j == NC - 1) umma_commit(&q_free[qb]);
        }
        __syncwarp();
        if (j > 0) issue_pv(g - 1, p_it, p_j, p_slot, p_round);
        p_it = it;
        p_j = j;
        p_slot = slot;
        p_round = round;
        if (++slot == FW_RING) {
          slot = 0;
          ++round;
        }
      }
    }
    if (g > 0) issue_pv(g - 1, p_it, p_j, p_slot, p_round);
  } else {
    // ---------------- softmax threads: thread r owns query row r of the tile ----------------
    const int r = tid;
    const uint32_t lane_base = tmem + (static_cast<uint32_t>(warp * 32) << 16);
    const float sl2 = a.scale_log2;
    const bool dropping = a.drop.thresh != 0u;
    const int keep_words = 8 * NQ;  // per (batch, head, query) row; the first 4 NQ are query-major
    const int last_keys = S - (NC - 1) * KC;  // keys in the (possibly partial) last chunk
    int prev_b = -1;
    int pending_qb = -1;  // thread 0: output store whose staging buffer has not been released yet
    uint32_t g = 0;
    ATACC_DECL;
    if (tid == 0) ATRACE_WHERE(11);
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int bh = t / NQ, qt = t - bh * NQ, b = bh / a.A, h = bh - b * a.A;
      const int q = qt * 128 + r;
      const int qb = it & 1;
      if (KB && b != prev_b) {
        named_bar_sync(1, FW_SOFT);  // nobody still reads the previous batch row's key bits
        for (int base = 0; base < NC * KC; base += FW_SOFT) {
          const int key = base + tid;
          const bool ok = key < S && (a.key_mask == nullptr || a.key_mask[static_cast<size_t>(b) * S + key] != 0);
          const uint32_t w = __ballot_sync(0xffffffffu, ok);
          if ((tid & 31) == 0 && key < NC * KC) kvalid[key >> 5] = w;
        }
        named_bar_sync(1, FW_SOFT);
        prev_b = b;
      }
      float m_run = -INFINITY, l = 0.0f;
      // Keep bits of the row over the keys of one chunk (3 words), fetched one chunk ahead of their use — across the
      // tile boundary too — by cp.async into shared memory: a register prefetch was spilled by the compiler right
      // after the load (168 registers), which exposed the whole load latency in every chunk.  A thread reads back
      // only what it copied itself, so cp.async.wait_group is the only synchronisation.
      auto fetch_keep = [&](int tt, int j, uint32_t buf) {
        if (dropping && tt < t1) {
          const int bh_ = tt / NQ, q_ = (tt - bh_ * NQ) * 128 + r;
          const uint32_t* row = a.keep + (static_cast<size_t>(bh_) * S + (q_ < S ? q_ : 0)) * keep_words;
#pragma unroll
          for (int k = 0; k < 3; ++k)
            if (3 * j + k < 4 * NQ) cp_async_4(kw_stage + (buf * 3 + k) * FW_SOFT + tid, row + 3 * j + k);
        }
        cp_async_commit();  // always: the group count tracks the chunk count
      };
      if (it == 0) fetch_keep(t, 0, g & 1u);
      for (int j = 0; j < NC; ++j, ++g) {
        const uint32_t bsel = g & 1u;
        const uint32_t s_addr = lane_base + bsel * KC;
        if (j + 1 < NC) fetch_keep(t, j + 1, bsel ^ 1u); else fetch_keep(t + 1, 0, bsel ^ 1u);
        ATACC(8);
        mbar_wait(&s_full[bsel], (g >> 1) & 1u);
        ATACC(1);
        tc_fence_after();
        // 32-key pieces of this chunk that hold keys at all (a partial last chunk has fewer; the columns beyond
        // them were not written by the score MMA and are neither read nor exponentiated)
        const int pieces = (KB && j == NC - 1) ? (last_keys + 31) >> 5 : 3;  // (without KB, S is a multiple of KC)
        uint32_t v[3][32];
#pragma unroll
        for (int k = 0; k < 3; ++k)
          if (k < pieces) tmem_ld_32x32(s_addr + k * 32, v[k]);
        tmem_wait_ld();
        // chunk maximum of the scaled scores over the keys that are attended.  Masked keys (padding, keys beyond S)
        // are SELECTED away by their bit in kvalid, never added to: a column beyond the score MMA's width may hold
        // anything.
        float cmax = -INFINITY;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (k >= pieces) continue;
          float raw = -INFINITY;
          if (KB && kvalid[3 * j + k] != 0xffffffffu) {
            const uint32_t vb = kvalid[3 * j + k];
#pragma unroll
            for (int i = 0; i < 32; ++i) raw = fmaxf(raw, ((vb >> i) & 1u) ? __uint_as_float(v[k][i]) : -INFINITY);
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) raw = fmax3(raw, __uint_as_float(v[k][i]), __uint_as_float(v[k][i + 1]));
          }
          cmax = fmaxf(cmax, raw * sl2);
        }
        ATACC(2);
        const float m_new = fmaxf(m_run, cmax);
        const bool grow = m_new > m_run + FW_RESCALE_LOG2;  // also true for the first finite maximum
        if (j == 0) {
          m_run = m_new;
        } else if (__any_sync(0xffffffffu, grow)) {
          // rescale the running sum and the output accumulator (whole warp: TMEM accesses are warp-wide)
          const float alpha = grow ? fast_exp2(m_run - m_new) : 1.0f;
          if (grow) m_run = m_new;
          l *= alpha;
          mbar_wait(&pv_done[(g - 1) & 1u], ((g - 1) >> 1) & 1u);  // P V of the previous chunk has landed
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(lane_base + FW_COL_O + c * 32, o);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(lane_base + FW_COL_O + c * 32, o);
          }
          tmem_wait_st();
        }
        ATACC(3);
        const float m_use = m_run == -INFINITY ? 0.0f : m_run;
        cp_async_wait<1>();  // this chunk's keep words have landed (the next chunk's may still be in flight)
        // P is stored unscaled; the dropout scale 1 / (1 - p) is applied with the final 1 / rowsum
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (k >= pieces) continue;
          const uint32_t bits = (dropping && 3 * j + k < 4 * NQ) ? kw_stage[(bsel * 3 + k) * FW_SOFT + tid] : 0xffffffffu;
          uint32_t pk[16];
          auto piece = [&](auto masked) {
            const uint32_t vb = decltype(masked)::value ? kvalid[3 * j + k] : 0xffffffffu;
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float p0 = fast_exp2(fmaf(__uint_as_float(v[k][i]), sl2, -m_use));
              float p1 = fast_exp2(fmaf(__uint_as_float(v[k][i + 1]), sl2, -m_use));
              if (decltype(masked)::value) {
                p0 = ((vb >> i) & 1u) ? p0 : 0.0f;
                p1 = ((vb >> (i + 1)) & 1u) ? p1 : 0.0f;
              }
              l += p0 + p1;
              pk[i >> 1] = pack_bf16x2(((bits >> i) & 1u) ? p0 : 0.0f, ((bits >> (i + 1)) & 1u) ? p1 : 0.0f);
            }
          };
          if (KB && kvalid[3 * j + k] != 0xffffffffu) piece(std::true_type{}); else piece(std::false_type{});
          tmem_st_32x16(s_addr + k * 16, pk);
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(&p_full[bsel]);
        ATACC(4);
        if (tid == 0 && pending_qb >= 0) {
          // the previous tile's output store has long finished reading its staging buffer: release it
          tma_wait_group_read<0>();
          mbar_arrive(&q_free[pending_qb]);
          pending_qb = -1;
        }
      }
      // ---- tile epilogue: O * (dropout scale / rowsum) -> bf16 -> swizzled smem (this tile's Q buffer: every
      // MMA of the tile has completed) -> one TMA store, which also clips the rows beyond S
      ATACC(5);
      mbar_wait(o_full, it & 1);
      ATACC(6);
      tc_fence_after();
      const float inv = l > 0.0f ? a.drop.scale / l : 0.0f;
      uint8_t* stage = sQ + qb * CH;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32(lane_base + FW_COL_O + c * 32, o);
        tmem_wait_ld();
        if (c == 1) {
          tc_fence_before();
          mbar_arrive(o_free);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint4 w;
          w.x = pack_bf16x2(__uint_as_float(o[8 * u]) * inv, __uint_as_float(o[8 * u + 1]) * inv);
          w.y = pack_bf16x2(__uint_as_float(o[8 * u + 2]) * inv, __uint_as_float(o[8 * u + 3]) * inv);
          w.z = pack_bf16x2(__uint_as_float(o[8 * u + 4]) * inv, __uint_as_float(o[8 * u + 5]) * inv);
          w.w = pack_bf16x2(__uint_as_float(o[8 * u + 6]) * inv, __uint_as_float(o[8 * u + 7]) * inv);
          *reinterpret_cast<uint4*>(stage + r * 128 + (((c * 4 + u) ^ (r & 7)) << 4)) = w;
        }
      }
      fence_proxy_async_smem();
      named_bar_sync(2, FW_SOFT);
      if (tid == 0) {
        tma_store_3d(&tm_ctx, stage, h * 64, qt * 128, b);
        tma_commit_group();
        pending_qb = qb;
      }
      // a row whose keys are all masked has l == 0: store +inf so that the backward recomputes P = 0
      if (q < S) a.lse2[(static_cast<size_t>(b) * a.A + h) * S + q] = l > 0.0f ? m_run + log2f(l) : INFINITY;
      ATACC(7);
    }
    if (tid == 0) {
      ATACC_WRITE();
      ATRACE_WHERE(12);
    }
    if (tid == 0) {
      if (pending_qb >= 0) {
        tma_wait_group_read<0>();
        mbar_arrive(&q_free[pending_qb]);
      }
      tma_wait_group<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == FW_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
  if (tid == 0) ATRACE(9);
}

// =========================================================================================
// backward pre-pass: D[b][h][q] = sum_d dO[b,q,h,d] * O[b,q,h,d]
// =========================================================================================
__global__ void __launch_bounds__(128)
attn_dsum_kernel(const bf16* __restrict__ ctx, const bf16* __restrict__ dctx, float* __restrict__ dsum, int B, int S,
                 int A, int H) {
  pdl_trigger();  // the next kernel (a GEMM) may start its prologue now; it waits for this grid before reading
  const int lane = threadIdx.x & 31;
  const int t = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (t >= B * S) return;
  const int b = t / S, q = t - b * S;
  for (int vbase = 0; vbase < H / 8; vbase += 32) {
    const int vi = vbase + lane;  // 16-byte vector index in the row; 8 vectors per head
    float acc = 0.0f;
    if (vi < H / 8) {
      const uint4 o = __ldg(reinterpret_cast<const uint4*>(ctx + static_cast<size_t>(t) * H) + vi);
      const uint4 d = __ldg(reinterpret_cast<const uint4*>(dctx + static_cast<size_t>(t) * H) + vi);
      const float2 o0 = unpack_bf16x2(o.x), o1 = unpack_bf16x2(o.y), o2 = unpack_bf16x2(o.z), o3 = unpack_bf16x2(o.w);
      const float2 d0 = unpack_bf16x2(d.x), d1 = unpack_bf16x2(d.y), d2 = unpack_bf16x2(d.z), d3 = unpack_bf16x2(d.w);
      acc = o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y + o2.x * d2.x + o2.y * d2.y + o3.x * d3.x +
            o3.y * d3.y;
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if ((lane & 7) == 0 && vi < H / 8) dsum[(static_cast<size_t>(b) * A + vi / 8) * S + q] = acc;
  }
}

// =========================================================================================
// backward, dQ
// =========================================================================================
// Same organisation as the forward: two CTAs per SM, each with 8 row warps (thread (r, hq) = query row r of a
// 128-query tile, half hq of the chunk's key columns), one MMA-issuing warp and one TMA warp; contiguous
// (batch, head, query-tile) ranges per CTA.  Keys stream through in chunks of 64 (K ring of 3 because the dQ
// MMA reads K again, V ring of 2); Q / dO tiles are double-buffered across tiles.  Per chunk g:
//   S = Q K_g^T, dP = dO V_g^T     tcgen05.mma into TMEM (64 + 64 fp32 columns, single-buffered: the row threads
//                                  pull them into registers and hand the columns back at once)
//   dS = P (keep dP c - D scale)   in registers; written as bf16 pairs into 32 further TMEM columns
//   dQ += dS K_g                   tcgen05.mma with the A operand (dS) read from TMEM, K the MN-major operand
// The score MMAs of chunk g + 1 are issued before the dQ MMA of chunk g is waited for.
constexpr int DQ_ROW = 256;
constexpr int DQ_THREADS = DQ_ROW + 64;
constexpr int DQ_MMA_WARP = 8, DQ_TMA_WARP = 9;
constexpr int DQ_KRING = 3, DQ_VRING = 2;
constexpr uint32_t DQ_COL_S = 0, DQ_COL_DP = 64, DQ_COL_DS = 128, DQ_COL_DQ = 192;

template <bool KB>
__global__ void __launch_bounds__(DQ_THREADS, 2)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_kv64,
                   const __grid_constant__ CUtensorMap tm_dctx, const __grid_constant__ CUtensorMap tm_dqkv,
                   const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  uint8_t* sQ = smem;                      // 2 x CH; a finished tile's buffer stages its dQ tile
  uint8_t* sdO = sQ + 2 * CH;              // 2 x CH
  uint8_t* sK = sdO + 2 * CH;              // DQ_KRING x HC
  uint8_t* sV = sK + DQ_KRING * HC;        // DQ_VRING x HC
  float* kbias = reinterpret_cast<float*>(sV + DQ_VRING * HC);  // [512]
  float* s_red = kbias + 512;              // [8][64] partial column sums of the drained tile
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_red + 512);
  uint64_t* qdo_full = bars + 0;   // [2]
  uint64_t* qdo_free = bars + 2;   // [2] last score MMAs of the tile done + dQ store has read the staging buffer
  uint64_t* k_full = bars + 4;     // [3]
  uint64_t* k_free = bars + 7;     // [3]
  uint64_t* v_full = bars + 10;    // [2]
  uint64_t* v_free = bars + 12;    // [2]
  uint64_t* s_full = bars + 14;
  uint64_t* s_free = bars + 15;
  uint64_t* ds_full = bars + 16;
  uint64_t* dq_done = bars + 17;   // dQ MMA of a chunk complete: dS columns free
  uint64_t* dq_full = bars + 18;   // tile's accumulator complete
  uint64_t* dq_free = bars + 19;   // accumulator drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
  uint32_t* kvalid = tmem_slot + 2;  // [16]

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx();
  const int S = a.S, NQ = a.nchunk, NC = a.nchunk64;
  const int tiles = a.B * a.A * NQ;
  const int t0 = static_cast<int>(static_cast<long long>(blockIdx.x) * tiles / gridDim.x);
  const int t1 = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * tiles / gridDim.x);
  if (tid == 0) ATRACE(0);

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_free[i], 2);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_free[i], 1);
    }
    for (int i = 0; i < DQ_KRING; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_free[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, DQ_ROW);
    mbar_init(ds_full, DQ_ROW);
    mbar_init(dq_done, 1);
    mbar_init(dq_full, 1);
    mbar_init(dq_free, DQ_ROW);
    fence_mbar_init();
  }
  if (warp == DQ_MMA_WARP) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  if (tid == DQ_TMA_WARP * 32) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_kv64);
    tma_prefetch_desc(&tm_dctx);
    tma_prefetch_desc(&tm_dqkv);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();     // launched with programmatic serialization: barrier set-up and TMEM allocation above overlapped the
                  // previous kernel's tail; its results are complete from here on
  pdl_trigger();  // the next kernel may start its prologue; it waits for this grid before reading

  if (warp == DQ_TMA_WARP) {
    // ---------------- TMA producer ----------------
    const bool leader = elect_one();
    uint32_t ks = 0, kr = 0, vs = 0, vr = 0;  // ring slots and rounds
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int bh = t / NQ, qt = t - bh * NQ, b = bh / a.A, h = bh - b * a.A;
      const int qb = it & 1;
      if (it >= 2) mbar_wait(&qdo_free[qb], ((it >> 1) - 1) & 1);
      if (leader) {
        mbar_arrive_expect_tx(&qdo_full[qb], 2 * CH);
        tma_load_3d(sQ + qb * CH, &tm_qkv, h * 64, qt * 128, b, &qdo_full[qb]);
        tma_load_3d(sdO + qb * CH, &tm_dctx, h * 64, qt * 128, b, &qdo_full[qb]);
      }
      __syncwarp();
      for (int j = 0; j < NC; ++j) {
        if (kr > 0) mbar_wait(&k_free[ks], (kr - 1) & 1);
        if (leader) {
          mbar_arrive_expect_tx(&k_full[ks], HC);
          tma_load_3d(sK + ks * HC, &tm_kv64, a.H + h * 64, j * 64, b, &k_full[ks]);
        }
        __syncwarp();
        if (vr > 0) mbar_wait(&v_free[vs], (vr - 1) & 1);
        if (leader) {
          mbar_arrive_expect_tx(&v_full[vs], HC);
          tma_load_3d(sV + vs * HC, &tm_kv64, 2 * a.H + h * 64, j * 64, b, &v_full[vs]);
        }
        __syncwarp();
        if (++ks == DQ_KRING) {
          ks = 0;
          ++kr;
        }
        if (++vs == DQ_VRING) {
          vs = 0;
          ++vr;
        }
      }
    }
  } else if (warp == DQ_MMA_WARP) {
    // ---------------- MMA issuer ----------------
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64, 0, 0);
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);
    const bool leader = elect_one();
    // dQ += dS K for chunk gp (the jp-th chunk of tile itp, K ring slot sp of round rp)
    auto issue_dq = [&](uint32_t gp, int itp, int jp, uint32_t sp) {
      mbar_wait(ds_full, gp & 1u);
      if (jp == 0 && itp > 0) mbar_wait(dq_free, (itp - 1) & 1);  // the previous tile's accumulator has been drained
      tc_fence_after();
      if (leader) {
        const uint32_t k_addr = smem_u32(sK + sp * HC);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_ts(tmem + DQ_COL_DQ, tmem + DQ_COL_DS + k * 8, umma_desc_mnmajor(k_addr + k * 2048, HC), idesc_o,
                       (jp > 0 || k > 0) ? 1u : 0u);
        umma_commit(dq_done);
        umma_commit(&k_free[sp]);
        if (jp == NC - 1) umma_commit(dq_full);
      }
      __syncwarp();
    };
    uint32_t g = 0, ks = 0, kr = 0, vs = 0, vr = 0;
    int p_it = 0, p_j = 0;
    uint32_t p_ks = 0;
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int qb = it & 1;
      for (int j = 0; j < NC; ++j, ++g) {
        // a new tile's first score MMAs must not delay the previous tile's last dQ MMA
        if (j == 0 && g > 0) issue_dq(g - 1, p_it, p_j, p_ks);
        if (j == 0) mbar_wait(&qdo_full[qb], (it >> 1) & 1);
        mbar_wait(&k_full[ks], kr & 1u);
        mbar_wait(&v_full[vs], vr & 1u);
        if (g > 0) mbar_wait(s_free, (g - 1) & 1u);  // the previous chunk's scores are in registers
        tc_fence_after();
        if (leader) {
          const uint32_t q_addr = smem_u32(sQ + qb * CH), do_addr = smem_u32(sdO + qb * CH);
          const uint32_t k_addr = smem_u32(sK + ks * HC), v_addr = smem_u32(sV + vs * HC);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem + DQ_COL_S, umma_desc_kmajor(q_addr + k * 32), umma_desc_kmajor(k_addr + k * 32), idesc_s,
                      k > 0 ? 1u : 0u);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem + DQ_COL_DP, umma_desc_kmajor(do_addr + k * 32), umma_desc_kmajor(v_addr + k * 32), idesc_s,
                      k > 0 ? 1u : 0u);
          umma_commit(s_full);
          umma_commit(&v_free[vs]);
          if (j == NC - 1) umma_commit(&qdo_free[qb]);
        }
        __syncwarp();
        if (j > 0) issue_dq(g - 1, p_it, p_j, p_ks);
        p_it = it;
        p_j = j;
        p_ks = ks;
        if (++ks == DQ_KRING) {
          ks = 0;
          ++kr;
        }
        if (++vs == DQ_VRING) {
          vs = 0;
          ++vr;
        }
      }
    }
    if (g > 0) issue_dq(g - 1, p_it, p_j, p_ks);
  } else {
    // ---------------- row threads ----------------
    const int r = tid & 127, hq = tid >> 7;
    const uint32_t lane_base = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    const float sl2 = a.scale_log2;
    const bool dropping = a.drop.thresh != 0u;
    const float c1 = a.drop.scale * a.scale;
    const int keep_words = 8 * NQ;
    int prev_b = -1;
    int pending_qb = -1;  // thread 0: dQ store whose staging buffer has not been released yet
    uint32_t g = 0;
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int bh = t / NQ, qt = t - bh * NQ, b = bh / a.A, h = bh - b * a.A;
      const int q = qt * 128 + r;
      const int qb = it & 1;
      if (KB && b != prev_b) {
        named_bar_sync(1, DQ_ROW);  // nobody still reads the previous batch row's bias
        for (int base = 0; base < NC * 64; base += DQ_ROW) {
          const int key = base + tid;
          const bool ok = key < S && (a.key_mask == nullptr || a.key_mask[static_cast<size_t>(b) * S + key] != 0);
          if (key < NC * 64) kbias[key] = ok ? 0.0f : -INFINITY;
          const uint32_t w = __ballot_sync(0xffffffffu, ok);
          if ((tid & 31) == 0 && key < NC * 64) kvalid[key >> 5] = w;
        }
        named_bar_sync(1, DQ_ROW);
        prev_b = b;
      }
      const bool q_ok = q < S;
      const size_t stat = (static_cast<size_t>(b) * a.A + h) * S + (q_ok ? q : 0);
      // ds = p * (keep ? dp : 0) * c1 - p * D * scale with p = exp2(s * sl2 - lse)
      const float nlse = q_ok ? -a.lse2[stat] : -INFINITY;
      const float nds = q_ok ? -a.dsum[stat] * a.scale : 0.0f;
      const uint32_t* keep_row = a.keep + stat * keep_words + hq;
      auto fetch_keep = [&](int j) -> uint32_t {
        return (dropping && j < NC && 2 * j + hq < 4 * NQ) ? ldg_issue_u32(keep_row + 2 * j) : 0xffffffffu;
      };
      uint32_t kw_next = fetch_keep(0);
      for (int j = 0; j < NC; ++j, ++g) {
        const uint32_t bits = kw_next;
        kw_next = fetch_keep(j + 1);
        mbar_wait(s_full, g & 1u);
        tc_fence_after();
        uint32_t vs[32], vd[32];
        tmem_ld_32x32(lane_base + DQ_COL_S + hq * 32, vs);
        tmem_ld_32x32(lane_base + DQ_COL_DP + hq * 32, vd);
        tmem_wait_ld();
        tc_fence_before();
        mbar_arrive(s_free);
        const int key0 = j * 64 + hq * 32;
        uint32_t pk[16];
        const uint64_t sl2p = f2_splat(sl2), c1p = f2_splat(c1), ndsp = f2_splat(nds);
        auto chunk = [&](auto biased) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float off0 = decltype(biased)::value ? kbias[key0 + i] + nlse : nlse;
            const float off1 = decltype(biased)::value ? kbias[key0 + i + 1] + nlse : nlse;
            // packed fp32 pairs: one issue slot for the two elements' multiply-adds (same operations per element)
            float a0, a1, ds0, ds1;
            f2_unpack(f2_fma(f2_pack(__uint_as_float(vs[i]), __uint_as_float(vs[i + 1])), sl2p, f2_pack(off0, off1)), a0, a1);
            const float p0 = fast_exp2(a0), p1 = fast_exp2(a1);
            const float dp0 = ((bits >> i) & 1u) ? __uint_as_float(vd[i]) : 0.0f;
            const float dp1 = ((bits >> (i + 1)) & 1u) ? __uint_as_float(vd[i + 1]) : 0.0f;
            f2_unpack(f2_mul(f2_pack(p0, p1), f2_fma(f2_pack(dp0, dp1), c1p, ndsp)), ds0, ds1);
            pk[i >> 1] = pack_bf16x2(ds0, ds1);
          }
        };
        // 32-key groups without a masked key (all of them unless KB) take the path without the bias
        if (KB && kvalid[j * 2 + hq] != 0xffffffffu) chunk(std::true_type{}); else chunk(std::false_type{});
        // the dS columns are needed only now: the previous chunk's dQ MMA has had the whole computation to finish
        if (g > 0) mbar_wait(dq_done, (g - 1) & 1u);
        tc_fence_after();
        tmem_st_32x16(lane_base + DQ_COL_DS + hq * 16, pk);
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(ds_full);
        if (tid == 0 && pending_qb >= 0) {
          // the previous tile's dQ store has long finished reading its staging buffer: release it
          tma_wait_group_read<0>();
          mbar_arrive(&qdo_free[pending_qb]);
          pending_qb = -1;
        }
      }
      // ---- drain: dQ tile -> bf16 -> swizzled smem (this tile's Q buffer: every MMA of the tile has completed)
      // -> one TMA store into the Q slab of dqkv (rows beyond S are clipped)
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      uint8_t* stage = sQ + qb * CH;
      {
        uint32_t o[32];
        tmem_ld_32x32(lane_base + DQ_COL_DQ + hq * 32, o);
        tmem_wait_ld();
        tc_fence_before();
        mbar_arrive(dq_free);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint4 w;
          w.x = pack_bf16x2(__uint_as_float(o[8 * u]), __uint_as_float(o[8 * u + 1]));
          w.y = pack_bf16x2(__uint_as_float(o[8 * u + 2]), __uint_as_float(o[8 * u + 3]));
          w.z = pack_bf16x2(__uint_as_float(o[8 * u + 4]), __uint_as_float(o[8 * u + 5]));
          w.w = pack_bf16x2(__uint_as_float(o[8 * u + 6]), __uint_as_float(o[8 * u + 7]));
          *reinterpret_cast<uint4*>(stage + r * 128 + (((hq * 4 + u) ^ (r & 7)) << 4)) = w;
        }
      }
      fence_proxy_async_smem();
      named_bar_sync(2, DQ_ROW);
      if (tid == 0) {
        tma_store_3d(&tm_dqkv, stage, h * 64, qt * 128, b);
        tma_commit_group();
        pending_qb = qb;
      }
      if (a.dbias != nullptr) {
        // gradient of the query bias: column sums of the staged tile (rows beyond S are exact zeros),
        // 4 row groups of 32 -> shared memory -> 64 atomics per tile
        s_red[(tid >> 6) * 64 + (tid & 63)] = staged_colsum(stage, (tid >> 6) * 32, 32, tid & 63);
        named_bar_sync(2, DQ_ROW);
        if (tid < 64) {
          float sum = 0.0f;
#pragma unroll
          for (int g4 = 0; g4 < 4; ++g4) sum += s_red[g4 * 64 + tid];
          atomicAdd(a.dbias + h * 64 + tid, sum);
        }
      }
    }
    if (tid == 0) {
      if (pending_qb >= 0) {
        tma_wait_group_read<0>();
        mbar_arrive(&qdo_free[pending_qb]);
      }
      tma_wait_group<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == DQ_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
  if (tid == 0) ATRACE(9);
}

// =========================================================================================
// backward, dK and dV (rows = keys)
// =========================================================================================
// The transposed problem (rows of the score tile are keys).  One CTA per SM with the whole tensor memory: 16 row
// warps (thread (r, qd) = key row r of a 128-key tile, quarter qd of a chunk's query columns), two MMA-issuing
// warps (score MMAs / gradient MMAs, on two scheduler partitions), one TMA warp; contiguous (batch, head, key-tile)
// ranges per CTA.  The queries stream through in chunks of 128 (Q / dO tiles in a 4-stage ring); K / V tiles are
// double-buffered across tiles and a finished tile's buffers stage its dK / dV output.  Per chunk g:
//   S^T = K Q_g^T, dP^T = V dO_g^T   tcgen05.mma into 128 + 128 fp32 columns; the row threads pull them into
//                                    registers and hand the columns back at once, so the score MMAs of chunk
//                                    g + 1 run while chunk g is still being computed
//   P^T, dS^T                        computed in registers, written as bf16 pairs into 64 + 64 further columns:
//                                    no shared-memory round trip, no proxy fence
//   dV += P^T dO_g, dK += dS^T Q_g   tcgen05.mma with the A operands read from TMEM
// TMEM: S 128, dP 128, P^T 64, dS^T 64, dV 64, dK 64 = 512 columns.  A tile's accumulators are drained lazily,
// after the next tile's first chunk (warps 0-7 take dV, warps 8-15 dK), so that nobody idles on the last MMAs.
constexpr int KV_ROW = 512;
constexpr int KV_THREADS = KV_ROW + 96;
constexpr int KV_MMA_WARP = 16, KV_MMA2_WARP = 17, KV_TMA_WARP = 18;
constexpr int KV_RING = 4;  // a stage lives from its load to the chunk's gradient MMAs: ~2 chunk periods + the load latency
constexpr uint32_t KV_COL_S = 0, KV_COL_DP = 128, KV_COL_PT = 256, KV_COL_DST = 320, KV_COL_DV = 384, KV_COL_DK = 448;

__global__ void __launch_bounds__(KV_THREADS, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_dctx,
                    const __grid_constant__ CUtensorMap tm_dqkv, const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1024(smem_raw);
  const int N = a.nchunk;  // key tiles per (batch, head) = query chunks per tile
  uint8_t* sK = smem;                 // 2 x CH; a finished tile's buffer stages its dK tile
  uint8_t* sV = sK + 2 * CH;          // 2 x CH; ... its dV tile
  uint8_t* sQ = sV + 2 * CH;          // KV_RING stages of 128 queries
  uint8_t* sdO = sQ + KV_RING * CH;   // KV_RING stages
  float* s_stat = reinterpret_cast<float*>(sdO + KV_RING * CH);  // [2 buffers][-lse2 | -D * scale][512]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_stat + 2 * 2 * 512);
  uint64_t* kv_full = bars + 0;    // [2]
  uint64_t* kv_free = bars + 2;    // [2] last score MMAs of the tile done + the eight output stores have read the buffers
  uint64_t* ld_full = bars + 4;    // [KV_RING]
  uint64_t* ld_free = bars + 8;    // [KV_RING] score MMAs + gradient MMAs of the chunk
  uint64_t* s_full = bars + 12;
  uint64_t* s_free = bars + 13;
  uint64_t* pt_full = bars + 14;
  uint64_t* pt_free = bars + 15;
  uint64_t* out_full = bars + 16;
  uint64_t* out_free = bars + 17;
  uint64_t* st_full = bars + 18;   // [2] per-query statistics of a (batch, head) staged by the TMA warp
  uint64_t* st_free = bars + 20;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 22);

  const int tid = threadIdx.x;
  const int warp = uniform_warp_idx();
  const int S = a.S;
  const int tiles = a.B * a.A * N;
  const int t0 = static_cast<int>(static_cast<long long>(blockIdx.x) * tiles / gridDim.x);
  const int t1 = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * tiles / gridDim.x);
  const uint32_t chunks = static_cast<uint32_t>(t1 - t0) * N;  // Q / dO chunks this CTA streams
  if (tid == 0) ATRACE(0);

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_free[i], 9);
      mbar_init(&st_full[i], 1);
      mbar_init(&st_free[i], KV_ROW);
    }
    for (int i = 0; i < KV_RING; ++i) {
      mbar_init(&ld_full[i], 1);
      mbar_init(&ld_free[i], 2);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, KV_ROW);
    mbar_init(pt_full, KV_ROW);
    mbar_init(pt_free, 1);
    mbar_init(out_full, 1);
    mbar_init(out_free, KV_ROW);
    fence_mbar_init();
  }
  if (warp == KV_MMA_WARP) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  if (tid == KV_TMA_WARP * 32) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_dctx);
    tma_prefetch_desc(&tm_dqkv);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();     // launched with programmatic serialization: barrier set-up and TMEM allocation above overlapped the
                  // previous kernel's tail; its results are complete from here on
  pdl_trigger();  // the next kernel may start its prologue; it waits for this grid before reading

  if (warp == KV_TMA_WARP) {
    // ---------------- TMA producer ----------------
    const bool leader = elect_one();
    const int lane = tid & 31;
    uint32_t g = 0;
    int nbh = 0;
    // per-query statistics of a (batch, head), staged for the row threads: -lse2 (p = exp2(s * sl2 - lse2)) and
    // -D * scale (ds = p * (dp * c1 - D * scale)); queries beyond S get p = 0.  Double-buffered: the row threads
    // never wait for each other at a (batch, head) boundary, only for this warp.  All 32 loads of a lane are in
    // flight together (one load latency per staging).
    auto stage_stats = [&](int bh_) {
      const int sb = nbh & 1;
      if (nbh >= 2) mbar_wait(&st_free[sb], ((nbh >> 1) - 1) & 1);
      float* dst = s_stat + sb * 1024;
      const size_t stat = static_cast<size_t>(bh_) * S;
      float vl[16], vd[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const int q = lane + 32 * u;
        vl[u] = q < S ? ldg_issue_f32(a.lse2 + stat + q) : INFINITY;
        vd[u] = q < S ? ldg_issue_f32(a.dsum + stat + q) : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        dst[lane + 32 * u] = -vl[u];
        dst[512 + lane + 32 * u] = -vd[u] * a.scale;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&st_full[sb]);
      ++nbh;
    };
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int bh = t / N, kt = t - bh * N, b = bh / a.A, h = bh - b * a.A;
      const int kvb = it & 1;
      if (it >= 2) mbar_wait(&kv_free[kvb], ((it >> 1) - 1) & 1);
      if (leader) {
        mbar_arrive_expect_tx(&kv_full[kvb], 2 * CH);
        tma_load_3d(sK + kvb * CH, &tm_qkv, a.H + h * 64, kt * 128, b, &kv_full[kvb]);
        tma_load_3d(sV + kvb * CH, &tm_qkv, 2 * a.H + h * 64, kt * 128, b, &kv_full[kvb]);
      }
      __syncwarp();
      if (it == 0) stage_stats(bh);
      for (int i = 0; i < N; ++i, ++g) {
        const uint32_t st = g % KV_RING;  // KV_RING is a power of two
        if (g >= KV_RING) mbar_wait(&ld_free[st], ((g / KV_RING) - 1u) & 1u);
        if (leader) {
          mbar_arrive_expect_tx(&ld_full[st], 2 * CH);
          tma_load_3d(sQ + st * CH, &tm_qkv, h * 64, i * 128, b, &ld_full[st]);
          tma_load_3d(sdO + st * CH, &tm_dctx, h * 64, i * 128, b, &ld_full[st]);
        }
        __syncwarp();
      }
      // the next (batch, head)'s statistics are staged one tile early, while the ring is full and this warp would
      // only wait: staging them at the tile boundary stalled the operand loads for several load latencies
      if (kt == N - 1 && t + 1 < t1) stage_stats(bh + 1);
    }
  } else if (warp == KV_MMA_WARP) {
    // ---------------- MMA issuer 1: score MMAs (S^T, dP^T of chunk g) ----------------
    // Two issuing warps on two scheduler partitions: the issue slots of a partition are shared with four row warps,
    // and a single warp doing all the waiting and issuing was the critical path of the kernel.  The two chains need
    // no mutual ordering beyond the barriers below.  Positions are tracked by counters (no divisions in the loop).
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
    const bool leader = elect_one();
    int it = 0, i = 0;
    uint32_t st = 0, round = 0;
    MTACC_DECL;
    for (uint32_t g = 0; g < chunks; ++g) {
      const int kvb = it & 1;
      if (i == 0) mbar_wait(&kv_full[kvb], (it >> 1) & 1);
      mbar_wait(&ld_full[st], round & 1u);
      MTACC(0);
      if (g >= 1) mbar_wait(s_free, (g - 1u) & 1u);  // the previous chunk's scores are in registers
      MTACC(1);
      tc_fence_after();
      if (leader) {
        const uint32_t k_addr = smem_u32(sK + kvb * CH), v_addr = smem_u32(sV + kvb * CH);
        const uint32_t q_addr = smem_u32(sQ + st * CH), do_addr = smem_u32(sdO + st * CH);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tmem + KV_COL_S, umma_desc_kmajor(k_addr + k * 32), umma_desc_kmajor(q_addr + k * 32), idesc_s,
                    k > 0 ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tmem + KV_COL_DP, umma_desc_kmajor(v_addr + k * 32), umma_desc_kmajor(do_addr + k * 32), idesc_s,
                    k > 0 ? 1u : 0u);
        umma_commit(s_full);
        umma_commit(&ld_free[st]);  // one of the two arrivals (the gradient MMAs read the stage too)
        if (i == N - 1) umma_commit(&kv_free[kvb]);
      }
      __syncwarp();
      MTACC(2);
      if (++st == KV_RING) {
        st = 0;
        ++round;
      }
      if (++i == N) {
        i = 0;
        ++it;
      }
    }
    MTACC_WRITE();
  } else if (warp == KV_MMA2_WARP) {
    // ---------------- MMA issuer 2: dV += P^T dO_g, dK += dS^T Q_g, A operands from TMEM ----------------
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);
    const bool leader = elect_one();
    int it = 0, i = 0;
    uint32_t st = 0, round = 0;
    for (uint32_t g = 0; g < chunks; ++g) {
      mbar_wait(pt_full, g & 1u);
      mbar_wait(&ld_full[st], round & 1u);  // long complete (the score MMAs read the same stage); orders this warp too
      if (i == 0 && it > 0) mbar_wait(out_free, (it - 1) & 1);  // the previous tile's accumulators have been drained
      tc_fence_after();
      if (leader) {
        const uint32_t q_addr = smem_u32(sQ + st * CH), do_addr = smem_u32(sdO + st * CH);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16_ts(tmem + KV_COL_DV, tmem + KV_COL_PT + k * 8, umma_desc_mnmajor(do_addr + k * 2048, CH), idesc_o,
                       (i > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16_ts(tmem + KV_COL_DK, tmem + KV_COL_DST + k * 8, umma_desc_mnmajor(q_addr + k * 2048, CH), idesc_o,
                       (i > 0 || k > 0) ? 1u : 0u);
        umma_commit(pt_free);
        umma_commit(&ld_free[st]);
        if (i == N - 1) umma_commit(out_full);
      }
      __syncwarp();
      if (++st == KV_RING) {
        st = 0;
        ++round;
      }
      if (++i == N) {
        i = 0;
        ++it;
      }
    }
  } else {
    // ---------------- row threads ----------------
    // thread (r, qd): key row r of the tile, queries [32 qd, 32 qd + 32) of every 128-query chunk
    const int r = tid & 127, qd = tid >> 7, ew = warp & 3;
    const uint32_t lane_base = tmem + (static_cast<uint32_t>(ew * 32) << 16);
    const float sl2 = a.scale_log2;
    const bool dropping = a.drop.thresh != 0u;
    const float c1 = a.drop.scale * a.scale;
    ATACC_DECL;
    // Inputs are fetched ahead of their use so that their global-memory latency hides behind the arithmetic:
    // the key's keep decisions over the chunk's queries (key-major layout, word 4 i + qd) one chunk ahead.
    auto fetch_keep = [&](uint32_t g) -> uint32_t {
      if (!dropping || g >= chunks) return 0xffffffffu;
      const int t = t0 + static_cast<int>(g / N), i = static_cast<int>(g % N);
      const int bh = t / N, kt = t - bh * N;
      const int key = kt * 128 + r;
      if (key >= S) return 0xffffffffu;
      return ldg_issue_u32(a.keep + (static_cast<size_t>(bh) * S + key) * (8 * N) + 4 * N + 4 * i + qd);
    };
    uint32_t keep_next = fetch_keep(0);
    int prev_bh = -1, nbh = 0;
    const float* s_lse = s_stat;
    const float* s_dsum = s_stat + 512;
    const int half = tid >> 8;                    // 0: drains dV, 1: drains dK
    const int hq = (tid >> 7) & 1;                // column half of the drained accumulator
    const bool storer = (tid & 255) == ew * 32;   // one storing thread per (half, lane quarter)
    int release_kvb = -1;  // storing threads: K / V buffer whose output store has not been waited for yet
    // Storing threads: hand a staged K / V buffer back to the producer once its output store has read it.  Called
    // before every wait that could depend on the producer (the store itself depends on nothing), one chunk or more
    // after the store was issued, so that it practically never blocks.
    auto release_staged = [&]() {
      if (storer && release_kvb >= 0) {
        tma_wait_group_read<0>();
        mbar_arrive(&kv_free[release_kvb]);
        release_kvb = -1;
      }
    };
    // ---- drain of tile it_d: accumulator -> bf16 -> swizzled smem (the tile's V / K buffer: its score MMAs have
    // completed) -> TMA stores, per (half, TMEM lane quarter): 2 warps, named barrier, box of 32 rows.  Rows of
    // masked keys hold garbage (their probabilities are not forced to zero in the loop): they get exact zeros.
    // P^T was accumulated without the dropout scale: dV takes it here.
    auto drain = [&](int it_d) {
      const int t = t0 + it_d;
      const int bh = t / N, kt = t - bh * N, b = bh / a.A, h = bh - b * a.A;
      const int kvb = it_d & 1;
      const int key = kt * 128 + r;
      const bool key_ok = key < S && (a.key_mask == nullptr || a.key_mask[static_cast<size_t>(b) * S + key] != 0);
      uint8_t* stage = (half == 0 ? sV : sK) + kvb * CH;
      const float mul = half == 0 ? a.drop.scale : 1.0f;
      release_staged();
      mbar_wait(out_full, it_d & 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32(lane_base + (half == 0 ? KV_COL_DV : KV_COL_DK) + hq * 32, v);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(out_free);
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = key_ok ? __uint_as_float(v[q4 * 8 + e]) * mul : 0.0f;
        st_tile16(stage, r, hq * 4 + q4, f);
      }
      fence_proxy_async_smem();
      const uint32_t bar_id = 2 + half * 4 + ew;  // the two warps of (half, lane quarter)
      named_bar_sync(bar_id, 64);
      // V block is the third H-wide slab of dqkv, K the second; rows beyond S are clipped by the map
      const int slab = half == 0 ? 2 * a.H : a.H;
      if (storer) {
        tma_store_3d(&tm_dqkv, stage + ew * 4096, slab + h * 64, kt * 128 + ew * 32, b);
        tma_commit_group();
        release_kvb = kvb;  // the buffer goes back to the producer once the store has read it (checked lazily)
      }
      if (a.dbias != nullptr) {
        // gradients of the value / key biases: column sums of the quarter's 32 staged rows (rows of masked or
        // out-of-range keys are exact zeros)
        const int col = (r & 31) + 32 * hq;
        atomicAdd(a.dbias + slab + h * 64 + col, staged_colsum(stage, ew * 32, 32, col));
        named_bar_sync(bar_id, 64);  // the quarter's column sums have been read before the buffer can be released
      }
      ATACC(6);
    };
    int pending = -1;  // tile whose accumulators have not been drained yet
    uint32_t g = 0;
    for (int t = t0, it = 0; t < t1; ++t, ++it) {
      const int bh = t / N;
      if (bh != prev_bh) {
        if (prev_bh >= 0) mbar_arrive(&st_free[(nbh - 1) & 1]);  // done with the previous (batch, head)'s statistics
        release_staged();  // (the producer stages the statistics after it has got this tile's K / V buffers back)
        mbar_wait(&st_full[nbh & 1], (nbh >> 1) & 1);
        s_lse = s_stat + (nbh & 1) * 1024;
        s_dsum = s_lse + 512;
        prev_bh = bh;
        ++nbh;
      }
      ATACC(7);
      for (int i = 0; i < N; ++i, ++g) {
        const uint32_t bits = keep_next;
        keep_next = fetch_keep(g + 1);
        release_staged();
        mbar_wait(s_full, g & 1u);
        ATACC(1);
        tc_fence_after();
        uint32_t vs[32], vd[32];
        tmem_ld_32x32(lane_base + KV_COL_S + qd * 32, vs);
        tmem_ld_32x32(lane_base + KV_COL_DP + qd * 32, vd);
        tmem_wait_ld();
        tc_fence_before();
        mbar_arrive(s_free);
        ATACC(2);
        const int q0 = i * 128 + qd * 32;
        uint32_t ppk[16], dpk[16];
        const uint64_t sl2p = f2_splat(sl2), c1p = f2_splat(c1);
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          const float4 l0 = *reinterpret_cast<const float4*>(s_lse + q0 + gq * 8);
          const float4 l1 = *reinterpret_cast<const float4*>(s_lse + q0 + gq * 8 + 4);
          const float4 d0 = *reinterpret_cast<const float4*>(s_dsum + q0 + gq * 8);
          const float4 d1 = *reinterpret_cast<const float4*>(s_dsum + q0 + gq * 8 + 4);
          const float nlse[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
          const float nds[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const int c = gq * 8 + e;
            const bool keep0 = (bits >> c) & 1u, keep1 = (bits >> (c + 1)) & 1u;
            // packed fp32 pairs: one issue slot for the two elements' multiply-adds (same operations per element)
            float a0, a1, ds0, ds1;
            f2_unpack(f2_fma(f2_pack(__uint_as_float(vs[c]), __uint_as_float(vs[c + 1])), sl2p, f2_pack(nlse[e], nlse[e + 1])), a0, a1);
            const float p0 = fast_exp2(a0), p1 = fast_exp2(a1);
            const float dp0 = keep0 ? __uint_as_float(vd[c]) : 0.0f;
            const float dp1 = keep1 ? __uint_as_float(vd[c + 1]) : 0.0f;
            ppk[c >> 1] = pack_bf16x2(keep0 ? p0 : 0.0f, keep1 ? p1 : 0.0f);
            f2_unpack(f2_mul(f2_pack(p0, p1), f2_fma(f2_pack(dp0, dp1), c1p, f2_pack(nds[e], nds[e + 1]))), ds0, ds1);
            dpk[c >> 1] = pack_bf16x2(ds0, ds1);
          }
        }
        ATACC(3);
        // the operand columns are needed only now: the previous chunk's gradient MMAs have had the whole
        // computation to finish
        if (g >= 1) mbar_wait(pt_free, (g - 1u) & 1u);
        ATACC(4);
        tc_fence_after();
        tmem_st_32x16(lane_base + KV_COL_PT + qd * 16, ppk);
        tmem_st_32x16(lane_base + KV_COL_DST + qd * 16, dpk);
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(pt_full);
        ATACC(5);
        // The previous tile is drained AFTER this tile's first chunk: by then its last gradient MMAs have certainly
        // completed, so nobody idles on them.
        if (pending >= 0) {
          drain(pending);
          pending = -1;
        }
      }
      pending = it;
    }
    if (pending >= 0) drain(pending);
    release_staged();
    if (tid == 0) ATACC_WRITE();
    if ((tid & 255) < 128 && (tid & 31) == 0) tma_wait_group<0>();  // the eight storing threads
  }

  tc_fence_before();
  __syncthreads();
  if (warp == KV_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
  if (tid == 0) ATRACE(9);
}

int fwd_smem_bytes() { return 2 * CH + 2 * FW_RING * KCB + 512 + 2 * 3 * FW_SOFT * 4 + 1024; }
int dq_smem_bytes() { return 4 * CH + (DQ_KRING + DQ_VRING) * HC + 2048 + 2048 + 512 + 1024; }
int dkv_smem_bytes() { return 4 * CH + 2 * KV_RING * CH + 8192 + 256 + 1024; }

AttnArgs make_args(const AttnPlan& p, const uint8_t* key_mask, DropKey drop) {
  AttnArgs a;
  a.B = p.shape.B;
  a.S = p.shape.S;
  a.A = p.shape.A;
  a.H = p.shape.H;
  a.nchunk = p.nchunk;
  a.nchunk64 = (p.shape.S + 63) / 64;
  a.nchunk96 = (p.shape.S + KC - 1) / KC;
  a.tmem_cols = p.tmem_cols;
  a.scale = 0.125f;  // 1 / sqrt(64)
  a.scale_log2 = 0.125f * 1.4426950408889634f;
  a.key_mask = key_mask;
  a.ctx = p.ctx;
  a.lse2 = p.lse2;
  a.dsum = p.dsum;
  a.dqkv = p.dqkv;
  a.keep = p.keep;
  a.dbias = nullptr;
  a.drop = drop;
  return a;
}

}  // namespace

int attn_plan(AttnPlan* plan, const AttnShape& s, const void* qkv, void* ctx, float* lse2, const void* dctx,
              float* dsum, void* dqkv, uint32_t* keep) {
  if (s.H != s.A * 64) {
    set_error("attention: head dim must be 64 (H=%d, A=%d)", s.H, s.A);
    return 1;
  }
  if (s.S < 1 || s.S > 512) {
    set_error("attention: sequence length %d outside [1, 512]", s.S);
    return 1;
  }
  plan->shape = s;
  plan->nchunk = (s.S + 127) / 128;
  // forward: the scores plus, when they leave room, 64 columns for the output accumulator
  const int cols = plan->nchunk * 128 > 448 ? 512 : plan->nchunk * 128 + 64;
  plan->tmem_cols = cols <= 128 ? 128 : (cols <= 256 ? 256 : 512);
  plan->qkv = reinterpret_cast<const bf16*>(qkv);
  plan->ctx = reinterpret_cast<bf16*>(ctx);
  plan->lse2 = lse2;
  plan->dctx = reinterpret_cast<const bf16*>(dctx);
  plan->dsum = dsum;
  plan->dqkv = reinterpret_cast<bf16*>(dqkv);
  plan->keep = keep;
  const uint64_t H = s.H;
  if (qkv != nullptr &&
      encode_tmap_3d(&plan->tm_qkv, qkv, 2, 3 * H, s.S, s.B, 3 * H, static_cast<uint64_t>(s.S) * 3 * H, 64, 128, 1))
    return 1;
  if (dctx != nullptr &&
      (encode_tmap_3d(&plan->tm_dctx, dctx, 2, H, s.S, s.B, H, static_cast<uint64_t>(s.S) * H, 64, 128, 1) ||
       encode_tmap_3d(&plan->tm_do64, dctx, 2, H, s.S, s.B, H, static_cast<uint64_t>(s.S) * H, 64, 64, 1) ||
       encode_tmap_3d(&plan->tm_q64, qkv, 2, 3 * H, s.S, s.B, 3 * H, static_cast<uint64_t>(s.S) * 3 * H, 64, 64, 1)))
    return 1;
  // the dK/dV kernel stores 32-row boxes (one per TMEM lane quarter); the dQ kernel 128-row ones
  if (dqkv != nullptr &&
      encode_tmap_3d(&plan->tm_dqkv32, dqkv, 2, 3 * H, s.S, s.B, 3 * H, static_cast<uint64_t>(s.S) * 3 * H, 64, 32, 1))
    return 1;
  if (qkv != nullptr &&
      encode_tmap_3d(&plan->tm_kv96, qkv, 2, 3 * H, s.S, s.B, 3 * H, static_cast<uint64_t>(s.S) * 3 * H, 64, KC, 1))
    return 1;
  if (ctx != nullptr &&
      encode_tmap_3d(&plan->tm_ctx, ctx, 2, H, s.S, s.B, H, static_cast<uint64_t>(s.S) * H, 64, 128, 1))
    return 1;
  if (dqkv != nullptr &&
      encode_tmap_3d(&plan->tm_dqkv, dqkv, 2, 3 * H, s.S, s.B, 3 * H, static_cast<uint64_t>(s.S) * 3 * H, 64, 128, 1))
    return 1;
  return 0;
}

int attn_make_keep_bits(const AttnPlan& plan, DropKey drop, cudaStream_t st) {
  if (drop.thresh == 0u) return 0;
  if (plan.keep == nullptr) {
    set_error("attention: dropout needs the keep-bit buffer");
    return 1;
  }
  const int BH = plan.shape.B * plan.shape.A;
  const int W = 4 * plan.nchunk;
  const long long items = static_cast<long long>(BH) * ((plan.shape.S + 31) / 32) * W;  // one warp each
  long long grid = (items + 7) / 8;
  if (grid > 148 * 16) grid = 148 * 16;
  attn_keep_bits_kernel<<<static_cast<int>(grid), 256, 0, st>>>(plan.keep, BH, plan.shape.S, W, drop);
  return check_launch("attn_keep_bits");
}

int attn_forward(const AttnPlan& plan, const uint8_t* key_mask, DropKey drop, cudaStream_t st, bool bits_ready) {
  static PerDeviceOnce once;
  int once_dev;
  if (once.needs_setup(&once_dev)) {
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        fwd_smem_bytes()));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        fwd_smem_bytes()));
    // two CTAs per SM need the whole shared-memory carve-out
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    once.mark_done(once_dev);
  }
  const AttnArgs a = make_args(plan, key_mask, drop);
  if (drop.thresh != 0u && !bits_ready && attn_make_keep_bits(plan, drop, st)) return 1;
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  const int tiles = plan.nchunk * plan.shape.A * plan.shape.B;
  const int grid = tiles < 2 * sms ? tiles : 2 * sms;  // persistent, two co-resident CTAs per SM
  const bool kb = key_mask != nullptr || (plan.shape.S % KC) != 0;
  if (kb)
    launch_pdl(attn_fwd_kernel<true>, grid, FW_THREADS, fwd_smem_bytes(), st, plan.tm_qkv, plan.tm_kv96, plan.tm_ctx, a);
  else
    launch_pdl(attn_fwd_kernel<false>, grid, FW_THREADS, fwd_smem_bytes(), st, plan.tm_qkv, plan.tm_kv96, plan.tm_ctx, a);
  return check_launch("attn_fwd");
}

int attn_backward(const AttnPlan& plan, const uint8_t* key_mask, DropKey drop, cudaStream_t st, bool dsum_ready,
                  float* dqkv_bias_grad) {
  static PerDeviceOnce once;
  int once_dev; if (once.needs_setup(&once_dev)) {
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dq_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        dq_smem_bytes()));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dq_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        dq_smem_bytes()));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dq_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dq_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        dkv_smem_bytes()));
    BQA_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    once.mark_done(once_dev);
  }
  AttnArgs a = make_args(plan, key_mask, drop);
  a.dbias = dqkv_bias_grad;
  if (drop.thresh != 0u && plan.keep == nullptr) {
    set_error("attention: dropout needs the keep-bit buffer written by the forward");
    return 1;
  }
  const int T = plan.shape.B * plan.shape.S;
  if (!dsum_ready) {
    attn_dsum_kernel<<<(T + 3) / 4, 128, 0, st>>>(plan.ctx, plan.dctx, plan.dsum, plan.shape.B, plan.shape.S,
                                                 plan.shape.A, plan.shape.H);
    if (check_launch("attn_dsum")) return 1;
  }
  int sms = 0;
  if (device_sm_count(&sms)) return 1;
  const int tiles = plan.nchunk * plan.shape.A * plan.shape.B;
  const int grid = tiles < 2 * sms ? tiles : 2 * sms;  // persistent, two co-resident CTAs per SM
  const bool kb = key_mask != nullptr || (plan.shape.S % 64) != 0;
  if (kb)
    launch_pdl(attn_bwd_dq_kernel<true>, grid, DQ_THREADS, dq_smem_bytes(), st, plan.tm_qkv, plan.tm_q64, plan.tm_dctx,
               plan.tm_dqkv, a);
  else
    launch_pdl(attn_bwd_dq_kernel<false>, grid, DQ_THREADS, dq_smem_bytes(), st, plan.tm_qkv, plan.tm_q64, plan.tm_dctx,
               plan.tm_dqkv, a);
  if (check_launch("attn_bwd_dq")) return 1;
  const int grid_kv = tiles < sms ? tiles : sms;  // one CTA per SM (it owns the whole tensor memory)
  launch_pdl(attn_bwd_dkv_kernel, grid_kv, KV_THREADS, dkv_smem_bytes(), st, plan.tm_qkv, plan.tm_dctx, plan.tm_dqkv32, a);
  return check_launch("attn_bwd_dkv");
}

}  // namespace bqa

#ifdef BQA_ATTN_TRACE
extern "C" __attribute__((visibility("default"))) int bqa_trace_read(unsigned long long* host_out) {
  return cudaMemcpyFromSymbol(host_out, bqa::g_attn_trace, sizeof(unsigned long long) * 4096 * 16) != cudaSuccess;
}
#endif
