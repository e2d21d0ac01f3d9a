// gemm_sm100.cu — warp-specialised, dynamically scheduled bf16 GEMM on tcgen05 tensor cores (sm_100a).
//
// Replaces the cuBLAS GEMMs the reference reaches through torch.nn.Linear in
// transformers' BertSelfAttention / BertSelfOutput / BertIntermediate / BertOutput
// (site-packages/transformers/models/bert/modeling_bert.py:179-181,294-298,339-342,352-356)
// and their autograd backward (reference trainer.py:197-204).
//
// CTA = 384 threads, one CTA per SM.  Large problems run as CTA pairs (cluster of 2, tcgen05
// cta_group::2, 256 x 256 tiles, 6-stage ring; 5 stages with the two-output epilogue), small ones
// as single CTAs (128 x 256 tiles, 4-stage ring).  One cluster is launched per tile; a cluster that
// runs asks the hardware to cancel a pending cluster and takes over its tile (cluster launch
// control), so the grid behaves like a persistent one whose tile list is dynamic.
//   warp 0             TMA producer: ring of {A 128x64, B 128x64 (pair) / 256x64} bf16 tiles, 128B swizzle
//   warp 1             MMA issuer  : tcgen05.mma kind::f16, N=256 K=16, fp32 accumulators in TMEM
//                      (warps 0/1 walk their loops warp-uniformly; one elected lane issues)
//   warp 2             TMEM allocator (512 columns = two 256-column accumulators, double-buffered)
//   warp 3             tile scheduler: clusterlaunchcontrol.try_cancel, look-ahead of one tile
//   warps 4-11         epilogue    : tcgen05.ld -> registers -> fused op -> own swizzled smem staging ->
//                      own TMA store / reduce-add (two warps per TMEM lane quarter, one column half each)
// The prologue (barrier init, TMEM allocation, descriptor prefetch) overlaps the previous kernel's
// tail through programmatic dependent launch.
// Operands may be K-major or MN-major (instruction-descriptor major bits + smem descriptor
// strides), so forward, dgrad and wgrad all read the tensors in the layout PyTorch gives them.
#include "gemm_sm100.cuh"

#include "bqa_host.cuh"

#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace bqa {

namespace {

constexpr int BLOCK_N = 256;
constexpr int BLOCK_K = 64;
constexpr int TMEM_COLS = 512;
constexpr int NUM_THREADS = 384;
constexpr int EPI_THREADS = 256;
constexpr int W_STAGE_BYTES = 32 * 128;  // one epilogue warp's staging buffer: 32 rows x 128 B (64 bf16 or 32 fp32)
constexpr int SMEM_BARRIER_BYTES = 320;
constexpr int QD = 2;  // tile-scheduler response ring depth

// NCTA = 1: one CTA computes a 128 x 256 tile.  NCTA = 2: a CTA pair (cluster of 2, cta_group::2)
// computes a 256 x 256 tile; each CTA stages its own 128 rows of A and 128 of the 256 rows of B,
// so operand traffic per SM drops by a third and the ring can be deeper.
template <int NCTA, int EPI>
struct Cfg {
  static constexpr int BLOCK_M = 128 * NCTA;             // rows of the (pair) tile
  static constexpr int B_ROWS = BLOCK_N / NCTA;          // B rows staged by one CTA
  static constexpr int A_STAGE_BYTES = 128 * BLOCK_K * 2;
  static constexpr int B_STAGE_BYTES = B_ROWS * BLOCK_K * 2;
  // per-warp epilogue staging: two 32 x 128 B buffers for the bf16 epilogues (the two outputs of the GELU epilogue;
  // otherwise the two 64-column blocks of the warp's slab, so that the side operand of both blocks can be fetched
  // by TMA while the main loop runs and a block's store never waits for the other's), one for the fp32 reduce-add
  static constexpr int WARP_STAGE_BYTES = W_STAGE_BYTES * (EPI == EPI_F32_ADD ? 1 : 2);
  static constexpr int C_BYTES = 8 * WARP_STAGE_BYTES;
  // TMA -> MMA ring depth: whatever the staging buffers leave of the 227 KiB
  static constexpr int STAGES = (NCTA == 2 ? 6 : 4) - (EPI == EPI_F32_ADD ? 0 : 1);
  static constexpr int SMEM_BYTES = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + C_BYTES + SMEM_BARRIER_BYTES;
  static constexpr int SMEM_ALLOC = SMEM_BYTES + 1024;   // slack for manual 1024-byte alignment
  static_assert(SMEM_ALLOC <= 232448, "exceeds the 227 KiB per-CTA shared memory limit");
};

template <int A_MN, int B_MN, int EPI, int NCTA>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                          const CUtensorMap& tmC2, const GemmArgs& args) {
  using C = Cfg<NCTA, EPI>;
  constexpr int STAGES = C::STAGES;
  constexpr int A_STAGE_BYTES = C::A_STAGE_BYTES, B_STAGE_BYTES = C::B_STAGE_BYTES, BLOCK_M = C::BLOCK_M;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + STAGES * A_STAGE_BYTES;
  uint8_t* sC = sB + STAGES * B_STAGE_BYTES;
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(sC + C::C_BYTES);
  uint64_t* bar_empty = bar_full + STAGES;
  uint64_t* bar_tfull = bar_empty + STAGES;
  uint64_t* bar_tempty = bar_tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_tempty + 2);
  // dynamic tile scheduling (cluster launch control): response ring + barriers
  uint64_t* bar_qfull = bar_tempty + 4;   // [QD] response landed (16 bytes), armed by each CTA's scheduler warp
  uint64_t* bar_qempty = bar_qfull + QD;  // [QD] on the even CTA: every consumer warp of the cluster has read it
  uint64_t* bar_qgo = bar_qempty + QD;    // on the even CTA: its producer has started another tile
  uint4* clc_resp = reinterpret_cast<uint4*>((reinterpret_cast<uintptr_t>(bar_qgo + 1) + 15) & ~uintptr_t(15));  // [QD]
  uint64_t* bar_side = reinterpret_cast<uint64_t*>(clc_resp + QD);  // [8] side operand of an epilogue warp has landed

  const int warp = uniform_warp_idx();  // warp-uniform value: role branches below are uniform control flow
  const int lane = threadIdx.x & 31;
  const uint32_t rank = NCTA == 2 ? cluster_ctarank() : 0u;  // 0 = the CTA that issues the MMAs

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    if (EPI == EPI_GELU_FWD) tma_prefetch_desc(&tmC2);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&bar_full[i], NCTA);   // pair: one arrival per CTA's producer, on the even CTA's barrier
      mbar_init(&bar_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_tfull[i], 1);
      mbar_init(&bar_tempty[i], NCTA * EPI_THREADS);  // pair: both CTAs' epilogues arrive on the even CTA
    }
    for (int i = 0; i < QD; ++i) {
      mbar_init(&bar_qfull[i], 1);
      mbar_init(&bar_qempty[i], NCTA * 9 + 1);  // producer + 8 epilogue warps per CTA, one MMA warp
    }
    mbar_init(bar_qgo, 1);
    for (int i = 0; i < 8; ++i) mbar_init(&bar_side[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (NCTA == 2) {
      tmem_alloc_pair(tmem_slot, TMEM_COLS);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_slot, TMEM_COLS);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // the prologue above overlapped the previous kernel's tail (programmatic dependent launch)
  pdl_wait();
  pdl_trigger();

  const int tiles_mn = args.m_tiles * args.n_tiles;
  // One cluster is launched per tile; the clusters that get to run take over the tiles of clusters that
  // have not started (cluster launch control), so the grid behaves like a persistent one with dynamic
  // scheduling: an SM that becomes free late (another kernel was running on it) simply joins in.
  // Every consumer warp walks the same sequence of responses: response k of the cluster is in slot k % QD.
  uint32_t qk = 0;  // responses consumed by this warp
  auto next_tile = [&](int& tile) -> bool {
    const uint32_t slot = qk % QD, ph = (qk / QD) & 1u;
    mbar_wait(&bar_qfull[slot], ph);
    uint32_t cta_x;
    const bool valid = clc_decode(&clc_resp[slot], cta_x);
    fence_proxy_async_smem();  // the slot is overwritten by the async proxy once released
    __syncwarp();
    if (lane == 0) {
      if (NCTA == 2 && rank != 0) mbar_arrive_cluster(&bar_qempty[slot], 0);
      else mbar_arrive(&bar_qempty[slot]);
    }
    ++qk;
    tile = static_cast<int>(cta_x) / NCTA;
    return valid;
  };
  const int tile0 = blockIdx.x / NCTA;

  if (warp == 0) {
    // ------------------------------ TMA producer (every CTA) ------------------------------
    // the whole warp walks the loop (uniform control flow and operands); one elected lane issues
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = tile0;;) {
      // starting a tile lets the scheduler ask for the one after it (look-ahead of one tile: a cluster never
      // holds more than one tile it has not started, which keeps short grids balanced)
      if (rank == 0 && leader) mbar_arrive(bar_qgo);
      const int split = tile / tiles_mn;
      const int mn = tile - split * tiles_mn;
      const int m0 = (mn / args.n_tiles) * BLOCK_M + static_cast<int>(rank) * 128;      // own A rows
      const int n0 = (mn % args.n_tiles) * BLOCK_N + static_cast<int>(rank) * C::B_ROWS;  // own B rows
      const int kb0 = split * args.kb_per_split;
      const int kb1 = min(kb0 + args.kb_per_split, args.kb_total);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&bar_empty[stage], phase ^ 1u);
        if (leader) {
        const int k0 = kb * BLOCK_K;
        uint8_t* a_dst = sA + stage * A_STAGE_BYTES;
        uint8_t* b_dst = sB + stage * B_STAGE_BYTES;
        uint64_t* fb = &bar_full[stage];
        if (NCTA == 1) mbar_arrive_expect_tx(fb, A_STAGE_BYTES + B_STAGE_BYTES);
        auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1) {
          if (NCTA == 2) tma_load_2d_pair(dst, tm, c0, c1, fb); else tma_load_2d(dst, tm, c0, c1, fb);
        };
        if (A_MN) {
#pragma unroll
          for (int i = 0; i < 2; ++i) load(a_dst + i * (BLOCK_K * 128), &tmA, m0 + i * 64, k0);
        } else {
          load(a_dst, &tmA, k0, m0);
        }
        if (B_MN) {
#pragma unroll
          for (int i = 0; i < C::B_ROWS / 64; ++i) load(b_dst + i * (BLOCK_K * 128), &tmB, n0 + i * 64, k0);
        } else {
          load(b_dst, &tmB, k0, n0);
        }
        if (NCTA == 2) {
          // both CTAs' bytes are credited to the even CTA's barrier, which expects them all
          if (rank == 0) mbar_arrive_expect_tx(fb, 2 * (A_STAGE_BYTES + B_STAGE_BYTES));
          else mbar_arrive_cluster(fb, 0);
        }
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
      if (!next_tile(tile)) break;
    }
  } else if (warp == 1 && rank == 0) {
    // ------------------------------ MMA issuer (even CTA only) ------------------------------
    constexpr uint32_t idesc = umma_idesc_bf16(BLOCK_M, BLOCK_N, A_MN, B_MN);
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    uint32_t it = 0;
    for (int tile = tile0;; ++it) {
      const int split = tile / tiles_mn;
      const int kb0 = split * args.kb_per_split;
      const int kb1 = min(kb0 + args.kb_per_split, args.kb_total);
      const uint32_t as = it & 1u;
      const uint32_t aphase = (it >> 1) & 1u;
      mbar_wait(&bar_tempty[as], aphase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&bar_full[stage], phase);
        tc_fence_after();
        if (leader) {
        const uint32_t a_addr = smem_u32(sA + stage * A_STAGE_BYTES);
        const uint32_t b_addr = smem_u32(sB + stage * B_STAGE_BYTES);
#pragma unroll
        for (int k = 0; k < BLOCK_K / 16; ++k) {
          const uint64_t adesc = A_MN ? umma_desc_mnmajor(a_addr + k * 2048, BLOCK_K * 128)
                                      : umma_desc_kmajor(a_addr + k * 32);
          const uint64_t bdesc = B_MN ? umma_desc_mnmajor(b_addr + k * 2048, BLOCK_K * 128)
                                      : umma_desc_kmajor(b_addr + k * 32);
          if (NCTA == 2) umma_bf16_pair(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          else umma_bf16(d_tmem, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        }
        if (NCTA == 2) {
          umma_commit_pair(&bar_empty[stage]);  // frees the stage in both CTAs
          if (kb == kb1 - 1) umma_commit_pair(&bar_tfull[as]);
        } else {
          umma_commit(&bar_empty[stage]);
          if (kb == kb1 - 1) umma_commit(&bar_tfull[as]);
        }
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
      if (!next_tile(tile)) break;
    }
  } else if (warp == 3) {
    // ------------------------------ tile scheduler (every CTA) ------------------------------
    // Each CTA's scheduler warp arms its own response barrier; the even CTA issues the cancellation
    // request, whose response is multicast to the pair.  A slot is reused only after every consumer warp
    // of the cluster has released it (they arrive on the even CTA's barrier).
    const bool leader = elect_one();
    for (uint32_t k = 0;; ++k) {
      const uint32_t slot = k % QD, ph = (k / QD) & 1u;
      if (rank == 0) {
        mbar_wait(bar_qgo, k & 1u);
        if (k >= QD) mbar_wait(&bar_qempty[slot], ((k / QD) - 1u) & 1u);
      }
      if (leader) {
        mbar_arrive_expect_tx(&bar_qfull[slot], 16);
        if (rank == 0) clc_try_cancel(&clc_resp[slot], &bar_qfull[slot], NCTA == 2);
      }
      __syncwarp();
      mbar_wait(&bar_qfull[slot], ph);
      uint32_t cta_x;
      if (!clc_decode(&clc_resp[slot], cta_x)) break;  // no cluster left: the consumers see the same and stop
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue ------------------------------
    // 8 independent warps.  Warp w may read TMEM lanes 32*(w%4)..+31; the two warps of a lane quarter
    // take one 128-column half of the tile each.  A warp drains its 32 x 128 block in 32-column pieces
    // through its own staging buffer (32 rows x 128 B) and issues its own TMA stores, so there is no
    // CTA-wide barrier in the loop and the warps hide each other's TMEM / global / store latencies.
    const int ew = warp & 3;                  // TMEM lane quarter
    const int cg = (warp - 4) >> 2;           // column half handled by this warp (0 / 1)
    const int row = ew * 32 + lane;           // accumulator row owned by this thread
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
    uint8_t* stage0 = sC + (warp - 4) * C::WARP_STAGE_BYTES;
    uint8_t* stage1 = stage0 + W_STAGE_BYTES;  // second output of the GELU epilogue
    uint8_t* srow = stage0 + lane * 128;
    const int sw = lane & 7;                  // SWIZZLE_128B: 16-byte unit j of row r lives at unit j ^ (r & 7)
    uint32_t it = 0;
    for (int tile = tile0;; ++it) {
      const int split = tile / tiles_mn;
      const int mn = tile - split * tiles_mn;
      const int m0 = (mn / args.n_tiles) * BLOCK_M + static_cast<int>(rank) * 128;  // rows held in this CTA's TMEM
      const int n0 = (mn % args.n_tiles) * BLOCK_N + cg * 128;                       // this warp's first column
      const uint32_t as = it & 1u;
      const uint32_t aphase = (it >> 1) & 1u;
      const int m = m0 + row;
      const bool row_ok = m < args.M;
      // Operand of the fused elementwise op (residual, gelu', or the rows the output is dotted with): its 32 x 128
      // block comes by TMA into the warp's two staging buffers — the exact places the output blocks are written to
      // afterwards — while the main loop of this tile still runs.  (Per-thread global loads one piece ahead left
      // a load latency exposed in each of the four pieces.)  Rows / columns outside the operand arrive as zeros.
      const bool dotting = EPI == EPI_LINEAR && args.dot_out != nullptr;  // side operand = dot_with, not added
      const bool has_side = (EPI == EPI_LINEAR && (args.residual != nullptr || dotting)) || EPI == EPI_GELU_BWD;
      float dot_acc = 0.0f;
      if (EPI != EPI_F32_ADD && EPI != EPI_GELU_FWD) {
        if (lane == 0) {
          tma_wait_group_read<0>();  // the previous tile's stores have read both buffers
          if (has_side) {
            mbar_arrive_expect_tx(&bar_side[warp - 4], 2 * W_STAGE_BYTES);
            tma_load_2d(stage0, &tmC2, n0, m0 + ew * 32, &bar_side[warp - 4]);
            tma_load_2d(stage1, &tmC2, n0 + 64, m0 + ew * 32, &bar_side[warp - 4]);
          }
        }
        __syncwarp();
      }
      // the dropout decisions of the thread's 128 elements depend only on their indices: hashed while the main loop
      // still runs instead of on the critical path after it
      uint32_t kbits[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
      if (EPI == EPI_LINEAR && args.drop.thresh != 0u) {
        const uint32_t base_idx = static_cast<uint32_t>(m) * static_cast<uint32_t>(args.N) + static_cast<uint32_t>(n0);
#pragma unroll
        for (int piece = 0; piece < 4; ++piece) kbits[piece] = keep_bits32(args.drop, base_idx + piece * 32);  // even index
      }
      mbar_wait(&bar_tfull[as], aphase);
      tc_fence_after();
#pragma unroll
      for (int piece = 0; piece < 4; ++piece) {
        const int n = n0 + piece * 32;
        uint32_t v[32];
        tmem_ld_32x32(lane_base + as * BLOCK_N + cg * 128 + piece * 32, v);
        // staging row of this piece's 64-column block (the GELU epilogue has one buffer per output instead)
        uint8_t* brow = (EPI == EPI_GELU_FWD || EPI == EPI_F32_ADD) ? srow : srow + (piece >> 1) * W_STAGE_BYTES;
        uint4 side[4];
        if (has_side) {
          if (piece == 0) mbar_wait(&bar_side[warp - 4], it & 1u);
#pragma unroll
          for (int q = 0; q < 4; ++q) side[q] = *reinterpret_cast<const uint4*>(brow + ((((piece & 1) * 4 + q) ^ sw) << 4));
        }
        const bool side_ok = has_side;
        tmem_wait_ld();
        if (piece == 3) {
          // accumulator fully drained into registers: hand the TMEM stage back to the MMA warp
          tc_fence_before();
          if (NCTA == 2 && rank != 0) mbar_arrive_cluster(&bar_tempty[as], 0);
          else mbar_arrive(&bar_tempty[as]);
        }
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);

        if (EPI == EPI_F32_ADD) {
          if (lane == 0) tma_wait_group_read<0>();
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 8; ++q)
            *reinterpret_cast<float4*>(srow + ((q ^ sw) << 4)) = make_float4(f[4 * q], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (n < args.N) tma_reduce_add_2d(&tmC, stage0, n, m0 + ew * 32);
            tma_commit_group();
          }
        } else {
          // bias, dropout and residual on packed fp32 pairs (one issue slot per two elements; the same operations and
          // roundings per element as in scalar form; a dropped element is multiplied by zero, as torch's dropout does)
          if (EPI != EPI_GELU_BWD && args.bias != nullptr && n < args.N) {
            const float4* bp = reinterpret_cast<const float4*>(args.bias + n);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 b = __ldg(bp + q);
              f2_unpack(f2_add(f2_pack(f[4 * q], f[4 * q + 1]), f2_pack(b.x, b.y)), f[4 * q], f[4 * q + 1]);
              f2_unpack(f2_add(f2_pack(f[4 * q + 2], f[4 * q + 3]), f2_pack(b.z, b.w)), f[4 * q + 2], f[4 * q + 3]);
            }
          }
          if (EPI == EPI_LINEAR) {
            if (args.drop.thresh != 0u) {
              const uint32_t kb = kbits[piece];
              const float sc = args.drop.scale;
#pragma unroll
              for (int i = 0; i < 32; i += 2)
                f2_unpack(f2_mul(f2_pack(f[i], f[i + 1]),
                                 f2_pack(((kb >> i) & 1u) ? sc : 0.0f, ((kb >> (i + 1)) & 1u) ? sc : 0.0f)),
                          f[i], f[i + 1]);
            }
            if (side_ok && !dotting) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 r = side[q];
                const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 rv = unpack_bf16x2(rw[e]);
                  f2_unpack(f2_add(f2_pack(f[8 * q + 2 * e], f[8 * q + 2 * e + 1]), f2_pack(rv.x, rv.y)), f[8 * q + 2 * e],
                            f[8 * q + 2 * e + 1]);
                }
              }
            }
          }
          if (EPI == EPI_GELU_BWD) {
            // aux holds gelu'(pre-activation), written by the forward GEMM's epilogue
            if (side_ok) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 r = side[q];
                const float2 u0 = unpack_bf16x2(r.x), u1 = unpack_bf16x2(r.y);
                const float2 u2 = unpack_bf16x2(r.z), u3 = unpack_bf16x2(r.w);
                f[8 * q] *= u0.x;
                f[8 * q + 1] *= u0.y;
                f[8 * q + 2] *= u1.x;
                f[8 * q + 3] *= u1.y;
                f[8 * q + 4] *= u2.x;
                f[8 * q + 5] *= u2.y;
                f[8 * q + 6] *= u3.x;
                f[8 * q + 7] *= u3.y;
              }
            }
          }
          if (EPI == EPI_GELU_BWD && args.colsum_out != nullptr) {
            // Column sums of this 32 x 32 piece over the warp's rows by a transpose-reduce: each step halves
            // the values a lane holds and doubles the rows they cover (31 shuffles), lane c ends with column c.
            // Rows beyond M are exact zeros (zero-filled operand rows).
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = f[i];
#pragma unroll
            for (int s = 16; s >= 1; s >>= 1) {
              const bool up = (lane & s) != 0;
#pragma unroll
              for (int i = 0; i < s; ++i) {
                const float send = up ? v[i] : v[i + s];
                const float recv = __shfl_xor_sync(0xffffffffu, send, s);
                v[i] = (up ? v[i + s] : v[i]) + recv;
              }
            }
            if (n + lane < args.N) atomicAdd(args.colsum_out + n + lane, v[0]);
          }
          // two pieces fill one 128-byte staging row; the buffer is reused once the previous store has
          // finished reading it (the other epilogues alternate between two buffers, checked once per tile)
          if (EPI == EPI_GELU_FWD && (piece & 1) == 0) {
            if (lane == 0) tma_wait_group_read<0>();
            __syncwarp();
          }
          if (EPI == EPI_GELU_FWD) {
            // one evaluation of the shared core gives the activation (second output) and its
            // derivative (first output, consumed by the dgrad epilogue instead of recomputing it)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float gv[8];
#pragma unroll
              for (int e = 0; e < 8; e += 2) {
                float dg0, dg1;
                gelu_and_grad2(f[8 * q + e], f[8 * q + e + 1], gv[e], gv[e + 1], dg0, dg1);
                f[8 * q + e] = dg0;
                f[8 * q + e + 1] = dg1;
              }
              uint4 o;
              o.x = pack_bf16x2(gv[0], gv[1]);
              o.y = pack_bf16x2(gv[2], gv[3]);
              o.z = pack_bf16x2(gv[4], gv[5]);
              o.w = pack_bf16x2(gv[6], gv[7]);
              *reinterpret_cast<uint4*>(srow + W_STAGE_BYTES + ((((piece & 1) * 4 + q) ^ sw) << 4)) = o;
            }
          }
          // bf16 pack, swizzled like the TMA store expects (SWIZZLE_128B)
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 o;
            o.x = pack_bf16x2(f[8 * q], f[8 * q + 1]);
            o.y = pack_bf16x2(f[8 * q + 2], f[8 * q + 3]);
            o.z = pack_bf16x2(f[8 * q + 4], f[8 * q + 5]);
            o.w = pack_bf16x2(f[8 * q + 6], f[8 * q + 7]);
            *reinterpret_cast<uint4*>(brow + ((((piece & 1) * 4 + q) ^ sw) << 4)) = o;
            if (dotting && side_ok) {
              // dot product of the values as stored (bf16) with the side operand, 64 columns = two pieces
              const uint4 r = side[q];
              const float2 a0 = unpack_bf16x2(o.x), a1 = unpack_bf16x2(o.y), a2 = unpack_bf16x2(o.z), a3 = unpack_bf16x2(o.w);
              const float2 b0 = unpack_bf16x2(r.x), b1 = unpack_bf16x2(r.y), b2 = unpack_bf16x2(r.z), b3 = unpack_bf16x2(r.w);
              dot_acc += a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y + a2.x * b2.x + a2.y * b2.y +
                         a3.x * b3.x + a3.y * b3.y;
            }
          }
          if (dotting && (piece & 1)) {
            if (row_ok && n < args.N) {
              const int blk = (n - 32) >> 6;  // 64-column block of the output row
              const int outer = m / args.dot_rows, inner = m - outer * args.dot_rows;
              args.dot_out[(static_cast<size_t>(outer) * (args.N >> 6) + blk) * args.dot_rows + inner] = dot_acc;
            }
            dot_acc = 0.0f;
          }
          if (piece & 1) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              const int nc = n - 32;
              if (nc < args.N) {
                tma_store_2d(&tmC, EPI == EPI_GELU_FWD ? stage0 : stage0 + (piece >> 1) * W_STAGE_BYTES, nc, m0 + ew * 32);
                if (EPI == EPI_GELU_FWD) tma_store_2d(&tmC2, stage1, nc, m0 + ew * 32);
              }
              tma_commit_group();
            }
          }
        }
      }
      if (!next_tile(tile)) break;
    }
    if (lane == 0) tma_wait_group<0>();
  }

  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();  // pair: the peer may still be reading barriers / TMEM
  if (warp == 2) {
    tc_fence_after();
    if (NCTA == 2) tmem_dealloc_pair(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int A_MN, int B_MN, int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC2,
            const GemmArgs args) {
  gemm_body<A_MN, B_MN, EPI, 1>(tmA, tmB, tmC, tmC2, args);
}

template <int A_MN, int B_MN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC2,
                 const GemmArgs args) {
  gemm_body<A_MN, B_MN, EPI, 2>(tmA, tmB, tmC, tmC2, args);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

template <int A_MN, int B_MN, int EPI>
int launch_variant(const GemmPlan& plan, cudaStream_t stream) {
  static PerDeviceOnce once;  // the shared-memory attribute is per device
  auto k1 = gemm_kernel<A_MN, B_MN, EPI>;
  auto k2 = gemm_pair_kernel<A_MN, B_MN, EPI>;
  int once_dev; if (once.needs_setup(&once_dev)) {
    cudaError_t e = cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<1, EPI>::SMEM_ALLOC);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<2, EPI>::SMEM_ALLOC);
    if (e != cudaSuccess) {
      set_error("gemm: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
      return 1;
    }
    once.mark_done(once_dev);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(plan.grid);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = plan.pair ? Cfg<2, EPI>::SMEM_ALLOC : Cfg<1, EPI>::SMEM_ALLOC;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see pdl_wait() in the kernel
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = plan.pair ? cudaLaunchKernelEx(&cfg, k2, plan.tmA, plan.tmB, plan.tmC, plan.tmC2, plan.args)
                            : cudaLaunchKernelEx(&cfg, k1, plan.tmA, plan.tmB, plan.tmC, plan.tmC2, plan.args);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("gemm: launch failed: %s", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

}  // namespace

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_err; }

int encode_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t cols, uint64_t rows,
                   uint64_t ld_elems, uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0 || ((ld_elems * elem_bytes) & 15u) != 0) {
    set_error("tensor map: base/stride must be 16-byte aligned (base=%p ld=%llu)", base,
              (unsigned long long)ld_elems);
    return 1;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * static_cast<uint64_t>(elem_bytes)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(2d) failed: CUresult %d (cols=%llu rows=%llu ld=%llu box=%ux%u)", (int)r,
              (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld_elems, box_cols,
              box_rows);
    return 1;
  }
  return 0;
}

int encode_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t d0, uint64_t d1, uint64_t d2,
                   uint64_t stride1_elems, uint64_t stride2_elems, uint32_t b0, uint32_t b1, uint32_t b2) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
    return 1;
  }
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {stride1_elems * static_cast<uint64_t>(elem_bytes),
                           stride2_elems * static_cast<uint64_t>(elem_bytes)};
  cuuint32_t box[3] = {b0, b1, b2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(3d) failed: CUresult %d", (int)r);
    return 1;
  }
  return 0;
}

int gemm_plan(GemmPlan* plan, const GemmProblem& p, int num_sms) {
  if (p.M <= 0 || p.N <= 0 || p.K <= 0) {
    set_error("gemm: empty problem %dx%dx%d", p.M, p.N, p.K);
    return 1;
  }
  if ((p.N % 64) != 0) {
    set_error("gemm: N=%d must be a multiple of 64", p.N);
    return 1;
  }
  if (p.k_splits > 1 && p.epilogue != EPI_F32_ADD) {
    set_error("gemm: split-K needs the fp32 reduce-add epilogue");
    return 1;
  }
  const int a_mn = p.a_mn_major ? 1 : 0, b_mn = p.b_mn_major ? 1 : 0;
  // CTA pairs (256-row tiles) whenever there are at least two 128-row tiles of M
  const bool pair = p.M > 128 && !p.force_single;
  plan->pair = pair ? 1 : 0;
  const int b_rows = pair ? BLOCK_N / 2 : BLOCK_N;
  int rc = 0;
  if (a_mn)
    rc |= encode_tmap_2d(&plan->tmA, p.A, 2, p.M, p.K, p.lda, 64, BLOCK_K);
  else
    rc |= encode_tmap_2d(&plan->tmA, p.A, 2, p.K, p.M, p.lda, BLOCK_K, 128);
  if (b_mn)
    rc |= encode_tmap_2d(&plan->tmB, p.B, 2, p.N, p.K, p.ldb, 64, BLOCK_K);
  else
    rc |= encode_tmap_2d(&plan->tmB, p.B, 2, p.K, p.N, p.ldb, BLOCK_K, b_rows);
  if (p.epilogue == EPI_F32_ADD)
    rc |= encode_tmap_2d(&plan->tmC, p.C, 4, p.N, p.M, p.ldc, 32, 32);
  else
    rc |= encode_tmap_2d(&plan->tmC, p.C, 2, p.N, p.M, p.ldc, 64, 32);
  if (p.epilogue == EPI_GELU_FWD) {
    if (p.C2 == nullptr) {
      set_error("gemm: EPI_GELU_FWD needs a second output");
      return 1;
    }
    rc |= encode_tmap_2d(&plan->tmC2, p.C2, 2, p.N, p.M, p.ldc, 64, 32);
  } else {
    // the other bf16 epilogues fetch their side operand (residual / rows to dot with / gelu') through this map: same
    // [M][N] shape and 64 x 32 boxes as the output
    const void* side = nullptr;
    int ld_side = 0;
    if (p.epilogue == EPI_LINEAR && p.residual != nullptr) side = p.residual, ld_side = p.ld_res;
    if (p.epilogue == EPI_LINEAR && p.dot_out != nullptr) side = p.dot_with, ld_side = p.ld_dot;
    if (p.epilogue == EPI_GELU_BWD) side = p.aux, ld_side = p.ld_aux;
    if (side != nullptr)
      rc |= encode_tmap_2d(&plan->tmC2, side, 2, p.N, p.M, ld_side, 64, 32);
    else
      plan->tmC2 = plan->tmC;
  }
  if (rc) return 1;

  GemmArgs& a = plan->args;
  a.M = p.M;
  a.N = p.N;
  a.K = p.K;
  const int block_m = pair ? 256 : 128;
  const int units = pair ? num_sms / 2 : num_sms;  // concurrently running tiles
  a.m_tiles = (p.M + block_m - 1) / block_m;
  a.n_tiles = (p.N + BLOCK_N - 1) / BLOCK_N;
  a.kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  int splits = p.k_splits < 1 ? 1 : p.k_splits;
  if (p.k_splits == 0 && p.epilogue == EPI_F32_ADD) {
    // automatic split-K: fill whole waves of `units` tiles, keep >= 8 k-blocks per split
    const int tiles = a.m_tiles * a.n_tiles;
    double best = -1.0;
    for (int sp = 1; sp <= 32 && a.kb_total / sp >= 8; ++sp) {
      const int total = tiles * sp;
      const int waves = (total + units - 1) / units;
      const double eff = static_cast<double>(total) / (static_cast<double>(waves) * units);
      if (eff > best + 0.02) {
        best = eff;
        splits = sp;
      }
    }
  }
  if (splits > a.kb_total) splits = a.kb_total;
  a.kb_per_split = (a.kb_total + splits - 1) / splits;
  a.k_splits = (a.kb_total + a.kb_per_split - 1) / a.kb_per_split;
  a.bias = p.bias;
  a.residual = reinterpret_cast<const bf16*>(p.residual);
  a.ld_res = p.ld_res;
  a.aux = reinterpret_cast<const bf16*>(p.aux);
  a.ld_aux = p.ld_aux;
  a.drop = p.drop;
  a.dot_with = reinterpret_cast<const bf16*>(p.dot_with);
  a.ld_dot = p.ld_dot;
  a.dot_out = p.dot_out;
  a.dot_rows = p.dot_rows;
  a.colsum_out = p.colsum_out;
  if (p.colsum_out != nullptr && p.epilogue != EPI_GELU_BWD) {
    set_error("gemm: the column-sum output is implemented for EPI_GELU_BWD only");
    return 1;
  }
  if (p.dot_out != nullptr &&
      (p.epilogue != EPI_LINEAR || p.residual != nullptr || p.dot_with == nullptr || p.dot_rows < 1 || p.N % 64 != 0)) {
    set_error("gemm: the row-dot output needs EPI_LINEAR without residual, a dot_with operand and N %% 64 == 0");
    return 1;
  }
  if (p.epilogue == EPI_GELU_BWD && p.aux == nullptr) {
    set_error("gemm: EPI_GELU_BWD needs aux");
    return 1;
  }
  plan->variant = a_mn | (b_mn << 1) | (p.epilogue << 2);
  const long tiles = static_cast<long>(a.m_tiles) * a.n_tiles * a.k_splits;
  (void)units;
  plan->grid = static_cast<int>(tiles) * (pair ? 2 : 1);  // one cluster per tile; running clusters take over the rest
  return 0;
}

int gemm_launch(const GemmPlan& plan, cudaStream_t stream) {
  switch (plan.variant) {
    case 0 | (0 << 1) | (EPI_LINEAR << 2):
      return launch_variant<0, 0, EPI_LINEAR>(plan, stream);
    case 0 | (0 << 1) | (EPI_GELU_FWD << 2):
      return launch_variant<0, 0, EPI_GELU_FWD>(plan, stream);
    case 0 | (1 << 1) | (EPI_LINEAR << 2):
      return launch_variant<0, 1, EPI_LINEAR>(plan, stream);
    case 0 | (1 << 1) | (EPI_GELU_BWD << 2):
      return launch_variant<0, 1, EPI_GELU_BWD>(plan, stream);
    case 1 | (1 << 1) | (EPI_F32_ADD << 2):
      return launch_variant<1, 1, EPI_F32_ADD>(plan, stream);
    case 0 | (0 << 1) | (EPI_F32_ADD << 2):
      return launch_variant<0, 0, EPI_F32_ADD>(plan, stream);
    default:
      set_error("gemm: unsupported variant a_mn=%d b_mn=%d epilogue=%d", plan.variant & 1, (plan.variant >> 1) & 1,
                plan.variant >> 2);
      return 1;
  }
}

}  // namespace bqa
