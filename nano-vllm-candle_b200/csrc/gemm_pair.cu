// K3 (M > 128): C[M,N] = A[M,K] * W[N,K]^T on CTA PAIRS (tcgen05.mma.cta_group::2).
// Replaces the same reference calls as gemm_tcgen05.cu (src/layers/linear.rs:33-38,72-77,
// 99-104,171-175,184-198; src/models/qwen3.rs:542-550) for prefill and for decode batches > 128.
//
// Output tile = 256 x 256 per CTA pair (two SMs of one TPC).  CTA h of the pair (h = cluster
// rank & 1) loads rows [128h, 128h+128) of the A tile AND rows [128h, 128h+128) of the W tile
// (its half of the N dimension); one tcgen05.mma.cta_group::2 (M=256, N=256, K=16) issued by the
// leader (h = 0) reads both halves of W and writes 128 accumulator rows x 256 columns into each
// CTA's tensor memory.  Per SM and k-block that is 32 KB of shared-memory ingest for 512 tensor
// cycles (64 B/clk) against 48 KB (96 B/clk) for the 1-CTA 128 x 256 tile.
//
// Two schedules share the kernel:
//  * S = 1, persistent: min(tiles, #pairs) clusters of 2 walk the tiles (n fastest); two
//    256-column accumulators in tensor memory (all 512 columns) so the epilogue of tile i
//    overlaps the main loop of tile i+1.
//  * S > 1, split-K inside a cluster of 2S CTAs (decode, M <= 256: too few tiles to fill 148
//    SMs): pair p of the cluster takes k-blocks [p*kb_per, ...) of ONE tile; partial accumulators
//    are exchanged through an L2-resident workspace between two cluster barriers (every CTA
//    publishes the 16-column chunks it does not own; CTA (p, h) sums chunk c, c % S == p, in
//    split order 0..S-1: deterministic, no atomics).  Distributed shared memory was measured
//    first and is too slow for 128 KB partial tiles (~17 B/clk per SM).
// Roles per CTA (192 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer (leader
// only), warps 2-5 epilogue (tcgen05.ld -> bf16/f32 -> 16-byte stores).
#include "common.cuh"
#include "ptx.cuh"
#include "internal.h"
#include "gemm_epilogue.cuh"

#include <cmath>
#include <cstdlib>

extern int g_nvc_debug[8];

namespace {

constexpr int kTileM = 256, kHalf = 128, BK = 64;
constexpr int kThreads = 192;
constexpr int kBytesA = kHalf * BK * 2;            // 16 KB: this CTA's 128 rows of the activation tile
constexpr int kTmemCols = 512;
constexpr int kRingBudget = 200 * 1024;

// Tile width NT.  NT = 256: one MMA (N = 256) per k-step, two accumulator stages (persistent
// schedule overlaps the epilogue).  NT in [128, 256): one narrower MMA per k-step, two accumulator stages - more,
// smaller tiles when 256-wide ones quantise badly onto the 74 pairs (96 tiles = two waves of which the second is 30 %
// full; 128 tiles of 192 = two full waves of 3/4 the work each).  (Wide tiles, NT in (256, 512] as two MMAs per k-step
// into one accumulator stage, measured no gain - 8B gate_up at batch 256: 53.1 vs 53.6 us - and were deleted in round 2.)
template <int NT>
struct PairCfg {
  static_assert(NT == 64 || (NT >= 128 && NT <= 256 && NT % 32 == 0), "tile width");
  static constexpr int NH = NT;                          // N of one tcgen05.mma
  static constexpr int kBoxRowsB = NH / 2;               // rows of W one CTA supplies per MMA (multiple of 8)
  static constexpr int kBytesB = (NT / 2) * BK * 2;      // this CTA's rows of the weight tile per k-block
  static constexpr int kStageBytes = kBytesA + kBytesB;
  // NT = 64 is the swapped-operand lm_head (NVC_EPI_TRANSPOSE): 20 KB stages, a deeper ring keeps as many bytes in flight
  static constexpr int kMaxStages = NT == 64 ? 10 : 6;
  static constexpr int kStages = kRingBudget / kStageBytes < kMaxStages ? kRingBudget / kStageBytes : kMaxStages;
  static constexpr int kAccStages = 2;
  static constexpr int kNumBars = 2 * kStages + 4;       // full[], empty[], tmem_full[2], tmem_empty[2]
  static constexpr int kOffStage = (kStages * kStageBytes + kNumBars * 8 + 16 + 127) & ~127;   // 4 epilogue warps x row staging
  // 4 KB per warp (32 rows x 128 B, RowStage); 8 KB for 256-wide tiles, whose QK-norm epilogue parks whole heads (32 rows x 256 B)
  static constexpr int kStageBuf = NT == 256 ? 8192 : 4096;
  static constexpr int kSmemBytes = kOffStage + 4 * kStageBuf + 1024;
  static_assert(kSmemBytes <= 232448, "shared memory per CTA");
  static constexpr int kChunks = NT / 16;                // 16-column accumulator chunks of a tile
  static constexpr int kPartF4 = kChunks * 4 * kHalf;    // float4s of one CTA's partial tile in the split-K workspace
  static_assert(kStageBytes % 1024 == 0 && (kBoxRowsB % 8) == 0, "swizzle atoms");
  // tensor-memory column of accumulator chunk c (accumulator stage 0)
  __device__ static constexpr uint32_t chunk_col(int c) {
    return (uint32_t)(c * 16);
  }
};

// ---- PTX used only by the pair kernel (cta_group::2 / cluster forms)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// this CTA's box -> its own shared memory; bytes complete on the LEADER's mbarrier (cluster address)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const CUtensorMap* m, uint32_t leader_bar,
                                                 int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once all tcgen05 ops issued so far by this thread have retired) on the mbarrier at this
// CTA-relative offset in every CTA of `cta_mask`
__device__ __forceinline__ void umma_commit_pair_mcast(uint32_t bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"(cta_mask) : "memory");
}
// Split-K exchange, executed by the 4 epilogue warps of every CTA of a cluster of 2S CTAs between
// two cluster barriers.  Partial accumulators go through an L2-resident workspace, not distributed
// shared memory (DSMEM moves ~17 B/clk per SM: measured 7.4k cycles for a 128 x 256 fp32 tile).
// Workspace layout per CTA: [chunk][float4 4][row 128], so a warp's 32 rows are 512 contiguous bytes.
// Ownership is per 32-column unit (two chunks: what the fused SiLU-mul epilogue needs in one thread):
// CTA (p, h) owns the units u with u % S == p and sums them in split order 0..S-1 (deterministic).
template <int NT, int S>
__device__ __forceinline__ void pair_publish(float4* __restrict__ my_ws, uint32_t taddr, int p, int r) {
  using cfg = PairCfg<NT>;
#pragma unroll 1
  for (int u = 0; u < cfg::kChunks / 2; ++u) {
    if (u % S == p) continue;           // my units stay in tensor memory
    uint32_t v[32];
    ptx::tmem_ld_32x32b_x16(taddr + cfg::chunk_col(2 * u), v);
    ptx::tmem_ld_32x32b_x16(taddr + cfg::chunk_col(2 * u + 1), v + 16);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j)
      __stcg(my_ws + (u * 8 + j) * kHalf + r, make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                                         __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])));
  }
}

template <int NT, int S>
__device__ __forceinline__ void pair_reduce(const float4* __restrict__ ws_cluster, uint32_t taddr, int p, int h, int r,
                                            void* __restrict__ C, int out_f32, int epi, int64_t ldc, int row0, int M,
                                            int n_base, int N, uint8_t* stage_buf) {
  using cfg = PairCfg<NT>;
  constexpr int kUnits = cfg::kChunks / 2;
  // G units per batch: 16-24 independent 16-byte L2 loads in flight per thread (one round trip per batch)
  constexpr int G = S == 2 ? 3 : 1;
  constexpr int kOwnedMax = (kUnits + S - 1) / S;
#pragma unroll 1
  for (int k0 = 0; k0 < kOwnedMax; k0 += G) {
    float4 t[G][S - 1][8];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int u = min(p + (k0 + g) * S, kUnits - 1);       // clamped: a redundant load instead of a branch
#pragma unroll
      for (int i = 0; i < S - 1; ++i) {
        const int z = i + (i >= p ? 1 : 0);                  // the i-th split that is not mine
        const float4* peer = ws_cluster + (size_t)(2 * z + h) * cfg::kPartF4 + u * 8 * kHalf + r;
#pragma unroll
        for (int j = 0; j < 8; ++j) t[g][i][j] = __ldcg(peer + j * kHalf);
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int u = p + (k0 + g) * S;
      if (u < kUnits) {                   // uniform over the CTA
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x16(taddr + cfg::chunk_col(2 * u), v);
        ptx::tmem_ld_32x32b_x16(taddr + cfg::chunk_col(2 * u + 1), v + 16);
        ptx::tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = 0.f;
        // split order 0..S-1 whoever reduces: peers i < p are splits 0..p-1, then mine, then splits p+1..
#pragma unroll
        for (int i = 0; i < S; ++i) {
          if (i == p) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] += __uint_as_float(v[j]);
          }
          if (i < S - 1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              f[4 * j] += t[g][i][j].x; f[4 * j + 1] += t[g][i][j].y;
              f[4 * j + 2] += t[g][i][j].z; f[4 * j + 3] += t[g][i][j].w;
            }
          }
        }
        // direct stores: a unit is only 32-128 bytes per row, staging it measured slower (16.0 vs 14.5 us on 8B o_proj)
        const int grow = row0 + (r & 31);
        if (grow < M) emit32(C, out_f32, epi, (int64_t)grow * ldc, n_base + u * 32, N, f);
      }
    }
  }
}

template <int NT>
__global__ void __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 void* __restrict__ C, int ldc, int out_f32, int M, int N, int k_blocks_total, int kb_per_split,
                 int num_tiles, int n_tiles, int prefetch_w, int epi, float4* __restrict__ ws,
                 long long* __restrict__ trace, const QkEpi qk) {
  using cfg = PairCfg<NT>;
  constexpr int kStages = cfg::kStages;
  constexpr int kStageBytes = cfg::kStageBytes;
  // optional phase timestamps of CTA 0 (tools/gemm_microbench.py); NULL in production
  const bool tr = trace != nullptr && blockIdx.x == 0;
  pdl_trigger();
  if (tr && threadIdx.x == 0) trace[0] = clock64();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + kStages * kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tmem_full_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };
  auto tmem_empty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * cfg::kNumBars;
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + kStages * kStageBytes + cfg::kNumBars * 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();
  const uint32_t csize = ptx::cluster_nctarank();
  const int S = (int)(csize >> 1);              // k-splits sharing one output tile
  const int p = (int)(crank >> 1);              // this pair's split
  const int h = (int)(crank & 1u);              // half of the tile rows (A) / of each MMA's columns (W)
  const uint32_t leader_rank = crank & ~1u;
  const uint16_t pair_mask = (uint16_t)(3u << leader_rank);
  const int cluster_id = blockIdx.x / csize;
  const int num_clusters = gridDim.x / csize;
  const int kb_begin = p * kb_per_split;
  const int num_kb = min(kb_per_split, k_blocks_total - kb_begin);

  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStages; ++s) {
        ptx::mbar_init(full_bar(s), 1);     // the leader's arrive.expect_tx; bytes of BOTH CTAs complete on the leader's barrier
        ptx::mbar_init(empty_bar(s), 1);    // the leader's tcgen05.commit, multicast to both CTAs
      }
      for (int a = 0; a < 2; ++a) {
        ptx::mbar_init(tmem_full_bar(a), 1);    // leader's commit after the last k-block of a tile (multicast)
        ptx::mbar_init(tmem_empty_bar(a), 8);   // 4 epilogue warps x 2 CTAs, on the leader's barrier
      }
      ptx::fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, kTmemCols);
    tmem_relinquish_pair();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;
  ptx::cluster_sync_all();       // every CTA's barriers exist before a peer's TMA / commit / arrive targets them
  if (tr && threadIdx.x == 0) trace[1] = clock64();

  // weights are never written by a kernel: pull this CTA's first k-blocks of W into L2 while the
  // predecessor is still running (decode only; the persistent schedule streams anyway)
  constexpr int kPrefetchKb = 16;
  if (warp == 0 && prefetch_w && cluster_id < num_tiles) {
    if (ptx::elect_one()) {
      const int n_tile = cluster_id % n_tiles;
      const int n = min(num_kb, kPrefetchKb);
      for (int kb = 0; kb < n; ++kb) {
#pragma unroll
        for (int i = 0; i < 1; ++i) {
          const int row = n_tile * NT + i * cfg::NH + h * cfg::kBoxRowsB;
          ptx::tma_prefetch_l2_2d(&tma_b, (kb_begin + kb) * BK, row);
        }
      }
    }
    __syncwarp();
  }
  pdl_wait();
  if (tr && threadIdx.x == 0) trace[2] = clock64();

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    const uint32_t leader_full0 = ptx::mapa_u32(full_bar(0), leader_rank);
    uint32_t it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m_tile = tile / n_tiles, n_tile = tile - m_tile * n_tiles;
      const int a_row = m_tile * kTileM + h * kHalf;
      const int b_row = n_tile * NT + h * cfg::kBoxRowsB;
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % kStages;
        if (it >= kStages) ptx::mbar_wait(empty_bar(s), ((it / kStages) & 1) ^ 1u);   // first lap: slots are free
        if (ptx::elect_one()) {
          if (h == 0) ptx::mbar_arrive_expect_tx(full_bar(s), 2 * kStageBytes);
          const uint32_t dst = smem_base + s * kStageBytes;
          const int k0 = (kb_begin + kb) * BK;
          tma_load_2d_pair(dst, &tma_a, leader_full0 + 8u * s, k0, a_row);
#pragma unroll
          for (int i = 0; i < 1; ++i) {   // this CTA's rows of each MMA's weight tile
            const uint32_t bdst = dst + kBytesA + i * (cfg::kBoxRowsB * BK * 2);
            tma_load_2d_pair(bdst, &tma_b, leader_full0 + 8u * s, k0, b_row + i * cfg::NH);
          }
          if (prefetch_w && kb + kPrefetchKb < num_kb) {
#pragma unroll
            for (int i = 0; i < 1; ++i) {
              ptx::tma_prefetch_l2_2d(&tma_b, (kb_begin + kb + kPrefetchKb) * BK, b_row + i * cfg::NH);
            }
          }
          if (tr && it < 32) trace[3 + it] = clock64();
        }
        __syncwarp();
      }
    }
  } else if (warp == 1 && h == 0) {
    // ------------------------------------------------------------ MMA issuer (leader CTA of the pair)
    constexpr uint32_t idesc = ptx::umma_idesc_bf16(kTileM, cfg::NH, 0, 0);
    constexpr uint32_t desc_hi = ptx::umma_desc_hi_sw128(1024);
    const uint32_t a_lo0 = ptx::umma_desc_lo(smem_base, 16);
    uint32_t it = 0;
    int ti = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++ti) {
      const int acc = ti & 1, use = ti >> 1;
      ptx::mbar_wait(tmem_empty_bar(acc), (use & 1) ^ 1u);    // both CTAs' epilogues drained this accumulator
      ptx::tcgen05_fence_after();
      const uint32_t d = tmem_base + (uint32_t)(acc * 256);
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % kStages;
        ptx::mbar_wait(full_bar(s), (it / kStages) & 1);
        ptx::tcgen05_fence_after();
        if (ptx::elect_one()) {
          if (tr && it < 32) trace[40 + it] = clock64();
          const uint32_t a_lo = a_lo0 + s * (kStageBytes >> 4);
          const uint32_t b_lo = a_lo + (kBytesA >> 4);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
#pragma unroll
            for (int i = 0; i < 1; ++i)
              umma_bf16_ss_pair(d + (uint32_t)(i * 256), ptx::umma_desc(a_lo + 2 * k, desc_hi),
                                ptx::umma_desc(b_lo + i * ((cfg::kBoxRowsB * BK * 2) >> 4) + 2 * k, desc_hi), idesc,
                                (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_pair_mcast(empty_bar(s), pair_mask);                             // stage reusable in both CTAs
          if (kb == num_kb - 1) umma_commit_pair_mcast(tmem_full_bar(acc), pair_mask);  // accumulator complete
          if (tr && kb == num_kb - 1 && ti == 0) trace[80] = clock64();
        }
        __syncwarp();
      }
    }
  } else if (warp >= 2 && S == 1) {
    // ------------------------------------------------------------ epilogue, persistent schedule
    const int q = warp & 3;                 // TMEM lane quadrant this warp may access
    const int r = q * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t leader_tmem_empty0 = ptx::mapa_u32(tmem_empty_bar(0), leader_rank);
    int ti = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++ti) {
      const int m_tile = tile / n_tiles, n_tile = tile - m_tile * n_tiles;
      const int acc = ti & 1, use = ti >> 1;
      const int grow = m_tile * kTileM + h * kHalf + r;
      const bool row_ok = grow < M;
      ptx::mbar_wait(tmem_full_bar(acc), use & 1);
      ptx::tcgen05_fence_after();
      if (tr && threadIdx.x == 64 && ti == 0) trace[81] = clock64();
      if (tr && threadIdx.x == 64 && tile + num_clusters >= num_tiles) trace[86] = clock64();   // last tile: exposed epilogue
      if constexpr (NT == 256) {
        if (epi == NVC_EPI_QK_NORM_ROPE) {
          // ---- K2 as the epilogue of the QKV projection: two heads of 128 columns per tile
          const int row_base = m_tile * kTileM + h * kHalf + q * 32;            // first row of this warp
          const uint32_t stage = smem_u32(smem_gen + cfg::kOffStage + q * cfg::kStageBuf);   // [32 rows][16 chunks of 16 B], chunk ^= row & 7
          const int l = lane & 7, sub = lane >> 3;
          const unsigned mask8 = 0xffu << (lane & 24);
          const int my_row = row_base + lane;
          const int pos_lane = my_row < M ? __ldg(qk.positions + my_row) : 0;
          const int slot_lane = (my_row < M && qk.slot_mapping != nullptr) ? __ldg(qk.slot_mapping + my_row) : -1;
          nvc_bf16* Cb = reinterpret_cast<nvc_bf16*>(C);
#pragma unroll 1
          for (int hh = 0; hh < 2; ++hh) {
            const int head = n_tile * 2 + hh;                                   // head index in the fused [q | k | v] row
            if (head * 128 >= N) break;
            // park the head's 128 accumulator columns as bf16 (what the plain epilogue would have stored)
#pragma unroll 1
            for (int c4 = 0; c4 < 4; ++c4) {
              uint32_t v[32];
              ptx::tmem_ld_32x32b_x16(taddr + (uint32_t)(acc * 256 + hh * 128 + c4 * 32), v);
              ptx::tmem_ld_32x32b_x16(taddr + (uint32_t)(acc * 256 + hh * 128 + c4 * 32 + 16), v + 16);
              ptx::tmem_ld_wait();
              float f[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint4 pk = pack8(f + 8 * k);
                const int chunk = c4 * 4 + k;
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(stage + lane * 256 + ((chunk ^ (lane & 7)) << 4)),
                             "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w) : "memory");
              }
            }
            __syncwarp();
            const bool is_q = head < qk.nh, is_k = !is_q && head < qk.nh + qk.kv;
            const uint4 w_lo = is_q ? __ldg(qk.qw + l) : (is_k ? __ldg(qk.kw + l) : make_uint4(0, 0, 0, 0));
            const uint4 w_hi = is_q ? __ldg(qk.qw + l + 8) : (is_k ? __ldg(qk.kw + l + 8) : make_uint4(0, 0, 0, 0));
#pragma unroll 4
            for (int pass = 0; pass < 8; ++pass) {
              const int rr = pass * 4 + sub;                                    // row inside the warp's 32: 8 lanes per row
              const int grow = row_base + rr;
              const bool ok = grow < M;
              const int pos = __shfl_sync(0xffffffffu, pos_lane, rr);
              const int slot = __shfl_sync(0xffffffffu, slot_lane, rr);
              uint4 ra, rb;
              asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(ra.x), "=r"(ra.y), "=r"(ra.z), "=r"(ra.w)
                           : "r"(stage + rr * 256 + ((l ^ (rr & 7)) << 4)) : "memory");
              asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(rb.x), "=r"(rb.y), "=r"(rb.z), "=r"(rb.w)
                           : "r"(stage + rr * 256 + (((l + 8) ^ (rr & 7)) << 4)) : "memory");
              uint4 o1 = ra, o2 = rb;
              if (is_q || is_k) {
                const float4* cs = qk.cos_sin + (int64_t)pos * 32;
                const float4 c0 = __ldg(cs + 2 * l), c1 = __ldg(cs + 2 * l + 1);
                const float4 s0 = __ldg(cs + 16 + 2 * l), s1 = __ldg(cs + 16 + 2 * l + 1);
                if (!ok) ra = rb = make_uint4(0, 0, 0, 0);
                qk_norm_rope8(ra, rb, mask8, true, w_lo, w_hi, qk.eps, c0, c1, s0, s1, o1, o2);
              }
              if (ok) {
                uint4* dst = reinterpret_cast<uint4*>(Cb + (int64_t)grow * ldc + head * 128);
                dst[l] = o1;
                dst[l + 8] = o2;
                if (!is_q && slot >= 0) {
                  const int kvh = is_k ? head - qk.nh : head - qk.nh - qk.kv;
                  uint4* cdst = (is_k ? qk.k_cache : qk.v_cache) + ((int64_t)slot * qk.kv + kvh) * 16;
                  cdst[l] = o1;
                  cdst[l + 8] = o2;
                }
              }
            }
            __syncwarp();                       // the staging buffer is reused by the next head
          }
          ptx::tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(leader_tmem_empty0 + 8u * acc);
          continue;
        }
      }
      if (epi == NVC_EPI_TRANSPOSE) {
        // accumulator row `grow` (a weight row) is column `grow` of the fp32 C; the warp's 32 rows are 128 contiguous bytes of
        // every C row (token) they touch
        float* Ct = reinterpret_cast<float*>(C) + grow;
#pragma unroll 1
        for (int c = 0; c < cfg::kChunks; c += 2) {
          uint32_t v[32];
          ptx::tmem_ld_32x32b_x16(taddr + (uint32_t)(acc * 256) + cfg::chunk_col(c), v);
          ptx::tmem_ld_32x32b_x16(taddr + (uint32_t)(acc * 256) + cfg::chunk_col(c + 1), v + 16);
          ptx::tmem_ld_wait();
          const int col0 = n_tile * NT + c * 16;
          if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < N) Ct[(int64_t)(col0 + j) * ldc] = __uint_as_float(v[j]);
          }
        }
        ptx::tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(leader_tmem_empty0 + 8u * acc);
        continue;
      }
      RowStage st = make_row_stage(smem_gen + cfg::kOffStage + q * cfg::kStageBuf, C, out_f32, epi, ldc,
                                   m_tile * kTileM + h * kHalf + q * 32, M, n_tile * NT, NT, N);
#pragma unroll 1
      for (int c = 0; c < cfg::kChunks; c += 2) {       // two 16-column chunks per tensor-memory round trip
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x16(taddr + (uint32_t)(acc * 256) + cfg::chunk_col(c), v);
        ptx::tmem_ld_32x32b_x16(taddr + (uint32_t)(acc * 256) + cfg::chunk_col(c + 1), v + 16);
        ptx::tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        st.emit32(lane, out_f32, epi, f);
      }
      st.finish(lane);
      ptx::tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(leader_tmem_empty0 + 8u * acc);
      if (tr && threadIdx.x == 64 && ti == 0) trace[83] = clock64();
      if (tr && threadIdx.x == 64 && tile + num_clusters >= num_tiles) trace[87] = clock64();
    }
  }

  if (S > 1) {
    // ------------------------------------------------------------ split-K exchange (one tile per cluster)
    const int tile = cluster_id;
    const int m_tile = tile / n_tiles, n_tile = tile - m_tile * n_tiles;
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    float4* ws_cluster = ws + (size_t)cluster_id * csize * cfg::kPartF4;
    if (warp >= 2) {
      ptx::mbar_wait(tmem_full_bar(0), 0);  // all MMAs of this pair have retired
      ptx::tcgen05_fence_after();
      if (tr && threadIdx.x == 64) trace[81] = clock64();
      float4* my_ws = ws_cluster + (size_t)crank * cfg::kPartF4;
      if (S == 2) pair_publish<NT, 2>(my_ws, taddr, p, r);
      else if (S == 3) pair_publish<NT, 3>(my_ws, taddr, p, r);
      else pair_publish<NT, 4>(my_ws, taddr, p, r);
    }
    if (tr && threadIdx.x == 64) trace[82] = clock64();
    ptx::cluster_sync_all();                // every partial is published (release / acquire at cluster scope)
    if (tr && threadIdx.x == 64) trace[85] = clock64();
    if (warp >= 2) {
      const int row0 = m_tile * kTileM + h * kHalf + q * 32;          // first row of this warp
      uint8_t* stage_buf = smem_gen + cfg::kOffStage + q * cfg::kStageBuf;
      if (S == 2) pair_reduce<NT, 2>(ws_cluster, taddr, p, h, r, C, out_f32, epi, ldc, row0, M, n_tile * NT, N, stage_buf);
      else if (S == 3) pair_reduce<NT, 3>(ws_cluster, taddr, p, h, r, C, out_f32, epi, ldc, row0, M, n_tile * NT, N, stage_buf);
      else pair_reduce<NT, 4>(ws_cluster, taddr, p, h, r, C, out_f32, epi, ldc, row0, M, n_tile * NT, N, stage_buf);
      if (tr && threadIdx.x == 64) trace[83] = clock64();
    }
  }
  ptx::tcgen05_fence_before();
  ptx::cluster_sync_all();       // nobody exits (or frees tensor memory) while a peer may still signal its barriers
  if (warp == 1) {
    ptx::tcgen05_fence_after();
    tmem_dealloc_pair(tmem_base, kTmemCols);
  }
  if (tr && threadIdx.x == 0) trace[84] = clock64();
}

int make_tma_2d(nvc_ctx* ctx, CUtensorMap* map, const void* base, uint64_t inner, uint64_t rows,
                uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_rows) {
  cuuint64_t gdim[2] = {inner, rows};
  cuuint64_t gstr[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    nvc_set_error(ctx, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return NVC_EINVAL;
  }
  return NVC_OK;
}

#define NVC_PAIR_WIDTHS(X) X(64) X(128) X(160) X(192) X(224) X(256)

template <int NT>
int launch_pair(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_f32, int M, int N,
                int K, int S, int epi, cudaStream_t stream, long long* trace, const QkEpi& qk = QkEpi{}, int ldw = 0) {
  using cfg = PairCfg<NT>;
  if (ldw == 0) ldw = K;
  CUtensorMap ta, tb;
  int rc = make_tma_2d(ctx, &ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK, kHalf);
  if (rc) return rc;
  rc = make_tma_2d(ctx, &tb, W, (uint64_t)K, (uint64_t)N, (uint64_t)ldw * 2, BK, cfg::kBoxRowsB);
  if (rc) return rc;
  const int kb_total = K / BK;
  const int m_tiles = (M + kTileM - 1) / kTileM, n_tiles = (N + NT - 1) / NT;
  const int tiles = m_tiles * n_tiles;
  if (S > 1 && (size_t)tiles * 2 * S * cfg::kPartF4 * sizeof(float4) > ctx->gemm_ws_bytes) S = 1;   // workspace too small
  int kb_per = (kb_total + S - 1) / S;
  S = (kb_total + kb_per - 1) / kb_per;      // no empty splits
  int clusters = tiles;
  if (S == 1) {
    const int avail = ctx->pair_clusters[1] > 0 ? ctx->pair_clusters[1] : ctx->num_sms / 2;
    // equal tiles per cluster where possible: ceil(tiles / waves) clusters, not a ragged last wave
    const int waves = (tiles + avail - 1) / avail;
    clusters = (tiles + waves - 1) / waves;
  }
  dim3 grid(clusters * 2 * S);
  // L2 prefetch of the first weight k-blocks before the dependency wait + 16 k-blocks ahead of the ring: pays for short,
  // latency-bound k-loops (decode at K <= 3072: 1 - 3 %), costs 4 - 9 % on the long ones (Qwen3-8B at batch 256, K >= 4096:
  // qkv 18.4 -> 16.8 us, o_proj 14.6 -> 13.9, gate_up 49.6 -> 47.4, down 30.3 -> 28.4 without it) - the prefetches compete
  // with the ring's own TMA loads for the same L2 / HBM queues (profiles/r02_pair_prefetch_ab.txt)
  const int prefetch_w = (M <= kTileM && kb_total <= 48) ? 1 : 0;
  NVC_CUDA_CHECK(ctx, nvc_launch_cluster(ctx, gemm_pair_kernel<NT>, grid, dim3(kThreads), cfg::kSmemBytes, stream, 2 * S,
                                         ta, tb, C, ldc, out_f32, M, N, kb_total, kb_per, tiles, n_tiles,
                                         prefetch_w, epi, reinterpret_cast<float4*>(ctx->gemm_ws), trace,
                                         qk));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

}  // namespace

// Co-resident clusters of 2S CTAs (S = 1..4), measured once per context: the schedule model needs
// them (cluster placement is per GPC: B200 holds 74 / 33 / 22 / 15 clusters of 2 / 4 / 6 / 8 CTAs).
extern "C" int nvc_gemm_pair_init(nvc_ctx* ctx) {
#define X(NT) \
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(gemm_pair_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, PairCfg<NT>::kSmemBytes));
  NVC_PAIR_WIDTHS(X)
#undef X
  for (int S = 1; S <= 4; ++S) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * S * 64);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = PairCfg<256>::kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2 * S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, gemm_pair_kernel<256>, &cfg);
    if (e != cudaSuccess || n <= 0) {
      (void)cudaGetLastError();
      n = ctx->num_sms / (2 * S);
    }
    ctx->pair_clusters[S] = n;
  }
  if (getenv("NVC_DEBUG"))
    fprintf(stderr, "nvc: pair GEMM co-resident clusters of 2/4/6/8 CTAs: %d %d %d %d\n", ctx->pair_clusters[1],
            ctx->pair_clusters[2], ctx->pair_clusters[3], ctx->pair_clusters[4]);
  return NVC_OK;
}

// Schedule choice.
static void choose_schedule(const nvc_ctx* ctx, int M, int N, int kb, int* nt_out, int* s_out) {
  const int m_tiles = (M + kTileM - 1) / kTileM;
  auto avail = [&](int S) { return ctx->pair_clusters[S] > 0 ? ctx->pair_clusters[S] : ctx->num_sms / (2 * S); };
  if (m_tiles == 1) {
    // Decode (one m-tile): tile width x in-cluster split-K from a linear cost model fitted to
    // tools/gemm_microbench.py --narrow on B200 (profiles/r01_gemm_narrow_sweep_s12.txt: 146 measured
    // configurations of 11 shapes, the model's pick is within 1 % of the best measured on every shape):
    //   t[us] = waves * (ceil(kb / S) * (0.0738 + 0.00105 * NT) + 6.23 + [S > 1] * (1.99 + 0.161 * S))
    // Narrow tiles win where 256-wide ones quantise badly onto the 74 pairs (8B gate_up: 96 tiles = two
    // waves, the second 30 % full, vs 128 tiles of 192 = two full waves of 3/4 the work) and where a
    // split-K cluster would otherwise leave a third of the SMs idle (8B o_proj: 16 tiles x 3 splits
    // = 96 SMs vs 22 tiles of 192 x 3 = 132).
    double best = 1e30;
    int best_s = 1, best_nt = 256;
    static const int kWidths[] = {256, 224, 192, 160, 128};
    for (int nt : kWidths) {
      const int tiles = (N + nt - 1) / nt;
      for (int S = 1; S <= 4 && S <= kb; ++S) {
        const double waves = ceil((double)tiles / avail(S));
        if (S > 1 && waves > 2) continue;            // not measured: several waves of split-K clusters
        const double kbs = ceil((double)kb / S);
        const double t = waves * (kbs * (0.0738 + 0.00105 * nt) + 6.23 + (S > 1 ? 1.99 + 0.161 * S : 0.0));
        if (t < best * 0.995) { best = t; best_s = S; best_nt = nt; }
      }
    }
    *nt_out = best_nt;
    *s_out = best_s;
    return;
  }
  // Prefill (several m-tiles): 256-wide tiles; cycle model fitted to tools/gemm_microbench.py --pair:
  //  * a k-block costs ~2 * NT tensor cycles per pair (~600 for NT = 256 once the weights stream from HBM);
  //  * the persistent schedule (S = 1) pays ~10k cycles once (prologue, first loads, last epilogue);
  //    a split-K cluster pays ~12k (prologue, drain, publish, cluster barrier, reduce);
  //  * waves = ceil(tiles / co-resident clusters of that size).
  double best = 1e30;
  int best_s = 1, best_nt = 256;
  for (int S = 1; S <= 4 && S <= kb; ++S) {
    const int tiles = m_tiles * ((N + 255) / 256);
    const double waves = ceil((double)tiles / avail(S));
    const double kbs = ceil((double)kb / S);
    const double t = S == 1 ? waves * kbs * 560.0 + 10000.0 : waves * (kbs * 600.0 + 12000.0);
    if (t < best * 0.98) { best = t; best_s = S; best_nt = 256; }
  }
  *nt_out = best_nt;
  *s_out = best_s;
}

// force_splits: 0 = choose; 1..4 = in-cluster split-K factor with 256-wide tiles; 10000 * S + 1000 + NT = tile width NT
extern "C" int nvc_gemm_pair(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_f32,
                             int M, int N, int K, int force_splits, int epi, void* stream, long long* trace) {
  // force_splits >= 1000: 10000 * S + 1000 + NT (S omitted = 1)
  int nt = 256, S = 1;
  if (force_splits >= 1000) { nt = force_splits % 10000 - 1000; S = force_splits / 10000; S = S < 1 ? 1 : (S > 4 ? 4 : S); }
  else if (force_splits > 0) { S = force_splits > 4 ? 4 : force_splits; }
  else choose_schedule(ctx, M, N, K / BK, &nt, &S);
#define X(NT) \
  if (nt == NT) return launch_pair<NT>(ctx, A, lda, W, C, ldc, out_f32, M, N, K, S, epi, nvc_stream(stream), trace);
  NVC_PAIR_WIDTHS(X)
#undef X
  NVC_REQUIRE(ctx, false, NVC_EINVAL, "unsupported pair tile width");
  return NVC_OK;
}

// lm_head at decode batch sizes with the operands swapped: logits[M tokens, N vocab] (fp32) = X [M, K] . W [N, K]^T computed
// as W . X^T - 256 weight rows per MMA (128 per CTA of the pair, both CTAs' epilogues busy), the tokens as its 64 / 128
// columns - and stored transposed (NVC_EPI_TRANSPOSE).  The 64 tokens of a decode step cost a quarter of the tensor time
// of a 256-row token tile, and the weight stream is what the kernel waits for.
extern "C" int nvc_gemm_pair_swapped(nvc_ctx* ctx, const nvc_bf16* X, int ldx, const nvc_bf16* W, float* C, int ldc, int M,
                                     int N, int K, void* stream, long long* trace) {
  if (M < 1 || M > 128 || K % BK != 0) return NVC_ENOTIMPL;
  if (M <= 64) return launch_pair<64>(ctx, W, K, X, C, ldc, 1, N, M, K, 1, NVC_EPI_TRANSPOSE, nvc_stream(stream), trace, QkEpi{}, ldx);
  return launch_pair<128>(ctx, W, K, X, C, ldc, 1, N, M, K, 1, NVC_EPI_TRANSPOSE, nvc_stream(stream), trace, QkEpi{}, ldx);
}

// K3 + K2 in one launch for the fused QKV projection of a prefill-sized step (src/models/qwen3.rs:205-234): returns
// NVC_ENOTIMPL (nothing launched) when the shape does not take the persistent 256-wide schedule - the caller then runs
// nvc_gemm_bf16 + nvc_qk_norm_rope_store.  N = (nh + 2 kv) * 128.
extern "C" int nvc_gemm_pair_qk(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, nvc_bf16* C, int ldc, int M, int N,
                                int K, const nvc_bf16* qw, const nvc_bf16* kw, float eps, const float* cos_sin,
                                const int32_t* positions, const int32_t* slot_mapping, nvc_bf16* k_cache, nvc_bf16* v_cache,
                                int nh, int kv, void* stream) {
  // below ~8k rows the last wave's longer epilogue is exposed and the fusion loses (4 096 x 4 096 x 1 024: 48.3 vs 46.7 us)
  if (M < 8192 || N != (nh + 2 * kv) * 128 || K % BK != 0) return NVC_ENOTIMPL;
  int nt = 256, S = 1;
  choose_schedule(ctx, M, N, K / BK, &nt, &S);
  if (nt != 256 || S != 1) return NVC_ENOTIMPL;
  QkEpi qk;
  qk.qw = reinterpret_cast<const uint4*>(qw);
  qk.kw = reinterpret_cast<const uint4*>(kw);
  qk.cos_sin = reinterpret_cast<const float4*>(cos_sin);
  qk.positions = positions;
  qk.slot_mapping = slot_mapping;
  qk.k_cache = reinterpret_cast<uint4*>(k_cache);
  qk.v_cache = reinterpret_cast<uint4*>(v_cache);
  qk.eps = eps;
  qk.nh = nh;
  qk.kv = kv;
  return launch_pair<256>(ctx, A, lda, W, C, ldc, 0, M, N, K, 1, NVC_EPI_QK_NORM_ROPE, nvc_stream(stream), nullptr, qk);
}

// Public form: the fused QKV projection followed by K2, one launch when the shape takes the persistent pair schedule
// (prefill-sized M), the two-kernel sequence otherwise - the results are bit-identical either way.
extern "C" int nvc_qkv_proj_norm_rope(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W_qkv, nvc_bf16* qkv, int M, int K,
                                      const nvc_bf16* q_norm_w, const nvc_bf16* k_norm_w, float eps, const float* cos_sin,
                                      const int32_t* positions, const int32_t* slot_mapping, nvc_bf16* k_cache,
                                      nvc_bf16* v_cache, int nh, int kv, int hd, void* stream) {
  NVC_REQUIRE(ctx, ctx && A && W_qkv && qkv && q_norm_w && k_norm_w && cos_sin && positions, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, slot_mapping == nullptr || (k_cache && v_cache), NVC_EINVAL, "slot_mapping without caches");
  NVC_REQUIRE(ctx, hd == 128 && nh > 0 && kv > 0 && M >= 0, NVC_ESHAPE, "head_dim must be 128");
  if (M == 0) return NVC_OK;
  const int N = (nh + 2 * kv) * hd;
  static const bool fused = !(getenv("NVC_QKV_FUSED") && atoi(getenv("NVC_QKV_FUSED")) == 0);
  int rc = (fused && g_nvc_debug[7] != 1)
               ? nvc_gemm_pair_qk(ctx, A, lda, W_qkv, qkv, N, M, N, K, q_norm_w, k_norm_w, eps, cos_sin, positions, slot_mapping,
                                  k_cache, v_cache, nh, kv, stream)
               : NVC_ENOTIMPL;
  if (rc != NVC_ENOTIMPL) return rc;
  rc = nvc_gemm_bf16(ctx, A, lda, W_qkv, qkv, N, NVC_DT_BF16, M, N, K, stream);
  if (rc) return rc;
  return nvc_qk_norm_rope_store(ctx, qkv, q_norm_w, k_norm_w, eps, cos_sin, positions, slot_mapping, k_cache, v_cache, M, nh, kv,
                                hd, stream);
}
