// K3: C[M,N] = A[M,K] * W[N,K]^T, bf16 x bf16 -> fp32 accumulate in tensor memory.
// Replaces candle_nn::Linear::forward under every *ParallelLinear::forward
// (src/layers/linear.rs:33-38,72-77,99-104,171-175,184-198) and compute_logits
// (src/models/qwen3.rs:542-550).
//
// One 128 x BN output tile per CTA, optional split-K over gridDim.z.  A pipeline stage holds
// KPS k-blocks of 64 (KPS x [128 x 64] A sub-tiles + KPS x [BN x 64] W sub-tiles, each a
// 128B-swizzled TMA box), so one mbarrier round trip (~130 cycles) is amortised over 4*KPS MMAs.
//   warp 0   : TMA producer for the activation sub-tiles (only the rows that exist when M < 128)
//   warp 2   : TMA producer for the weight sub-tiles, then joins the epilogue
//   warp 1   : allocates TMEM; one elected lane issues tcgen05.mma (M=128, N=BN, K=16);
//              tcgen05.commit frees the stage / publishes the accumulator
//   warps 2-5: epilogue - tcgen05.ld 32 lanes x 16 columns, convert, 16-byte stores
// All role loops are executed by the whole warp with one elected lane issuing, which keeps the
// TMA / UMMA operands in uniform registers (back-to-back UTMALDG / UTCHMMA in SASS).
// Split-K: each split writes its fp32 partial tile to a workspace; the CTA that takes the last
// ticket (acq_rel atomic) sums the partials in split order (deterministic, coalesced).
// Rows >= M and columns >= N of edge tiles are zero-filled by TMA and masked at the store.
#include "common.cuh"
#include "ptx.cuh"
#include "internal.h"
#include "gemm_epilogue.cuh"

#include <cstdlib>

static int g_nvc_gemm_prefetch = getenv("NVC_GEMM_PREFETCH") ? atoi(getenv("NVC_GEMM_PREFETCH")) : 1;

static int g_nvc_gemm_ks = getenv("NVC_GEMM_KS") ? atoi(getenv("NVC_GEMM_KS")) : 1;
static int g_nvc_gemm_pair = getenv("NVC_GEMM_PAIR") ? atoi(getenv("NVC_GEMM_PAIR")) : 1;
// smallest N whose fp32 output at M <= 128 goes through the swapped-operand pair schedule (lm_head at decode batch sizes)
static int g_nvc_swap_min_n = getenv("NVC_SWAP_MIN_N") ? atoi(getenv("NVC_SWAP_MIN_N")) : 32768;

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 192;
constexpr int kEpilogueThreads = 128;
constexpr int kSmemBudget = 200 * 1024;

template <int BN, int KPS>
struct Cfg {
  static constexpr int kBytesA = BM * BK * 2;                 // one A sub-tile
  static constexpr int kBytesB = BN * BK * 2;                 // one W sub-tile
  static constexpr int kStageBytes = KPS * (kBytesA + kBytesB);
  static constexpr int kStages = kSmemBudget / kStageBytes < 8 ? kSmemBudget / kStageBytes : 8;
  static constexpr int kTmemCols = BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256));   // power of two >= 32
  // tiles + (2*stages + 1) mbarriers + tmem slot + ticket, plus 1024 for manual alignment
  static constexpr int kSmemBytes = kStages * kStageBytes + (2 * kStages + 1) * 8 + 16 + 1024;
  static_assert(kStages >= 2, "need at least a double buffer");
};

__device__ __forceinline__ int atomic_add_acq_rel_gpu(int* p, int v) {
  int old;
  asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

template <int BN, int KPS, int CL>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         void* __restrict__ C, int ldc, int out_f32, int M, int N, int k_blocks_total,
                         int k_blocks_per_split, int a_box_rows, int prefetch_w, int ks_cluster, int epi, float* __restrict__ ws,
                         int* __restrict__ counters, long long* __restrict__ trace) {
  using cfg = Cfg<BN, KPS>;
  // optional phase timestamps of CTA (0,0,0) (tools/gemm_microbench.py); NULL in production
  const bool tr = trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  pdl_trigger();   // the next kernel may start its own prologue
  if (tr && threadIdx.x == 0) trace[0] = clock64();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + cfg::kStages * cfg::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (cfg::kStages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * cfg::kStages);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + cfg::kStages * cfg::kStageBytes + (2 * cfg::kStages + 1) * 8);
  volatile int* ticket_gen = reinterpret_cast<volatile int*>(const_cast<uint32_t*>(tmem_slot_gen) + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x, m_tile = blockIdx.y, split = blockIdx.z;
  const int num_splits = gridDim.z;
  const int kb_begin = split * k_blocks_per_split;
  const int kb_end = min(k_blocks_total, kb_begin + k_blocks_per_split);
  const int num_kb = kb_end - kb_begin;
  const int num_it = (num_kb + KPS - 1) / KPS;       // pipeline iterations (stages consumed)
  // CL > 1: the CL CTAs of a cluster work on CL adjacent n-tiles of the same m-tile / k-range and share
  // the activation tile: each loads 1/CL of its rows and multicasts them to all (a_box_rows = rows per CTA)
  const uint32_t cta_rank = CL > 1 ? ptx::cluster_ctarank() : 0u;
  constexpr uint16_t kMask = (uint16_t)((1u << CL) - 1u);

  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < cfg::kStages; ++s) {
        ptx::mbar_init(full_bar(s), 2);   // two producers: warp 0 (A sub-tiles) and warp 2 (W sub-tiles)
        ptx::mbar_init(empty_bar(s), CL);   // released by the MMA warps of all CTAs that received the multicast
      }
      ptx::mbar_init(tmem_full_bar, 1);
      ptx::fence_barrier_init();
    }
    __syncwarp();
    ptx::tmem_alloc(tmem_slot, cfg::kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;
  if (CL > 1) ptx::cluster_sync_all();   // peers' barriers are initialised before anyone multicasts into them
  // Weights are never written by a kernel, so their first kPrefetchKb k-blocks are pulled into L2
  // BEFORE the dependency wait (overlapping the predecessor), and the producer keeps prefetching
  // that far ahead of its shared-memory ring: a CTA's ring (~190 KB) covers only ~60 B/clk at HBM
  // latency, the L2 prefetch decouples the bytes in flight from the ring size.
  constexpr int kPrefetchKb = BN >= 256 ? 8 : (BN >= 128 ? 16 : 32);
  // Decode (prefetch_w): the W sub-tiles of the first ring lap are not only prefetched but LOADED into shared
  // memory before the wait - a K = 1024 GEMM has its whole weight tile on chip when the activations arrive, and
  // only the A boxes (the larger share of the ingest at small BN) are left on the critical path.
  const int w_pre = prefetch_w ? min(num_it, cfg::kStages) : 0;
  if (warp == 2 && prefetch_w) {
    if (ptx::elect_one()) {
      for (int it = 0; it < w_pre; ++it) {
        const int nsub = min(KPS, num_kb - it * KPS);
        ptx::mbar_arrive_expect_tx(full_bar(it), (uint32_t)cfg::kBytesB * nsub);
        const uint32_t dst = smem_base + it * cfg::kStageBytes + (uint32_t)(KPS * cfg::kBytesA);
#pragma unroll
        for (int j = 0; j < KPS; ++j)
          if (j < nsub) ptx::tma_load_2d(dst + j * cfg::kBytesB, &tma_b, full_bar(it), (kb_begin + it * KPS + j) * BK, n_tile * BN);
      }
      const int k0 = w_pre * KPS, k1 = min(num_kb, k0 + kPrefetchKb);
      for (int kb = k0; kb < k1; ++kb) ptx::tma_prefetch_l2_2d(&tma_b, (kb_begin + kb) * BK, n_tile * BN);
    }
    __syncwarp();
  }
  pdl_wait();      // everything above (barrier init, TMEM alloc, descriptor prefetch, first W loads) overlapped the predecessor
  if (tr && threadIdx.x == 0) trace[1] = clock64();

  if (warp == 0 || warp == 2) {
    // ------------------------------------------------------------ TMA producers
    const bool is_a = warp == 0;
    // bytes ONE box delivers into this CTA per k-block (A: all CL slices land here)
    const uint32_t sub_bytes = is_a ? (uint32_t)a_box_rows * CL * (BK * 2) : (uint32_t)cfg::kBytesB;
    const uint32_t sub_stride = is_a ? (uint32_t)cfg::kBytesA : (uint32_t)cfg::kBytesB;
    const CUtensorMap* map = is_a ? &tma_a : &tma_b;
    const int row0 = is_a ? m_tile * BM + (int)cta_rank * a_box_rows : n_tile * BN;
    const uint32_t off = is_a ? cta_rank * (uint32_t)a_box_rows * (BK * 2) : (uint32_t)(KPS * cfg::kBytesA);
    for (int it = is_a ? 0 : w_pre; it < num_it; ++it) {
      const int s = it % cfg::kStages;
      if (it >= cfg::kStages) ptx::mbar_wait(empty_bar(s), ((it / cfg::kStages) & 1) ^ 1u);   // first lap: slots are free
      if (ptx::elect_one()) {
        const int nsub = min(KPS, num_kb - it * KPS);
        ptx::mbar_arrive_expect_tx(full_bar(s), sub_bytes * nsub);
        const uint32_t dst = smem_base + s * cfg::kStageBytes + off;
        const int k0 = (kb_begin + it * KPS) * BK;
#pragma unroll
        for (int j = 0; j < KPS; ++j) {
          if (j < nsub) {
            if (CL > 1 && is_a) ptx::tma_load_2d_mcast(dst + j * sub_stride, map, full_bar(s), k0 + j * BK, row0, kMask);
            else ptx::tma_load_2d(dst + j * sub_stride, map, full_bar(s), k0 + j * BK, row0);
          }
        }
        if (!is_a && prefetch_w) {       // stay kPrefetchKb k-blocks ahead in L2
#pragma unroll
          for (int j = 0; j < KPS; ++j) {
            const int kb = it * KPS + j + kPrefetchKb;
            if (kb < num_kb) ptx::tma_prefetch_l2_2d(&tma_b, (kb_begin + kb) * BK, row0);
          }
        }
        if (tr && is_a && it < 32) trace[2 + it] = clock64();
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN, 0, 0);
    constexpr uint32_t desc_hi = ptx::umma_desc_hi_sw128(1024);
    const uint32_t a_lo0 = ptx::umma_desc_lo(smem_base, 16);
    const uint32_t b_lo0 = ptx::umma_desc_lo(smem_base + KPS * cfg::kBytesA, 16);
    for (int it = 0; it < num_it; ++it) {
      const int s = it % cfg::kStages;
      ptx::mbar_wait(full_bar(s), (it / cfg::kStages) & 1);
      ptx::tcgen05_fence_after();
      if (ptx::elect_one()) {
        if (tr && it < 32) trace[40 + it] = clock64();
        const int nsub = min(KPS, num_kb - it * KPS);
        const uint32_t a_lo = a_lo0 + s * (cfg::kStageBytes >> 4);
        const uint32_t b_lo = b_lo0 + s * (cfg::kStageBytes >> 4);
#pragma unroll
        for (int j = 0; j < KPS; ++j) {
          if (j < nsub) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              // 16 bf16 = 32 bytes along K inside the 128-byte swizzle span: +2 in the address field
              ptx::umma_bf16_ss(tmem_base, ptx::umma_desc(a_lo + j * (cfg::kBytesA >> 4) + 2 * k, desc_hi),
                                ptx::umma_desc(b_lo + j * (cfg::kBytesB >> 4) + 2 * k, desc_hi), idesc,
                                (it | j | k) != 0 ? 1u : 0u);
            }
          }
        }
        if (CL > 1) ptx::umma_commit_mcast(empty_bar(s), kMask);   // every CTA of the cluster wrote into this stage
        else ptx::umma_commit(empty_bar(s));       // stage reusable once these MMAs have read it
        if (it == num_it - 1) {
          ptx::umma_commit(tmem_full_bar);    // accumulator complete
          if (tr) trace[80] = clock64();
        }
      }
      __syncwarp();
    }
  }
  if (warp >= 2) {
    // ------------------------------------------------------------ epilogue (warps 2..5)
    const int q = warp & 3;                 // TMEM lane quadrant this warp may access
    const int r = q * 32 + lane;            // row inside the tile
    const int grow = m_tile * BM + r;
    const bool row_ok = grow < M;
    ptx::mbar_wait(tmem_full_bar, 0);
    ptx::tcgen05_fence_after();
    if (tr && threadIdx.x == 64) trace[81] = clock64();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int tile_id = m_tile * gridDim.x + n_tile;

    if (num_splits == 1) {
      if constexpr (BN >= 32) {
#pragma unroll 1
        for (int c = 0; c < BN / 16; c += 2) {          // two 16-column chunks per tensor-memory round trip
          uint32_t v[32];
          ptx::tmem_ld_32x32b_x16(taddr + c * 16, v);
          ptx::tmem_ld_32x32b_x16(taddr + c * 16 + 16, v + 16);
          ptx::tmem_ld_wait();
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (row_ok) emit32(C, out_f32, epi, (int64_t)grow * ldc, n_tile * BN + c * 16, N, f);
        }
      } else {
        uint32_t v[16];
        ptx::tmem_ld_32x32b_x16(taddr, v);
        ptx::tmem_ld_wait();
        float f[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
        if (row_ok && n_tile * BN < N) store16(C, out_f32, (int64_t)grow * ldc + n_tile * BN, f);
      }
    } else if (ks_cluster) {
      // ---- split-K inside a cluster (1,1,S): park this CTA's fp32 partial tile in its own, now idle,
      // operand ring as [float4 column group][row] (conflict-free); summed after the cluster barrier below
      ptx::fence_proxy_async_smem();
#pragma unroll 1
      for (int c = 0; c < BN / 16; ++c) {
        uint32_t v[16];
        ptx::tmem_ld_32x32b_x16(taddr + c * 16, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j)
          ptx::st_shared_f4(smem_base + (uint32_t)(((c * 4 + j) * BM + r) * 16), __uint_as_float(v[4 * j]),
                            __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
      }
    } else {
      // ---- split-K: publish the partial, last arriver reduces in split order
      float* my = ws + ((int64_t)tile_id * num_splits + split) * (BM * BN) + (int64_t)r * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 16; ++c) {
        uint32_t v[16];
        ptx::tmem_ld_32x32b_x16(taddr + c * 16, v);
        ptx::tmem_ld_wait();
        if (row_ok) {
          float4* dst = reinterpret_cast<float4*>(my + c * 16);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            __stcg(dst + j, make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                        __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])));
        }
      }
      // CTA barrier orders every thread's stores before thread 64's release; the acquire half of the
      // same atomic plus the second barrier orders the last CTA's loads after all other CTAs' stores.
      ptx::named_bar_sync(1, kEpilogueThreads);
      if (threadIdx.x == 64) *ticket_gen = atomic_add_acq_rel_gpu(counters + tile_id, 1);
      ptx::named_bar_sync(1, kEpilogueThreads);
      if (*ticket_gen == num_splits - 1) {
        // coalesced reduction over the valid rows of the tile: thread t takes float4 #(t + 128 i);
        // the partials of all splits are summed in split order (deterministic)
        const float* tile_ws = ws + (int64_t)tile_id * num_splits * (BM * BN);
        const int rows_valid = min(BM, M - m_tile * BM);
        const int cols_valid = min(BN, N - n_tile * BN);
        constexpr int kVecPerRow = BN / 4;
        const int et = threadIdx.x - 64;
#pragma unroll 4
        for (int v = et; v < rows_valid * kVecPerRow; v += kEpilogueThreads) {
          const int rr = v / kVecPerRow, c4 = v - rr * kVecPerRow;
          if (c4 * 4 >= cols_valid) continue;
          const float4* src = reinterpret_cast<const float4*>(tile_ws + (int64_t)rr * BN) + c4;
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          int z = 0;
          for (; z + 4 <= num_splits; z += 4) {
            const float4 f0 = __ldcg(src + (int64_t)(z + 0) * (BM * BN / 4));
            const float4 f1 = __ldcg(src + (int64_t)(z + 1) * (BM * BN / 4));
            const float4 f2 = __ldcg(src + (int64_t)(z + 2) * (BM * BN / 4));
            const float4 f3 = __ldcg(src + (int64_t)(z + 3) * (BM * BN / 4));
            acc.x = (((acc.x + f0.x) + f1.x) + f2.x) + f3.x;
            acc.y = (((acc.y + f0.y) + f1.y) + f2.y) + f3.y;
            acc.z = (((acc.z + f0.z) + f1.z) + f2.z) + f3.z;
            acc.w = (((acc.w + f0.w) + f1.w) + f2.w) + f3.w;
          }
          for (; z < num_splits; ++z) {
            const float4 f = __ldcg(src + (int64_t)z * (BM * BN / 4));
            acc.x += f.x; acc.y += f.y; acc.z += f.z; acc.w += f.w;
          }
          const int64_t o = (int64_t)(m_tile * BM + rr) * ldc + n_tile * BN + c4 * 4;
          if (out_f32) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(C) + o) = acc;
          } else {
            uint2 pk;
            pk.x = pack_bf16(acc.x, acc.y);
            pk.y = pack_bf16(acc.z, acc.w);
            *reinterpret_cast<uint2*>(reinterpret_cast<nvc_bf16*>(C) + o) = pk;
          }
        }
        if (threadIdx.x == 64) counters[tile_id] = 0;   // re-arm for the next launch
      }
    }
    if (tr && threadIdx.x == 64 && !ks_cluster) trace[82] = clock64();
    ptx::tcgen05_fence_before();
  }
  if (ks_cluster) {
    // Distributed-shared-memory reduction: CTA `split` of the cluster owns the float4 column groups
    // g with g % S == split and sums them over the S partial tiles in split order (deterministic);
    // no global workspace, no atomics.
    ptx::cluster_sync_all();             // all partials parked (release / acquire at cluster scope)
    if (warp >= 2) {
      // work items = (owned column group, valid row), spread over all 128 epilogue threads (the partials sit in
      // shared memory, so any thread can sum any element; consecutive threads take consecutive rows of a group:
      // contiguous DSMEM reads).  Two items per round trip: DSMEM latency, not bandwidth, bounds this phase.
      const int et = threadIdx.x - 64;
      const int rows_valid = min(BM, M - m_tile * BM);
      const int cols_valid = min(BN, N - n_tile * BN);
      const int owned = (BN / 4 - split + num_splits - 1) / num_splits;        // groups split, split + S, ...
      const int items = owned * rows_valid;
      for (int i0 = et; i0 < items; i0 += 2 * kEpilogueThreads) {
        float4 t[2][8];
        int rr[2], gg[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int idx = min(i0 + u * kEpilogueThreads, items - 1);           // clamped: a redundant load, no branch
          const int gk = idx / rows_valid;
          rr[u] = idx - gk * rows_valid;
          gg[u] = split + gk * num_splits;
          const uint32_t mine = smem_base + (uint32_t)((gg[u] * BM + rr[u]) * 16);
#pragma unroll
          for (int z = 0; z < 8; ++z)
            if (z < num_splits) t[u][z] = ptx::ld_cluster_f4(ptx::mapa_u32(mine, (uint32_t)z));
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (i0 + u * kEpilogueThreads >= items || gg[u] * 4 >= cols_valid) continue;
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int z = 0; z < 8; ++z)                                          // split order: deterministic
            if (z < num_splits) { acc.x += t[u][z].x; acc.y += t[u][z].y; acc.z += t[u][z].z; acc.w += t[u][z].w; }
          const int64_t o = (int64_t)(m_tile * BM + rr[u]) * ldc + n_tile * BN + gg[u] * 4;
          if (out_f32) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(C) + o) = acc;
          } else {
            uint2 pk;
            pk.x = pack_bf16(acc.x, acc.y);
            pk.y = pack_bf16(acc.z, acc.w);
            *reinterpret_cast<uint2*>(reinterpret_cast<nvc_bf16*>(C) + o) = pk;
          }
        }
      }
      if (tr && threadIdx.x == 64) trace[82] = clock64();
    }
  }
  __syncthreads();
  if (CL > 1 || ks_cluster) ptx::cluster_sync_all();   // nobody exits while a peer may still arrive on its barriers / read its smem
  if (warp == 1) {
    ptx::tcgen05_fence_after();
    ptx::tmem_dealloc(tmem_base, cfg::kTmemCols);
  }
  if (tr && threadIdx.x == 0) trace[83] = clock64();
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

template <int BN, int KPS, int CL>
int launch(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_f32, int M,
           int N, int K, int splits, int ks_cluster, int epi, cudaStream_t stream, long long* trace) {
  using cfg = Cfg<BN, KPS>;
  CUtensorMap ta, tb;
  // a single partial m-tile only needs the rows that exist (rounded to the 8-row swizzle atom per
  // cluster CTA); each CTA of a cluster loads a_rows / CL of them
  const int a_rows = M >= BM ? BM : ((M + 8 * CL - 1) / (8 * CL)) * (8 * CL);
  int rc = make_tma_2d(ctx, &ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK, a_rows / CL);
  if (rc) return rc;
  rc = make_tma_2d(ctx, &tb, W, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, BK, BN);
  if (rc) return rc;
  const int kb_total = K / BK;
  const int kb_per = (kb_total + splits - 1) / splits;
  const int eff_splits = (kb_total + kb_per - 1) / kb_per;   // no empty splits
  const int n_tiles = (N + BN - 1) / BN;
  NVC_REQUIRE(ctx, n_tiles % CL == 0, NVC_ESHAPE, "n-tiles must divide by the cluster size");
  dim3 grid(n_tiles, (M + BM - 1) / BM, eff_splits);
  // ks_cluster: the eff_splits CTAs of one output tile form a cluster (1,1,S) and reduce over DSMEM
  const int ks = (ks_cluster && CL == 1 && eff_splits > 1 && eff_splits <= 8) ? 1 : 0;
  NVC_REQUIRE(ctx, !ks || (size_t)BN * BM * 4 <= (size_t)cfg::kStages * cfg::kStageBytes, NVC_EINVAL, "ring too small for the partial tile");
  NVC_CUDA_CHECK(ctx, nvc_launch_cluster3(ctx, gemm_bf16_tcgen05_kernel<BN, KPS, CL>, grid, dim3(kThreads), cfg::kSmemBytes,
                                          stream, dim3(CL, 1, ks ? eff_splits : 1), ta, tb, C, ldc, out_f32, M, N, kb_total,
                                          kb_per, a_rows / CL,
                                          // first ring lap of W before the dependency wait + L2 prefetch ahead of the ring:
                                          // helps the short, latency-bound k-loops of decode (K <= 3072: 5.1 vs 5.7 us), costs
                                          // 2 - 8 % on the long ones (K 12288 at M 64: 23.2 -> 21.6 us without it) and at M >= 256
                                          (g_nvc_gemm_prefetch && M <= BM && kb_total <= 48) ? 1 : 0,
                                          ks, epi, ctx->gemm_ws, ctx->gemm_counters, trace));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// (BN, KPS) instantiations.  KPS > 1 (several k-blocks per mbarrier round trip) is for short,
// latency-bound CTAs (decode); KPS = 1 keeps the deepest ring for long compute-bound k-loops.
// CL > 1: cluster of CL CTAs along N sharing the activation tile through TMA multicast (decode).
#define NVC_GEMM_CONFIGS(X)                                                                          \
  X(256, 1, 1) X(128, 1, 1) X(128, 2, 1) X(64, 2, 1) X(64, 4, 1) X(32, 4, 1)                       \
  X(16, 4, 8) X(16, 4, 4) X(32, 4, 4) X(32, 4, 8) X(64, 4, 4) X(64, 2, 4) X(96, 2, 4) X(128, 2, 4) X(128, 1, 4) X(256, 1, 2)

// Tile choice, fitted to tools/gemm_microbench.py (CUDA-graph timing + CTA phase traces) on B200:
//  * a CTA ingests ~60 B/clk through TMA when the operands stream from HBM, so a k-block costs
//    max(2*BN [tcgen05 floor], (rows + BN) * 128 / 60 [TMA]) cycles, plus ~2500 cycles of
//    prologue + first-load latency + epilogue per CTA;
//  * decode (M <= 128): one wave, as many CTAs as the N dimension gives (up to ~all SMs), several
//    k-blocks per pipeline stage; split-K through the L2 workspace costs >= 6 us (ticket atomic +
//    reduction round trips) and never won in the sweep, so it is only used when forced.
void choose_config(const nvc_ctx* ctx, int M, int N, int K, bool allow_split, int* bn_out, int* kps_out, int* splits_out,
                   int* cl_out, int* ks_out) {
  *splits_out = 1;
  *cl_out = 1;
  *ks_out = 0;
  if (M <= BM) {
    // Decode (one m-tile).  Cycle model fitted to CTA phase traces (tools/gemm_microbench.py --trace-decode):
    // the TMA unit of an SM moves ~90 B/clk plus ~30 cycles per box (two loads + one L2 prefetch per k-block),
    // whatever the source (the L2 prefetch before the PDL wait hides HBM), and the activation rows are re-read
    // by every n-tile, so a k-block costs (rows + BN) * 128 / 90 + 90 cycles; ~2200 cycles of prologue +
    // first-load latency; epilogue 300 + 6 * BN, or, with split-K inside a cluster (1,1,S), ~2300 + the DSMEM
    // exchange at ~17 B/clk.  The winners are always ONE wave of ~96-132 CTAs with the widest tile that allows it
    // (co-resident: 148 CTAs, 74 clusters of 2, 33 clusters of 4).
    const int rows = ((M + 7) / 8) * 8;
    const int kb = K / BK;
    const int bns[4] = {32, 64, 128, 256};
    const int ss[3] = {1, 2, 4};
    double best = 1e30;
    int best_bn = 32, best_s = 1;
    const double hbm_floor = (double)N * K * 2 / 3400.0 + 3000.0;     // 6.5 TB/s at 1.9 GHz
    for (int i = 0; i < 4; ++i) {
      for (int j = 0; j < 3; ++j) {
        const int bn = bns[i], S = ss[j];
        if (S > 1 && (!allow_split || !g_nvc_gemm_ks || kb / S < 4)) continue;
        if (bn / 2 >= N) continue;
        const int tiles = (N + bn - 1) / bn;
        const int cap = S == 1 ? ctx->num_sms : (S == 2 ? 2 * (ctx->pair_clusters[1] > 0 ? ctx->pair_clusters[1] : ctx->num_sms / 2)
                                                       : 4 * (ctx->pair_clusters[2] > 0 ? ctx->pair_clusters[2] : ctx->num_sms / 5));
        const double waves = ceil((double)tiles * S / cap);
        const double per_kb = (double)(rows + bn) * 128.0 / 90.0 + 90.0;
        const double tail = S == 1 ? 300.0 + 6.0 * bn : 2300.0 + (double)(S - 1) * rows * (bn / S) * 4.0 / 17.0;
        const double t = waves * (2200.0 + per_kb * ((kb + S - 1) / S) + tail);
        // below the HBM floor every candidate takes the floor: rank those by the modeled CTA time (most slack)
        const double key = t > hbm_floor ? t : hbm_floor * (1.0 + 1e-3 * t / hbm_floor);
        if (key < best * 0.98 || (key < best && t < hbm_floor)) { best = key; best_bn = bn; best_s = S; }
      }
    }
    *bn_out = best_bn;
    *kps_out = best_bn == 256 ? 1 : (best_bn == 128 ? 2 : 4);
    *splits_out = best_s;
    *ks_out = best_s > 1 ? 1 : 0;
    return;
  }
  const int kb = K / BK;
  const int m_tiles = (M + BM - 1) / BM;
  const int cands[3] = {256, 128, 64};
  double best = 1e30;
  int best_bn = 256;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (bn / 2 >= N) continue;                       // tile mostly empty
    const double tiles = (double)m_tiles * ((N + bn - 1) / bn);
    const double waves = ceil(tiles / ctx->num_sms);
    const double per_kb = fmax(2.0 * bn, (BM + bn) * 128.0 / 60.0);
    const double t = waves * (kb * per_kb + 2500.0 + bn * 18.0);
    if (t < best * 0.999) { best = t; best_bn = bn; }
  }
  *bn_out = best_bn;
  *kps_out = best_bn == 64 ? 2 : 1;
}

}  // namespace

extern "C" int nvc_gemm_init(nvc_ctx* ctx) {
#define X(BN, KPS, CL)                                                               \
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, KPS, CL>,    \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN, KPS>::kSmemBytes));
  NVC_GEMM_CONFIGS(X)
#undef X
  return NVC_OK;
}

static int gemm_impl(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                     int out_dtype, int M, int Nout, int K, int force_bn, int force_splits, int epi,
                     void* stream, long long* trace) {
  NVC_REQUIRE(ctx, ctx && A && W && C, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, M >= 0 && Nout > 0 && K > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, K % 64 == 0, NVC_ESHAPE, "K must be a multiple of 64");
  NVC_REQUIRE(ctx, Nout % 16 == 0, NVC_ESHAPE, "Nout must be a multiple of 16");
  NVC_REQUIRE(ctx, lda % 8 == 0 && lda >= K, NVC_ESHAPE, "lda must be >= K and a multiple of 8");
  NVC_REQUIRE(ctx, ldc >= (epi == NVC_EPI_SILU_MUL ? Nout / 2 : Nout) && ldc % 8 == 0, NVC_ESHAPE,
              "ldc must be >= the output width and a multiple of 8");
  NVC_REQUIRE(ctx, epi == NVC_EPI_NONE || (epi == NVC_EPI_SILU_MUL && Nout % 32 == 0 && out_dtype == NVC_DT_BF16),
              NVC_ESHAPE, "fused SiLU-mul epilogue needs Nout % 32 == 0 and bf16 output");
  NVC_REQUIRE(ctx, out_dtype == NVC_DT_BF16 || out_dtype == NVC_DT_F32, NVC_EINVAL, "bad out_dtype");
  NVC_REQUIRE(ctx, ((uintptr_t)A % 16) == 0 && ((uintptr_t)W % 16) == 0 && ((uintptr_t)C % 16) == 0, NVC_EINVAL,
              "pointers must be 16-byte aligned");
  NVC_REQUIRE(ctx, ctx->encode_tiled != nullptr, NVC_EDEVICE, "cuTensorMapEncodeTiled unavailable");
  if (M == 0) return NVC_OK;
  // M > 128: CTA-pair kernel (gemm_pair.cu); force_bn = 512 selects it explicitly, any other force_bn the 1-CTA tiles
  // lm_head at decode batch sizes (vocabulary-sized N, short K): hundreds of one-k-loop tiles per SM pair leave the
  // 1-CTA kernel's per-tile prologue / epilogue exposed; the pair kernel's persistent schedule overlaps the epilogue
  // with the next tile (M 64, N 151936, K 1024: 75.1 -> 67.3 us; slower at M = 1 and at K = 4096,
  // profiles/r02_lm_head_sweep_pair.txt)
  const bool vocab_pair = force_bn == 0 && force_splits == 0 && epi == NVC_EPI_NONE && M >= 32 && M <= BM && Nout >= 65536 &&
                          K <= 1024 && g_nvc_gemm_pair;
  // ... and with fp32 logits the operands go in swapped (weights as the 256-row MMA operand, transposed store): force_bn 513
  // (fp32 logits for up to 128 tokens, any K: 64 x 151 936 x 1 024 68.0 -> 64.1 us, M = 1 62.3 -> 53.8, K = 4 096 224.7 ->
  // 202.5 = 0.97 of the HBM peak; profiles/r02_lm_head_sweep_swapped.txt)
  // Down to vocabulary shards of TP 4 for M <= 64 (37 984 columns: 25.5 -> 20.5 us); at 65..128 tokens only where it measured
  // ahead (full vocabulary, or K >= 2 048 from 65 536 columns); the TP 8 shard (18 992) stays on the 1-CTA kernel (11.2 vs 13.2).
  const bool swap_n = M <= 64 ? Nout >= g_nvc_swap_min_n : (Nout >= 131072 || (Nout >= 65536 && K >= 2048));
  const bool vocab_swapped = force_bn == 0 && force_splits == 0 && epi == NVC_EPI_NONE && out_dtype == NVC_DT_F32 && M <= BM &&
                             swap_n && g_nvc_gemm_pair;
  if (vocab_swapped || (force_bn == 513 && out_dtype == NVC_DT_F32 && epi == NVC_EPI_NONE && M <= BM))
    return nvc_gemm_pair_swapped(ctx, A, lda, W, reinterpret_cast<float*>(C), ldc, M, Nout, K, stream, trace);
  if (force_bn == 512 || vocab_pair || (force_bn == 0 && M > BM && g_nvc_gemm_pair))
    return nvc_gemm_pair(ctx, A, lda, W, C, ldc, out_dtype == NVC_DT_F32, M, Nout, K, force_splits, epi, stream, trace);
  int bn, kps, splits, cl = 1, ks = 0;
  choose_config(ctx, M, Nout, K, epi == NVC_EPI_NONE, &bn, &kps, &splits, &cl, &ks);
  if (force_bn) {          // force_bn = BN + 1000 * KPS + 10000 * CL (KPS 0: the first instantiation of that BN)
    bn = force_bn % 1000;
    kps = (force_bn / 1000) % 10;
    cl = force_bn / 10000;
    if (cl == 0) cl = 1;
    if (kps == 0) kps = bn >= 128 ? 1 : (bn == 64 ? 2 : 4);
    ks = 0;
    splits = 1;
  }
  // force_splits: 1..99 = split-K through the global workspace; 100 + S = split-K inside a cluster (1,1,S), S <= 8
  if (force_splits >= 100) { splits = force_splits - 100; ks = 1; cl = 1; }
  else if (force_splits) { splits = force_splits; ks = 0; }
  NVC_REQUIRE(ctx, !ks || splits <= 8, NVC_EINVAL, "cluster split-K supports at most 8 splits");
  if (epi != NVC_EPI_NONE) {       // the fused epilogue lives in the direct-store path: no split-K, tiles of >= 32 columns
    NVC_REQUIRE(ctx, force_splits == 0 || force_splits == 1, NVC_EINVAL, "fused epilogue cannot be combined with split-K");
    splits = 1;
    ks = 0;
    if (bn < 32) { bn = 32; kps = 4; cl = 1; }
  }
  if (splits > K / BK) splits = K / BK;
  const int tiles = ((M + BM - 1) / BM) * ((Nout + bn - 1) / bn);
  if (splits > 1 && !ks) {
    NVC_REQUIRE(ctx, (size_t)tiles * splits * BM * bn * 4 <= ctx->gemm_ws_bytes && tiles <= ctx->gemm_num_counters,
                NVC_ENOMEM, "split-K workspace too small");
  }
  cudaStream_t st = nvc_stream(stream);
  const int f32 = out_dtype == NVC_DT_F32;
#define X(BN, KPS, CL) \
  if (bn == BN && kps == KPS && cl == CL) return launch<BN, KPS, CL>(ctx, A, lda, W, C, ldc, f32, M, Nout, K, splits, ks, epi, st, trace);
  NVC_GEMM_CONFIGS(X)
#undef X
  NVC_REQUIRE(ctx, false, NVC_EINVAL, "unsupported tile configuration");
  return NVC_OK;
}

extern "C" int nvc_gemm_bf16_ex(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                                int out_dtype, int M, int Nout, int K, int force_bn, int force_splits,
                                void* stream) {
  return gemm_impl(ctx, A, lda, W, C, ldc, out_dtype, M, Nout, K, force_bn, force_splits, NVC_EPI_NONE, stream, nullptr);
}

// Debug: same GEMM, CTA (0,0,0) writes clock64() phase stamps to trace_d[0..128) (device memory).
extern "C" int nvc_gemm_bf16_trace(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                                   int out_dtype, int M, int Nout, int K, int force_bn, int force_splits,
                                   void* stream, long long* trace_d) {
  return gemm_impl(ctx, A, lda, W, C, ldc, out_dtype, M, Nout, K, force_bn, force_splits, NVC_EPI_NONE, stream, trace_d);
}

extern "C" int nvc_gemm_bf16(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                             int out_dtype, int M, int Nout, int K, void* stream) {
  return nvc_gemm_bf16_ex(ctx, A, lda, W, C, ldc, out_dtype, M, Nout, K, 0, 0, stream);
}

// K3 + K6 fused (Qwen3MLP::forward, src/models/qwen3.rs:323-327): act = silu(A Wg^T) * (A Wu^T) with the
// gate/up rows of W interleaved in groups of 16 (see include/nvc_b200.h).
extern "C" int nvc_gemm_bf16_silu_mul_ex(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W_gu16, nvc_bf16* act,
                                         int ldc, int M, int I, int K, int force_bn, int force_splits, void* stream) {
  NVC_REQUIRE(ctx, I > 0 && I % 16 == 0, NVC_ESHAPE, "I must be a multiple of 16");
  return gemm_impl(ctx, A, lda, W_gu16, act, ldc, NVC_DT_BF16, M, 2 * I, K, force_bn, force_splits, NVC_EPI_SILU_MUL,
                   stream, nullptr);
}

extern "C" int nvc_gemm_bf16_silu_mul(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W_gu16, nvc_bf16* act,
                                      int ldc, int M, int I, int K, void* stream) {
  return nvc_gemm_bf16_silu_mul_ex(ctx, A, lda, W_gu16, act, ldc, M, I, K, 0, 0, stream);
}
