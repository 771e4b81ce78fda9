// K4, second generation: varlen causal flash attention (src/models/qwen3.rs:236-277) on tcgen05, built so that TWO
// CTAs are resident per SM and one CTA's serial chain (TMA -> S MMA -> softmax -> P V MMA) runs under the other's:
//   * PERSISTENT grid (2 CTAs per SM): every CTA walks work items (q tile, head, sequence) in longest-first order
//     (the last q tile of a causal sequence has the most kv tiles), keeping its barriers and tensor memory; the
//     producer warp requests the next item's Q / K while the softmax warps still drain the current one, so the
//     ~6 us of per-CTA start-up that a one-item grid pays per ~5 tiles at T = 512 is paid once per CTA;
//   * kv tiles of 64 tokens: Q 32 KB + K ring 2 x 16 KB + V ring 2 x 16 KB + 16 KB of output staging = 112 KB;
//   * the output rows leave through a per-warp staging buffer (RowStage, gemm_epilogue.cuh): one thread owns one row, so
//     direct 16-byte stores touched 32 lines per instruction and the LSU throttled the epilogue (14 % of the stall
//     samples at T = 512); staged, every store instruction covers 4 rows x 128 contiguous bytes;
//   * S is double-buffered in tensor memory; P goes BACK into tensor memory (bf16 pairs over the first 32 columns
//     of the S buffer it came from) and is the A operand of the P V MMA from there: no shared-memory round trip,
//     no proxy fence per tile (measured +8 % on every shape, profiles/r02_prefill_attn_variants_tmem_poly.json);
//   * the output accumulator stays in tensor memory across kv tiles; the softmax threads rescale it with
//     tcgen05.ld / tcgen05.st only when some row's running maximum grew by more than 2^8 (the exponentials are taken
//     against a possibly stale maximum: p <= 256, exact after the final division by l) - with the plain "maximum
//     moved" rule 74 % of the tiles of a random batch paid the 128-column round trip;
//   * the row maximum and the row sum are reduced over four independent chains and the exponentials are issued as
//     one block of independent MUFU operations (the serial fmax / fadd chains were 14 % of the stall samples);
//   * barrier init / TMEM allocation / descriptor prefetch run before the PDL dependency wait.
// A share of the exponentials on the FMA pipe (degree-3 polynomial) was measured and is not used: no gain at 1/4,
// a loss at 1/2 (same profile file) - the kernel is bound by dependent latency per tile, not by the MUFU unit.
// Included by prefill_attn.cu (shares PagedKV, make_tma_heads, fast_exp2).
#pragma once

namespace v2 {

constexpr int kThreads = 192;                  // TMA warp + MMA warp + 4 softmax warps
constexpr int kKT = 64;                        // kv tokens per tile
constexpr int kQBytes = 128 * 128 * 2;         // 32 KB: two [128 rows][64 dims] swizzled halves
constexpr int kHalfKV = kKT * 128;             // 8 KB: [64 tokens][64 dims]
constexpr int kKBytes = 2 * kHalfKV;           // 16 KB: one K or V tile
constexpr int kKStages = 2, kVStages = 2;
constexpr int kOffQ = 0;
constexpr int kOffK = kQBytes;
constexpr int kOffV = kOffK + kKStages * kKBytes;
constexpr int kOffStage = kOffV + kVStages * kKBytes;   // 4 softmax warps x 4 KB row staging for the output
constexpr int kOffBar = kOffStage + 4 * 4096;           // 114688
// q_full, q_empty, k_full[2], k_empty[2], v_full[2], v_empty[2], s_full[2], p_full, o_full, o_free
constexpr int kNumBars = 2 + 2 * kKStages + 2 * kVStages + 2 + 3;
// two CTAs per SM: 2 x (request + 1 KB reserved) <= 228 KB; the dynamic shared memory is requested 1024-byte aligned
// (no static shared memory in this kernel: the pad is 0 in practice), a pad of up to 768 B fits, beyond that the kernel traps
constexpr int kSmemBytes = 115712;
static_assert(kOffBar + kNumBars * 8 + 16 + 768 <= kSmemBytes, "shared memory budget for two CTAs per SM");
constexpr int kTmemCols = 256;                 // S / P (double-buffered): columns [0, 64) and [64, 128); O: [128, 256)
constexpr float kRescaleLog2 = 8.0f;           // rescale O only when a row's maximum grew by more than 2^8

// work item -> (q tile, head, sequence): q tiles in DESCENDING order (longest first), heads fastest (the heads of a GQA
// group run side by side and share their K / V tiles through L2)
// n / d for n < 2^32 by a host-made multiplier: floor(n * ceil(2^(32+s) / d) / 2^(32+s)), s = ceil(log2 d) (exact: Granlund -
// Montgomery); the runtime divisions of the item decomposition were 12 % of the stall samples at T = 512
struct FastDiv {
  unsigned long long mul;
  int shift, d;
  __device__ __forceinline__ int div(int n) const { return (int)(((unsigned long long)(unsigned)n * mul) >> shift); }
};
inline FastDiv make_fastdiv(int d) {
  FastDiv f;
  int s = 0;
  while ((1ll << s) < d) ++s;
  f.shift = 32 + s;
  f.mul = ((1ull << f.shift) + (unsigned long long)d - 1) / (unsigned long long)d;
  f.d = d;
  return f;
}
struct Sched { FastDiv per, nh, group; int QT, num_items; };   // per = nh * B
struct Item { int qt, h, b, kvh; };
__device__ __forceinline__ Item item_of(int it, const Sched& sc) {
  const int q = sc.per.div(it), rem = it - q * sc.per.d;
  Item x;
  x.qt = sc.QT - 1 - q;
  x.b = sc.nh.div(rem);
  x.h = rem - x.b * sc.nh.d;
  x.kvh = sc.group.div(x.h);
  return x;
}

template <bool kPaged>
__global__ void __launch_bounds__(192, 2)
prefill_attn_v2_kernel(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
                       const __grid_constant__ CUtensorMap tma_v, const int32_t* __restrict__ cu_seqlens, const Sched sc,
                       float scale_log2e, nvc_bf16* __restrict__ out, const PagedKV pk) {
  pdl_trigger();
  const int num_items = sc.num_items, nh = sc.nh.d;
  extern __shared__ __align__(1024) uint8_t smem_v2[];
  const uint32_t base = (smem_u32(smem_v2) + 1023u) & ~1023u;   // no-op when the alignment is honoured; pad <= 896 otherwise
  uint8_t* gen = smem_v2 + (base - smem_u32(smem_v2));
  if ((base - smem_u32(smem_v2)) + kOffBar + kNumBars * 8 + 16 > (uint32_t)kSmemBytes) __trap();   // alignment not honoured
  const uint32_t bar0 = base + kOffBar;
  const uint32_t q_full = bar0, q_empty = bar0 + 8u;
  // K and V have their own rings: a K slot is free as soon as S = Q K^T has read it (one tile ahead of the
  // softmax), a V slot only after P V - so K(j+1) is requested a whole tile earlier than with (K, V) stages
  auto k_full = [&](int s) { return bar0 + 8u * (2 + s); };
  auto k_empty = [&](int s) { return bar0 + 8u * (2 + kKStages + s); };
  auto v_full = [&](int s) { return bar0 + 8u * (2 + 2 * kKStages + s); };
  auto v_empty = [&](int s) { return bar0 + 8u * (2 + 2 * kKStages + kVStages + s); };
  auto s_full = [&](int s) { return bar0 + 8u * (2 + 2 * kKStages + 2 * kVStages + s); };
  const uint32_t p_full = bar0 + 8u * (kNumBars - 3), o_full = bar0 + 8u * (kNumBars - 2), o_free = bar0 + 8u * (kNumBars - 1);
  const uint32_t tmem_slot = bar0 + 8u * kNumBars;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen + kOffBar + kNumBars * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // geometry of one item (every role derives the same numbers): q row i of the tile is context position q0 + i and sees
  // keys 0 .. q0 + i; n_tiles = 0 marks a q tile past the end of its sequence (skipped by every role)
  struct Geo { int seq_start, seq_len, kv_len, q0, n_tiles; };
  auto geo_of = [&](const Item& x) {
    Geo g;
    g.seq_start = __ldg(cu_seqlens + x.b);
    g.seq_len = __ldg(cu_seqlens + x.b + 1) - g.seq_start;
    g.kv_len = kPaged ? __ldg(pk.context_lens + x.b) : g.seq_len;
    g.q0 = g.kv_len - g.seq_len + x.qt * 128;
    g.n_tiles = x.qt * 128 >= g.seq_len ? 0 : (min(g.kv_len, g.q0 + 128) + kKT - 1) / kKT;
    return g;
  };

  if (warp == 0) {
    if (lane == 0) {
      ptx::tma_prefetch_desc(&tma_q);
      ptx::tma_prefetch_desc(&tma_k);
      ptx::tma_prefetch_desc(&tma_v);
      ptx::mbar_init(q_full, 1);
      ptx::mbar_init(q_empty, 1);
      for (int s = 0; s < kKStages; ++s) { ptx::mbar_init(k_full(s), 1); ptx::mbar_init(k_empty(s), 1); }
      for (int s = 0; s < kVStages; ++s) { ptx::mbar_init(v_full(s), 1); ptx::mbar_init(v_empty(s), 1); }
      ptx::mbar_init(s_full(0), 1);
      ptx::mbar_init(s_full(1), 1);
      ptx::mbar_init(p_full, 128);
      ptx::mbar_init(o_full, 1);
      ptx::mbar_init(o_free, 128);
      ptx::fence_barrier_init();
    }
    __syncwarp();
  } else if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot_gen;
  const uint32_t tmem_s = tmem, tmem_o = tmem + 128;
  pdl_wait();      // q / k / v (or the cache rows of this step) are written by the predecessor

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (whole warp loops, one lane issues)
    // g counts kv tiles over all items of this CTA: ring slots and barrier phases follow it
    auto load_kv = [&](const Item& x, const Geo& ge, int j, uint32_t g, bool is_v) {
      const int s = is_v ? (int)(g % kVStages) : (int)(g % kKStages);
      const uint32_t lap = is_v ? g / kVStages : g / kKStages;
      ptx::mbar_wait(is_v ? v_empty(s) : k_empty(s), (lap & 1) ^ 1u);
      if (ptx::elect_one()) {
        const uint32_t dst = base + (is_v ? kOffV : kOffK) + s * kKBytes;
        const uint32_t bar = is_v ? v_full(s) : k_full(s);
        const CUtensorMap* map = is_v ? &tma_v : &tma_k;
        const int kvh = x.kvh;
        if constexpr (kPaged) {
          // boxes of box_rows tokens; only those that start inside the context are loaded (the rest of
          // the slot keeps stale bytes: masked in S, zeroed in V by the softmax warps)
          const int rows = pk.box_rows, tok0 = j * kKT;
          const int nbox = min(kKT / rows, (ge.kv_len - tok0 + rows - 1) / rows);
          ptx::mbar_arrive_expect_tx(bar, (uint32_t)(nbox * rows * 256));   // two 64-wide halves
          const int32_t* bt = pk.block_tables + (int64_t)x.b * pk.bt_stride;
          for (int i = 0; i < nbox; ++i) {
            const int tok = tok0 + i * rows;
            const int slot = __ldg(bt + tok / pk.block_size) * pk.block_size + tok % pk.block_size;
            const uint32_t o = (uint32_t)(i * rows * 128);
            ptx::tma_load_3d(dst + o, map, bar, 0, kvh, slot);
            ptx::tma_load_3d(dst + kHalfKV + o, map, bar, 64, kvh, slot);
          }
        } else {
          ptx::mbar_arrive_expect_tx(bar, kKBytes);
          const int row = ge.seq_start + j * kKT;
          ptx::tma_load_3d(dst, map, bar, 0, kvh, row);
          ptx::tma_load_3d(dst + kHalfKV, map, bar, 64, kvh, row);
        }
      }
      __syncwarp();
    };
    uint32_t g = 0, n_done = 0;
    Item xn = item_of(blockIdx.x, sc);               // gridDim.x <= num_items
    Geo gn = geo_of(xn);
    for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
      const Item x = xn;
      const Geo ge = gn;
      if (it + (int)gridDim.x < num_items) {          // the next item's geometry: its loads fly under this item
        xn = item_of(it + gridDim.x, sc);
        gn = geo_of(xn);
      }
      if (ge.n_tiles == 0) continue;
      // Q: the previous item's last S MMA has read it (first item: the barrier is fresh, the wait falls through)
      ptx::mbar_wait(q_empty, (n_done & 1) ^ 1u);
      if (ptx::elect_one()) {
        ptx::mbar_arrive_expect_tx(q_full, kQBytes);
        ptx::tma_load_3d(base + kOffQ, &tma_q, q_full, 0, x.h, ge.seq_start + x.qt * 128);
        ptx::tma_load_3d(base + kOffQ + 16384, &tma_q, q_full, 64, x.h, ge.seq_start + x.qt * 128);
      }
      __syncwarp();
      load_kv(x, ge, 0, g, false);
      for (int j = 0; j < ge.n_tiles; ++j) {
        if (j + 1 < ge.n_tiles) load_kv(x, ge, j + 1, g + j + 1, false);   // K runs one tile ahead of V, as S runs ahead of P V
        load_kv(x, ge, j, g + j, true);
      }
      g += ge.n_tiles;
      ++n_done;
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (whole warp loops, one lane issues)
    constexpr uint32_t idesc_qk = ptx::umma_idesc_bf16(128, kKT, 0, 0);
    constexpr uint32_t idesc_pv = ptx::umma_idesc_bf16(128, 128, 0, 1);   // B = V is MN-major
    constexpr uint32_t hi_k = ptx::umma_desc_hi_sw128(1024);
    const uint32_t q_lo = ptx::umma_desc_lo(base + kOffQ, 16);
    // S(g) = Q K^T into S buffer g & 1 : 8 k-steps of 16 over head_dim (two 64-wide swizzled halves).  The buffer last
    // held S / P of tile g - 2, whose P V MMA was issued earlier by this thread: tcgen05.mma executes in issue order.
    auto issue_s = [&](uint32_t g, bool last_of_item) {
      const int s = (int)(g % kKStages);
      ptx::mbar_wait(k_full(s), (g / kKStages) & 1);
      ptx::tcgen05_fence_after();
      if (ptx::elect_one()) {
        const uint32_t k_lo = ptx::umma_desc_lo(base + kOffK + s * kKBytes, 16);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t qoff = ((ks >> 2) * 16384 + (ks & 3) * 32) >> 4;
          const uint32_t koff = ((ks >> 2) * kHalfKV + (ks & 3) * 32) >> 4;
          ptx::umma_bf16_ss(tmem_s + (g & 1) * kKT, ptx::umma_desc(q_lo + qoff, hi_k), ptx::umma_desc(k_lo + koff, hi_k), idesc_qk,
                            ks != 0);
        }
        ptx::umma_commit(s_full(g & 1));
        ptx::umma_commit(k_empty(s));          // the K slot is free as soon as these MMAs have read it
        if (last_of_item) ptx::umma_commit(q_empty);   // ... and so is Q after the item's last S
      }
      __syncwarp();
    };
    uint32_t g = 0, n_done = 0;
    Item xn = item_of(blockIdx.x, sc);               // gridDim.x <= num_items
    Geo gn = geo_of(xn);
    for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
      const Item x = xn;
      const Geo ge = gn;
      if (it + (int)gridDim.x < num_items) {          // the next item's geometry: its loads fly under this item
        xn = item_of(it + gridDim.x, sc);
        gn = geo_of(xn);
      }
      if (ge.n_tiles == 0) continue;
      const int n = ge.n_tiles;
      ptx::mbar_wait(q_full, n_done & 1);
      issue_s(g, n == 1);
      for (int j = 0; j < n; ++j) {
        // S is double-buffered: S(j+1) is computed while the softmax threads work on S(j)
        if (j + 1 < n) issue_s(g + j + 1, j + 2 == n);
        // P(j) is in tensor memory, O has been rescaled for tile j (both by the softmax threads)
        ptx::mbar_wait(p_full, (g + j) & 1);
        const int vs = (int)((g + j) % kVStages);
        ptx::mbar_wait(v_full(vs), ((g + j) / kVStages) & 1);
        // the first P V of an item overwrites O: the previous item's output rows have been read out of it
        if (j == 0 && n_done > 0) ptx::mbar_wait(o_free, (n_done - 1) & 1);
        ptx::tcgen05_fence_after();
        if (ptx::elect_one()) {
          // O += P V_j: A = P from tensor memory (8 columns per 16 kv tokens); B = V (MN-major: LBO = 8 KB between the
          // two 64-wide head-dim halves, SBO = 1 KB between groups of 8 token rows, 16 token rows = 2048 bytes per step)
          const uint32_t v_lo = ptx::umma_desc_lo(base + kOffV + vs * kKBytes, (uint32_t)kHalfKV);
#pragma unroll
          for (int ks = 0; ks < kKT / 16; ++ks)
            ptx::umma_bf16_ts(tmem_o, tmem_s + ((g + j) & 1) * kKT + ks * 8, ptx::umma_desc(v_lo + ks * (2048 >> 4), hi_k),
                              idesc_pv, (j | ks) != 0);
          ptx::umma_commit(o_full);
          ptx::umma_commit(v_empty(vs));
        }
        __syncwarp();
      }
      g += n;
      ++n_done;
    }
  } else {
    // ------------------------------------------------------------ softmax / correction / output warps
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                 // row in the q tile
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint32_t g = 0;
    Item xn = item_of(blockIdx.x, sc);               // gridDim.x <= num_items
    Geo gn = geo_of(xn);
    for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
      const Item x = xn;
      const Geo ge = gn;
      if (it + (int)gridDim.x < num_items) {          // the next item's geometry: its loads fly under this item
        xn = item_of(it + gridDim.x, sc);
        gn = geo_of(xn);
      }
      if (ge.n_tiles == 0) continue;
      const int n = ge.n_tiles;
      float m = -INFINITY, l = 0.f;                 // m: the maximum the exponentials are taken against (may lag the true one)
      for (int j = 0; j < n; ++j) {
        const uint32_t gi = g + j;
        const uint32_t s_addr = tmem_s + (gi & 1) * kKT + lane_addr;
        ptx::mbar_wait(s_full(gi & 1), (gi >> 1) & 1);
        ptx::tcgen05_fence_after();
        uint32_t sv[kKT];
        ptx::tmem_ld_32x32b_x32(s_addr, sv);
        ptx::tmem_ld_32x32b_x32(s_addr + 32, sv + 32);
        ptx::tmem_ld_wait();
        // diagonal tiles only (the tile holds keys past some row's position): columns past this row's position
        // become -inf, i.e. p = exp2(-inf) = 0; every other tile takes the mask-free path below
        if (j * kKT + kKT - 1 > ge.q0) {
          const int lim = ge.q0 + r - j * kKT;            // last visible column of this row
#pragma unroll
          for (int i = 0; i < kKT; ++i)
            if (i > lim) sv[i] = 0xff800000u;
        }
        // row maximum over four independent chains
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < kKT; i += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(sv[i]));
          mx1 = fmaxf(mx1, __uint_as_float(sv[i + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(sv[i + 2]));
          mx3 = fmaxf(mx3, __uint_as_float(sv[i + 3]));
        }
        const float mt = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * scale_log2e;   // finite on tile 0: key 0 is visible to every row
        // lazy rescale: move the reference maximum (and pay the O round trip) only when some row of this warp outgrew it
        // by more than 2^kRescaleLog2; m = -inf on the first tile, so that one always moves
        const bool move = __any_sync(0xffffffffu, mt > m + kRescaleLog2);
        float alpha = 1.0f;
        if (move) {
          const float m_new = fmaxf(m, mt);
          alpha = fast_exp2(m - m_new);
          m = m_new;
        }
        // exponentials: 64 independent FFMA, then 64 independent MUFU, in place
#pragma unroll
        for (int i = 0; i < kKT; ++i) sv[i] = __float_as_uint(fmaf(__uint_as_float(sv[i]), scale_log2e, -m));
#pragma unroll
        for (int i = 0; i < kKT; ++i) sv[i] = __float_as_uint(fast_exp2(__uint_as_float(sv[i])));
        float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
        uint32_t pk2[kKT / 2];
#pragma unroll
        for (int i = 0; i < kKT; i += 4) {
          const float p0 = __uint_as_float(sv[i]), p1 = __uint_as_float(sv[i + 1]);
          const float p2 = __uint_as_float(sv[i + 2]), p3 = __uint_as_float(sv[i + 3]);
          ps0 += p0; ps1 += p1; ps2 += p2; ps3 += p3;
          pk2[i >> 1] = pack_bf16(p0, p1);
          pk2[(i >> 1) + 1] = pack_bf16(p2, p3);
        }
        l = l * alpha + ((ps0 + ps1) + (ps2 + ps3));
        if (j > 0) {
          // P V(j-1) has completed: O holds the sum over tiles < j.  Waited for on EVERY tile, moved or not: a parity
          // wait is only meaningful within one phase of the barrier, so the softmax threads must observe each o_full
          // phase in turn (skipping them let the final wait take an older phase for the last one).  The MMA was issued a
          // whole softmax pass ago, so the wait falls through.
          ptx::mbar_wait(o_full, (gi - 1) & 1);
        }
        if (j > 0 && move) {
          // O *= alpha (warp-collective; alpha = 1 for rows that did not move)
          ptx::tcgen05_fence_after();
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t v[32];
            ptx::tmem_ld_32x32b_x32(tmem_o + lane_addr + c * 32, v);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            ptx::tmem_st_32x32b_x32(tmem_o + lane_addr + c * 32, v);
          }
        }
        // P -> bf16 pairs over the S columns this thread has just read (its own lane): the A operand of the second MMA
        ptx::tmem_st_32x32b_x32(s_addr, pk2);
        ptx::tmem_st_wait();
        if constexpr (kPaged) {
          // V rows at / past the end of the context (last tile only): zero them (every P column there is 0,
          // but the cache bytes may be NaN) once the slot's TMA has landed; the proxy fence orders these
          // generic stores before the P V MMA.
          if (j * kKT + kKT > ge.kv_len) {
            const int vs = (int)(gi % kVStages);
            ptx::mbar_wait(v_full(vs), (gi / kVStages) & 1);
            if (r < kKT && j * kKT + r >= ge.kv_len) {
              uint8_t* v_row = gen + kOffV + vs * kKBytes + r * 128;
#pragma unroll
              for (int k16 = 0; k16 < 8; ++k16) {
                *reinterpret_cast<uint4*>(v_row + (k16 << 4)) = make_uint4(0, 0, 0, 0);
                *reinterpret_cast<uint4*>(v_row + kHalfKV + (k16 << 4)) = make_uint4(0, 0, 0, 0);
              }
            }
            ptx::fence_proxy_async_smem();
          }
        }
        ptx::tcgen05_fence_before();
        ptx::mbar_arrive(p_full);
      }
      // ---- O / l -> bf16 rows of the output
      ptx::mbar_wait(o_full, (g + n - 1) & 1);
      ptx::tcgen05_fence_after();
      const float inv = 1.0f / l;
      // this warp's 32 rows x 128 head dims leave through its staging buffer: 8 lanes per row, 128 contiguous bytes each
      RowStage st = make_row_stage(gen + kOffStage + quad * 4096, out, 0, NVC_EPI_NONE, (int64_t)nh * 128,
                                   ge.seq_start + x.qt * 128 + quad * 32, ge.seq_start + ge.seq_len, x.h * 128, 128, nh * 128);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x32(tmem_o + lane_addr + c * 32, v);
        ptx::tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * inv;
        st.emit32(lane, 0, NVC_EPI_NONE, f);
      }
      st.finish(lane);
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(o_free);                      // the next item's first P V may overwrite O
      g += n;
    }
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tcgen05_fence_after();
    ptx::tmem_dealloc(tmem, kTmemCols);
  }
}

}  // namespace v2
