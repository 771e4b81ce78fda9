// K4, second generation: the same varlen causal flash attention as prefill_attn_kernel
// (src/models/qwen3.rs:236-277), restructured so that TWO CTAs are resident per SM and one CTA's
// serial chain (TMA -> S MMA -> softmax -> P V MMA) runs under the other's:
//   * kv tiles of 64 tokens: Q 32 KB + 2 x (K 16 KB + V 16 KB) + P 16 KB = 112 KB of shared memory;
//   * the output accumulator stays in TENSOR MEMORY across kv tiles (P V accumulates in place); the
//     softmax threads rescale it by alpha with tcgen05.ld / tcgen05.st only when some row's running
//     maximum moved, so no thread holds 128 fp32 accumulators (236 -> ~100 registers);
//   * S is read from tensor memory once (64 columns stay in registers between the max and exp passes);
//   * the diagonal costs 1.5 kv tiles of 64 instead of one full 128-wide tile;
//   * barrier init / TMEM allocation / descriptor prefetch run before the PDL dependency wait
//     (cu_seqlens / context_lens / block_tables are step inputs, not kernel outputs).
// Included by prefill_attn.cu (shares PagedKV, make_tma_heads, fast_exp2).
#pragma once

namespace v2 {

constexpr int kThreads = 192;                  // TMA warp + MMA warp + 4 softmax warps
constexpr int kKT = 64;                        // kv tokens per tile
constexpr int kQBytes = 128 * 128 * 2;         // 32 KB: two [128 rows][64 dims] swizzled halves
constexpr int kHalfKV = kKT * 128;             // 8 KB: [64 tokens][64 dims]
constexpr int kKBytes = 2 * kHalfKV;           // 16 KB
constexpr int kStageBytes = 2 * kKBytes;       // K tile + V tile
constexpr int kStages = 2;
constexpr int kPBytes = 128 * kKT * 2;         // 16 KB: [128 rows][64 kv tokens], one swizzled half
constexpr int kOffQ = 0;
constexpr int kOffKV = kQBytes;
constexpr int kOffP = kOffKV + kStages * kStageBytes;
constexpr int kOffBar = kOffP + kPBytes;       // 114688
constexpr int kNumBars = 13;                   // q_full, k_full[2], k_empty[2], v_full[2], v_empty[2], s_full[2], p_full, o_full
// two CTAs per SM: 2 x (request + 1 KB reserved) <= 228 KB; the alignment pad is <= 896 B for a 128-byte aligned base
constexpr int kSmemBytes = 115712;
static_assert(kOffBar + kNumBars * 8 + 16 + 896 <= kSmemBytes, "shared memory budget for two CTAs per SM");
constexpr int kTmemCols = 256;                 // S (double-buffered): columns [0, 64) and [64, 128); O: columns [128, 256)

// 2^x for x <= 0 on the FMA pipe (Cody-Waite split + degree-3 minimax polynomial, relative error 9e-5 - below half a
// bf16 ulp, and P is rounded to bf16): the MUFU unit (16 ex2 per clock per SM) is as loaded as the tensor pipe in this
// kernel (512 cycles each per 128 x 64 tile), so a fraction of the exponentials is computed here instead.
__device__ __forceinline__ float exp2_poly(float x) {
  x = fmaxf(x, -126.0f);
  const float xr = __fadd_rd(x, 12582912.0f);          // 1.5 * 2^23: floor(x) lands in the low mantissa bits
  const float f = x - (xr - 12582912.0f);               // [0, 1)
  float p = fmaf(f, 0.0771190897f, 0.227564394f);
  p = fmaf(p, f, 0.695146143f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(xr) << 23));
}

// kPT: P goes back into tensor memory (bf16 pairs over the first 32 columns of the S buffer it came from) and is the
// A operand of the P V MMA from there - no shared-memory round trip, no proxy fence per tile.
// kPoly = n > 0: every n-th exponential of a row on the FMA pipe (exp2_poly), the others on the MUFU unit.
template <bool kPaged, bool kPT, int kPoly>
__global__ void __launch_bounds__(192, 2)
prefill_attn_v2_kernel(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
                       const __grid_constant__ CUtensorMap tma_v, const int32_t* __restrict__ cu_seqlens,
                       int nh, int group, float scale_log2e, nvc_bf16* __restrict__ out, const PagedKV pk) {
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  pdl_trigger();
  const int seq_start = __ldg(cu_seqlens + b);
  const int seq_len = __ldg(cu_seqlens + b + 1) - seq_start;
  if (qt * 128 >= seq_len) return;            // whole CTA exits before any allocation
  const int kvh = h / group;
  // q row i of the tile is context position off + qt*128 + i and sees keys 0..that position
  const int kv_len = kPaged ? __ldg(pk.context_lens + b) : seq_len;
  const int off = kPaged ? kv_len - seq_len : 0;
  const int q0 = off + qt * 128;               // context position of row 0
  const int n_tiles = (min(kv_len, q0 + 128) + kKT - 1) / kKT;

  extern __shared__ __align__(1024) uint8_t smem_v2[];
  const uint32_t base = (smem_u32(smem_v2) + 1023u) & ~1023u;   // no-op when the alignment is honoured; pad <= 896 otherwise
  uint8_t* gen = smem_v2 + (base - smem_u32(smem_v2));
  const uint32_t bar0 = base + kOffBar;
  const uint32_t q_full = bar0;
  // K and V have their own rings: a K slot is free as soon as S = Q K^T has read it (one tile ahead of the
  // softmax), a V slot only after P V - so K(j+1) is requested a whole tile earlier than with (K, V) stages
  auto k_full = [&](int s) { return bar0 + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bar0 + 8u * (3 + s); };
  auto v_full = [&](int s) { return bar0 + 8u * (5 + s); };
  auto v_empty = [&](int s) { return bar0 + 8u * (7 + s); };
  auto s_full = [&](int s) { return bar0 + 8u * (9 + s); };
  const uint32_t p_full = bar0 + 8u * 11, o_full = bar0 + 8u * 12;
  const uint32_t tmem_slot = bar0 + 8u * kNumBars;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen + kOffBar + kNumBars * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // one K or V tile (is_v) of kv tile j into its ring slot
  auto load_tile = [&](int j, bool is_v) {
    const int s = j & 1;
    ptx::mbar_wait(is_v ? v_empty(s) : k_empty(s), ((j >> 1) & 1) ^ 1u);
    if (ptx::elect_one()) {
      const uint32_t dst = base + kOffKV + s * kStageBytes + (is_v ? kKBytes : 0);
      const uint32_t bar = is_v ? v_full(s) : k_full(s);
      const CUtensorMap* map = is_v ? &tma_v : &tma_k;
      if constexpr (kPaged) {
        // boxes of box_rows tokens; only those that start inside the context are loaded (the rest of
        // the slot keeps stale bytes: masked in S, zeroed in V by the softmax warps)
        const int rows = pk.box_rows, tok0 = j * kKT;
        const int nbox = min(kKT / rows, (kv_len - tok0 + rows - 1) / rows);
        ptx::mbar_arrive_expect_tx(bar, (uint32_t)(nbox * rows * 256));   // two 64-wide halves
        const int32_t* bt = pk.block_tables + (int64_t)b * pk.bt_stride;
        for (int i = 0; i < nbox; ++i) {
          const int tok = tok0 + i * rows;
          const int slot = __ldg(bt + tok / pk.block_size) * pk.block_size + tok % pk.block_size;
          const uint32_t o = (uint32_t)(i * rows * 128);
          ptx::tma_load_3d(dst + o, map, bar, 0, kvh, slot);
          ptx::tma_load_3d(dst + kHalfKV + o, map, bar, 64, kvh, slot);
        }
      } else {
        ptx::mbar_arrive_expect_tx(bar, kKBytes);
        const int row = seq_start + j * kKT;
        ptx::tma_load_3d(dst, map, bar, 0, kvh, row);
        ptx::tma_load_3d(dst + kHalfKV, map, bar, 64, kvh, row);
      }
    }
    __syncwarp();
  };
  if (warp == 0) {
    // The producer warp initialises the barriers itself and requests Q and K(0) right away, while warp 1
    // allocates tensor memory: the first loads do not wait for the CTA-wide sync.
    if (lane == 0) {
      ptx::tma_prefetch_desc(&tma_q);
      ptx::tma_prefetch_desc(&tma_k);
      ptx::tma_prefetch_desc(&tma_v);
      ptx::mbar_init(q_full, 1);
      for (int s = 0; s < kStages; ++s) {
        ptx::mbar_init(k_full(s), 1); ptx::mbar_init(k_empty(s), 1);
        ptx::mbar_init(v_full(s), 1); ptx::mbar_init(v_empty(s), 1);
      }
      ptx::mbar_init(s_full(0), 1);
      ptx::mbar_init(s_full(1), 1);
      ptx::mbar_init(p_full, 128);
      ptx::mbar_init(o_full, 1);
      ptx::fence_barrier_init();
    }
    __syncwarp();
    pdl_wait();    // q / k / v (or the cache rows of this step) are written by the predecessor
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(q_full, kQBytes);
      ptx::tma_load_3d(base + kOffQ, &tma_q, q_full, 0, h, seq_start + qt * 128);
      ptx::tma_load_3d(base + kOffQ + 16384, &tma_q, q_full, 64, h, seq_start + qt * 128);
    }
    __syncwarp();
    load_tile(0, false);
  } else if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot_gen;
  const uint32_t tmem_s = tmem, tmem_o = tmem + 128;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (whole warp loops, one lane issues)
    for (int j = 0; j < n_tiles; ++j) {
      if (j + 1 < n_tiles) load_tile(j + 1, false);   // K runs one tile ahead of V, as S = Q K^T runs ahead of P V
      load_tile(j, true);
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (whole warp loops, one lane issues)
    constexpr uint32_t idesc_qk = ptx::umma_idesc_bf16(128, kKT, 0, 0);
    constexpr uint32_t idesc_pv = ptx::umma_idesc_bf16(128, 128, 0, 1);   // B = V is MN-major
    constexpr uint32_t hi_k = ptx::umma_desc_hi_sw128(1024);
    const uint32_t q_lo = ptx::umma_desc_lo(base + kOffQ, 16);
    const uint32_t p_lo = ptx::umma_desc_lo(base + kOffP, 16);
    auto issue_s = [&](int j) {   // S = Q K_j^T : 8 k-steps of 16 over head_dim (two 64-wide swizzled halves)
      const int s = j & 1;
      ptx::mbar_wait(k_full(s), (j >> 1) & 1);
      ptx::tcgen05_fence_after();
      if (ptx::elect_one()) {
        const uint32_t k_lo = ptx::umma_desc_lo(base + kOffKV + s * kStageBytes, 16);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t qoff = ((ks >> 2) * 16384 + (ks & 3) * 32) >> 4;
          const uint32_t koff = ((ks >> 2) * kHalfKV + (ks & 3) * 32) >> 4;
          ptx::umma_bf16_ss(tmem_s + s * kKT, ptx::umma_desc(q_lo + qoff, hi_k), ptx::umma_desc(k_lo + koff, hi_k), idesc_qk,
                            ks != 0);
        }
        ptx::umma_commit(s_full(s));
        ptx::umma_commit(k_empty(s));          // the K slot is free as soon as these MMAs have read it
      }
      __syncwarp();
    };
    ptx::mbar_wait(q_full, 0);
    issue_s(0);
    for (int j = 0; j < n_tiles; ++j) {
      // S is double-buffered: S(j+1) goes into the buffer S(j-1) was read from (p_full(j-1) was awaited in the
      // previous iteration), so the tensor pipe computes it while the softmax threads work on S(j)
      if (j + 1 < n_tiles) issue_s(j + 1);
      // P(j) is in shared memory, S(j) has been read, O has been rescaled for tile j (all by the softmax threads)
      ptx::mbar_wait(p_full, j & 1);
      ptx::mbar_wait(v_full(j & 1), (j >> 1) & 1);
      ptx::tcgen05_fence_after();
      if (ptx::elect_one()) {
        // O += P V_j: A = P (K-major, 16 kv tokens per step); B = V (MN-major: LBO = 8 KB between the two
        // 64-wide head-dim halves, SBO = 1 KB between groups of 8 token rows, 16 token rows = 2048 bytes per step)
        const uint32_t v_lo = ptx::umma_desc_lo(base + kOffKV + (j & 1) * kStageBytes + kKBytes, (uint32_t)kHalfKV);
#pragma unroll
        for (int ks = 0; ks < kKT / 16; ++ks) {
          if constexpr (kPT)     // P(j) sits in the first 32 columns of S buffer j & 1: 8 columns per 16 kv tokens
            ptx::umma_bf16_ts(tmem_o, tmem_s + (j & 1) * kKT + ks * 8, ptx::umma_desc(v_lo + ks * (2048 >> 4), hi_k), idesc_pv,
                              (j | ks) != 0);
          else
            ptx::umma_bf16_ss(tmem_o, ptx::umma_desc(p_lo + ((ks * 32) >> 4), hi_k), ptx::umma_desc(v_lo + ks * (2048 >> 4), hi_k),
                              idesc_pv, (j | ks) != 0);
        }
        ptx::umma_commit(o_full);
        ptx::umma_commit(v_empty(j & 1));
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------ softmax / correction / output warps
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                 // row in the q tile
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    float m = -INFINITY, l = 0.f;
    uint8_t* p_row = gen + kOffP + r * 128;
    for (int j = 0; j < n_tiles; ++j) {
      ptx::mbar_wait(s_full(j & 1), (j >> 1) & 1);
      ptx::tcgen05_fence_after();
      uint32_t sv[kKT];
      ptx::tmem_ld_32x32b_x32(tmem_s + (j & 1) * kKT + lane_addr, sv);
      ptx::tmem_ld_32x32b_x32(tmem_s + (j & 1) * kKT + lane_addr + 32, sv + 32);
      ptx::tmem_ld_wait();
      // diagonal tiles only (the tile holds keys past some row's position): columns past this row's position
      // become -inf, i.e. p = exp2(-inf) = 0; every other tile takes the mask-free path below
      if (j * kKT + kKT - 1 > q0) {
        const int lim = q0 + r - j * kKT;               // last visible column of this row
#pragma unroll
        for (int i = 0; i < kKT; ++i)
          if (i > lim) sv[i] = 0xff800000u;
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < kKT; ++i) mx = fmaxf(mx, __uint_as_float(sv[i]));
      const float m_new = fmaxf(m, mx * scale_log2e);     // finite from tile 0 on: key 0 is visible to every row
      const float alpha = fast_exp2(m - m_new);
      float psum = 0.f;
      uint32_t pk2[kKT / 2];
#pragma unroll
      for (int i = 0; i < kKT; i += 2) {
        const float x0 = __uint_as_float(sv[i]) * scale_log2e - m_new, x1 = __uint_as_float(sv[i + 1]) * scale_log2e - m_new;
        const float p0 = (kPoly > 0 && i % kPoly == kPoly - 1) ? exp2_poly(x0) : fast_exp2(x0);
        const float p1 = (kPoly > 0 && (i + 1) % kPoly == kPoly - 1) ? exp2_poly(x1) : fast_exp2(x1);
        psum += p0 + p1;
        pk2[i >> 1] = pack_bf16(p0, p1);
      }
      l = l * alpha + psum;
      m = m_new;
      if (j > 0) {
        // P V(j-1) has completed: the P buffer is free and O holds the sum over tiles < j
        ptx::mbar_wait(o_full, (j - 1) & 1);
        ptx::tcgen05_fence_after();
        if (__any_sync(0xffffffffu, alpha != 1.0f)) {      // some row's maximum moved: O *= alpha (warp-collective)
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t v[32];
            ptx::tmem_ld_32x32b_x32(tmem_o + lane_addr + c * 32, v);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            ptx::tmem_st_32x32b_x32(tmem_o + lane_addr + c * 32, v);
          }
          ptx::tmem_st_wait();
        }
      }
      if constexpr (kPT) {
        // P -> bf16 pairs over the S columns this thread has just read (its own lane): the A operand of the second MMA
        ptx::tmem_st_32x32b_x32(tmem_s + (j & 1) * kKT + lane_addr, pk2);
        ptx::tmem_st_wait();
      } else {
        // P -> bf16, swizzled store (16-byte chunk ^= row % 8): the A operand of the second MMA
#pragma unroll
        for (int k8 = 0; k8 < kKT / 8; ++k8)
          *reinterpret_cast<uint4*>(p_row + ((k8 ^ (r & 7)) << 4)) =
              make_uint4(pk2[4 * k8], pk2[4 * k8 + 1], pk2[4 * k8 + 2], pk2[4 * k8 + 3]);
      }
      if constexpr (kPaged) {
        // V rows at / past the end of the context (last tile only): zero them (every P column there is 0,
        // but the cache bytes may be NaN) once the slot's TMA has landed; the proxy fence below orders these
        // generic stores before the P V MMA.
        if (j * kKT + kKT > kv_len) ptx::mbar_wait(v_full(j & 1), (j >> 1) & 1);
        if (r < kKT && j * kKT + r >= kv_len) {
          uint8_t* v_row = gen + kOffKV + (j & 1) * kStageBytes + kKBytes + r * 128;
#pragma unroll
          for (int k16 = 0; k16 < 8; ++k16) {
            *reinterpret_cast<uint4*>(v_row + (k16 << 4)) = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(v_row + kHalfKV + (k16 << 4)) = make_uint4(0, 0, 0, 0);
          }
        }
      }
      if (!kPT || kPaged) ptx::fence_proxy_async_smem();    // generic shared-memory stores (P, zeroed V rows) before the MMA reads them
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(p_full);
    }
    // ---- O / l -> bf16 rows of the output
    ptx::mbar_wait(o_full, (n_tiles - 1) & 1);
    ptx::tcgen05_fence_after();
    const int row_in_seq = qt * 128 + r;
    const float inv = 1.0f / l;
    uint4* dst = reinterpret_cast<uint4*>(out + ((int64_t)(seq_start + row_in_seq) * nh + h) * 128);
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t v[32];
      ptx::tmem_ld_32x32b_x32(tmem_o + lane_addr + c * 32, v);
      ptx::tmem_ld_wait();
      if (row_in_seq < seq_len) {
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) {
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = __uint_as_float(v[k8 * 8 + i]) * inv;
          dst[c * 4 + k8] = pack8(o);
        }
      }
    }
    ptx::tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    ptx::tcgen05_fence_after();
    ptx::tmem_dealloc(tmem, kTmemCols);
  }
}

}  // namespace v2
