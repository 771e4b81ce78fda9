// K4: varlen causal flash attention for prefill on tcgen05 / TMEM.
// Replaces the dense attention of Qwen3Attention::forward (src/models/qwen3.rs:236-277):
// GQA expand (:240-256), QK^T*scale (:258-259), host-built -1e9 mask + H2D (:260-270),
// f32 softmax (:271-273), P*V and head merge (:275-277).  No mask tensor, no materialised
// K/V copies, no [B,nh,T,T] scores.
//
// CTA = (128-row q tile, q head, sequence).  Roles:
//   warp 0   : TMA producer (Q once; K and V tiles of 128 tokens, 2-stage ring), 3-D tensor
//              maps {head_dim, head, token} so q/k/v are read in place from the fused qkv rows
//   warp 1   : TMEM allocator + single-thread tcgen05.mma issuer
//              S = Q K^T  (A, B K-major, 128B swizzle) and O_j = P V_j (B = V is MN-major)
//   warps 2-5: one thread per q row: tcgen05.ld S, online softmax (log2 domain, fp32),
//              P -> bf16 -> swizzled smem (A operand of the second MMA), then
//              acc = acc * alpha + O_j with the output accumulator in registers
// FLOPs per (sequence, head): 4 * 128 * T^2 / 2 (causal).
//
// kPaged = true (prefix caching): the q rows of sequence b are its LAST q_len
// positions of a context of context_lens[b] tokens; K / V tiles are read from the paged cache
// through block_tables (one TMA box per block, or per 128 tokens when block_size >= 128), after
// qk_norm_rope_store has scattered this step's k / v there.  Cache rows past the context may
// be uninitialised memory: their S columns are masked by the causal rule and their V rows are
// zeroed in shared memory before the P V product (0 * NaN would poison the row).
#include "common.cuh"
#include "ptx.cuh"
#include "internal.h"

int g_nvc_debug[8] = {0, 0, 0, 0, 0, 0, 0, 0};
extern "C" int nvc_debug_set(int key, int value) {
  if (key < 0 || key >= 8) return NVC_EINVAL;
  g_nvc_debug[key] = value;
  return NVC_OK;
}

namespace {

constexpr int kThreads = 192;
constexpr int kQBytes = 128 * 128 * 2;        // 32 KB: two [128][64] swizzled halves
constexpr int kKVStageBytes = 2 * kQBytes;    // K tile + V tile
constexpr int kOffQ = 0;
constexpr int kOffKV = kQBytes;
constexpr int kOffP = kOffKV + 2 * kKVStageBytes;
constexpr int kOffBar = kOffP + kQBytes;
constexpr int kNumBars = 9;                   // q_full, kv_full[2], kv_empty[2], s_full, p_full, o_full, o_empty
constexpr int kSmemBytes = kOffBar + kNumBars * 8 + 16 + 1024;
constexpr int kTmemCols = 256;                // S: cols [0,128), O: cols [128,256)

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct PagedKV {                              // kPaged only
  const int32_t* block_tables;                // [B, bt_stride], -1 padded
  const int32_t* context_lens;                // [B]
  int bt_stride;
  int block_size;
  int box_rows;                               // token rows per K / V TMA box = min(128, block_size)
};

template <bool kPaged>
__global__ void __launch_bounds__(kThreads, 1)
prefill_attn_kernel(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
                    const __grid_constant__ CUtensorMap tma_v, const int32_t* __restrict__ cu_seqlens,
                    int nh, int group, float scale_log2e, nvc_bf16* __restrict__ out, int v_lbo, int v_sbo,
                    const PagedKV pk) {
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  pdl_trigger();
  pdl_wait();      // cu_seqlens is read right away, so there is no prologue to overlap here
  const int seq_start = __ldg(cu_seqlens + b);
  const int seq_len = __ldg(cu_seqlens + b + 1) - seq_start;
  if (qt * 128 >= seq_len) return;            // whole CTA exits before any allocation
  const int kvh = h / group;
  // kPaged: q row i of the tile is context position off + qt*128 + i and sees keys 0..that position
  const int kv_len = kPaged ? __ldg(pk.context_lens + b) : seq_len;
  const int off = kPaged ? kv_len - seq_len : 0;
  const int n_tiles = kPaged ? (min(kv_len, off + qt * 128 + 128) + 127) / 128
                             : qt + 1;        // causal: kv tiles 0..qt

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar0 = base + kOffBar;
  const uint32_t q_full = bar0;
  auto kv_full = [&](int s) { return bar0 + 8u * (1 + s); };
  auto kv_empty = [&](int s) { return bar0 + 8u * (3 + s); };
  const uint32_t s_full = bar0 + 8u * 5, p_full = bar0 + 8u * 6, o_full = bar0 + 8u * 7, o_empty = bar0 + 8u * 8;
  const uint32_t tmem_slot = bar0 + 8u * kNumBars;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen + kOffBar + kNumBars * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tma_q);
    ptx::tma_prefetch_desc(&tma_k);
    ptx::tma_prefetch_desc(&tma_v);
  }
  if (warp == 1) {
    if (lane == 0) {
      ptx::mbar_init(q_full, 1);
      for (int s = 0; s < 2; ++s) { ptx::mbar_init(kv_full(s), 1); ptx::mbar_init(kv_empty(s), 1); }
      ptx::mbar_init(s_full, 1);
      ptx::mbar_init(p_full, 128);
      ptx::mbar_init(o_full, 1);
      ptx::mbar_init(o_empty, 128);
      ptx::fence_barrier_init();
    }
    __syncwarp();
    ptx::tmem_alloc(tmem_slot, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot_gen;
  const uint32_t tmem_s = tmem, tmem_o = tmem + 128;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (whole warp loops, one lane issues)
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(q_full, kQBytes);
      ptx::tma_load_3d(base + kOffQ, &tma_q, q_full, 0, h, seq_start + qt * 128);
      ptx::tma_load_3d(base + kOffQ + 16384, &tma_q, q_full, 64, h, seq_start + qt * 128);
    }
    __syncwarp();
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j & 1;
      ptx::mbar_wait(kv_empty(s), ((j >> 1) & 1) ^ 1u);
      if (ptx::elect_one()) {
        const uint32_t kdst = base + kOffKV + s * kKVStageBytes, vdst = kdst + kQBytes;
        if constexpr (kPaged) {
          // boxes of box_rows tokens; only those that start inside the context are loaded (the
          // rest of the stage keeps stale bytes: masked in S, zeroed in V by the softmax warps)
          const int rows = pk.box_rows, tok0 = j * 128;
          const int nbox = min(128 / rows, (kv_len - tok0 + rows - 1) / rows);
          ptx::mbar_arrive_expect_tx(kv_full(s), (uint32_t)(nbox * rows * 512));   // K + V, two 64-wide halves each
          const int32_t* bt = pk.block_tables + (int64_t)b * pk.bt_stride;
          for (int i = 0; i < nbox; ++i) {
            const int tok = tok0 + i * rows;
            const int slot = __ldg(bt + tok / pk.block_size) * pk.block_size + tok % pk.block_size;
            const uint32_t o = (uint32_t)(i * rows * 128);
            ptx::tma_load_3d(kdst + o, &tma_k, kv_full(s), 0, kvh, slot);
            ptx::tma_load_3d(kdst + 16384 + o, &tma_k, kv_full(s), 64, kvh, slot);
            ptx::tma_load_3d(vdst + o, &tma_v, kv_full(s), 0, kvh, slot);
            ptx::tma_load_3d(vdst + 16384 + o, &tma_v, kv_full(s), 64, kvh, slot);
          }
        } else {
          ptx::mbar_arrive_expect_tx(kv_full(s), kKVStageBytes);
          const int row = seq_start + j * 128;
          ptx::tma_load_3d(kdst, &tma_k, kv_full(s), 0, kvh, row);
          ptx::tma_load_3d(kdst + 16384, &tma_k, kv_full(s), 64, kvh, row);
          ptx::tma_load_3d(vdst, &tma_v, kv_full(s), 0, kvh, row);
          ptx::tma_load_3d(vdst + 16384, &tma_v, kv_full(s), 64, kvh, row);
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (whole warp loops, one lane issues)
    constexpr uint32_t idesc_qk = ptx::umma_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t idesc_pv = ptx::umma_idesc_bf16(128, 128, 0, 1);   // B = V is MN-major
    constexpr uint32_t hi_k = ptx::umma_desc_hi_sw128(1024);
    const uint32_t hi_v = ptx::umma_desc_hi_sw128((uint32_t)v_sbo);
    const uint32_t q_lo = ptx::umma_desc_lo(base + kOffQ, 16);
    const uint32_t p_lo = ptx::umma_desc_lo(base + kOffP, 16);
    auto issue_s = [&](int j) {   // S = Q K_j^T : 8 k-steps of 16 over head_dim (two 64-wide swizzled halves)
      const int s = j & 1;
      ptx::mbar_wait(kv_full(s), (j >> 1) & 1);
      ptx::tcgen05_fence_after();
      if (ptx::elect_one()) {
        const uint32_t k_lo = ptx::umma_desc_lo(base + kOffKV + s * kKVStageBytes, 16);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t off = ((ks >> 2) * 16384 + (ks & 3) * 32) >> 4;
          ptx::umma_bf16_ss(tmem_s, ptx::umma_desc(q_lo + off, hi_k), ptx::umma_desc(k_lo + off, hi_k), idesc_qk, ks != 0);
        }
        ptx::umma_commit(s_full);
      }
      __syncwarp();
    };
    ptx::mbar_wait(q_full, 0);
    issue_s(0);
    for (int j = 0; j < n_tiles; ++j) {
      ptx::mbar_wait(p_full, j & 1);           // P(j) is in smem and S(j) has been consumed
      ptx::tcgen05_fence_after();
      if (j + 1 < n_tiles) issue_s(j + 1);
      if (j > 0) {
        ptx::mbar_wait(o_empty, (j - 1) & 1);  // softmax threads are done with O(j-1)
        ptx::tcgen05_fence_after();
      }
      if (ptx::elect_one()) {
        // O_j = P V_j: A = P (K-major, 16 kv tokens per step); B = V (MN-major: 16 token rows = 2048 bytes per step)
        const uint32_t v_lo = ptx::umma_desc_lo(base + kOffKV + (j & 1) * kKVStageBytes + kQBytes, (uint32_t)v_lbo);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t poff = ((ks >> 2) * 16384 + (ks & 3) * 32) >> 4;
          ptx::umma_bf16_ss(tmem_o, ptx::umma_desc(p_lo + poff, hi_k), ptx::umma_desc(v_lo + ks * (2048 >> 4), hi_v),
                            idesc_pv, ks != 0);
        }
        ptx::umma_commit(o_full);
        ptx::umma_commit(kv_empty(j & 1));
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------ softmax / output warps
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                 // row in the q tile
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    float acc[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) acc[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    uint8_t* p_row = gen + kOffP + r * 128;
    for (int j = 0; j < n_tiles; ++j) {
      ptx::mbar_wait(s_full, j & 1);
      ptx::tcgen05_fence_after();
      // diag: the tile holds keys past some row's position; lim = last visible column of this row
      const bool diag = kPaged ? (j * 128 + 127 > off + qt * 128) : (j == qt);
      const int lim = kPaged ? off + qt * 128 + r - j * 128 : r;
      // pass 1: row max
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x32(tmem_s + lane_addr + c * 32, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float x = __uint_as_float(v[i]);
          if (!diag || (c * 32 + i) <= lim) mx = fmaxf(mx, x);
        }
      }
      const float m_new = fmaxf(m, mx * scale_log2e);
      const float alpha = fast_exp2(m - m_new);
      float psum = 0.f;
      // pass 2: p = exp2(s*scale - m), bf16, swizzled store (chunk ^= row % 8)
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x32(tmem_s + lane_addr + c * 32, v);
        ptx::tmem_ld_wait();
        float p[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float x = __uint_as_float(v[i]) * scale_log2e - m_new;
          p[i] = (!diag || (c * 32 + i) <= lim) ? fast_exp2(x) : 0.f;
          psum += p[i];
        }
        // columns [c*32, c*32+32) = half (c>>1), 16-byte chunks ((c&1)*4 .. +4) of this row
        uint8_t* half_base = p_row + (c >> 1) * 16384;
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const int chunk = (c & 1) * 4 + k4;
          *reinterpret_cast<uint4*>(half_base + ((chunk ^ (r & 7)) << 4)) = pack8(p + 8 * k4);
        }
      }
      l = l * alpha + psum;
      m = m_new;
      if constexpr (kPaged) {
        // V rows at / past the end of the context: zero them (every P column there is 0, but the
        // cache bytes may be NaN).  s_full(j) implies the stage's TMA has landed; the proxy fence
        // below orders these generic stores before the P V MMA.
        if (j * 128 + r >= kv_len) {
          uint8_t* v_row = gen + kOffKV + (j & 1) * kKVStageBytes + kQBytes + r * 128;
#pragma unroll
          for (int k16 = 0; k16 < 8; ++k16) {
            *reinterpret_cast<uint4*>(v_row + (k16 << 4)) = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(v_row + 16384 + (k16 << 4)) = make_uint4(0, 0, 0, 0);
          }
        }
      }
      ptx::fence_proxy_async_smem();
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(p_full);
      // O_j = P V_j
      ptx::mbar_wait(o_full, j & 1);
      ptx::tcgen05_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x32(tmem_o + lane_addr + c * 32, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[c * 32 + i] = acc[c * 32 + i] * alpha + __uint_as_float(v[i]);
      }
      ptx::tcgen05_fence_before();
      ptx::mbar_arrive(o_empty);
    }
    const int row_in_seq = qt * 128 + r;
    if (row_in_seq < seq_len) {
      const float inv = 1.0f / l;
      uint4* dst = reinterpret_cast<uint4*>(out + ((int64_t)(seq_start + row_in_seq) * nh + h) * 128);
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = acc[c * 8 + i] * inv;
        dst[c] = pack8(o);
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

int make_tma_heads(nvc_ctx* ctx, CUtensorMap* map, const void* base, int heads, int64_t tokens, int ld, int box_rows = 128) {
  cuuint64_t gdim[3] = {128, (cuuint64_t)heads, (cuuint64_t)tokens};
  cuuint64_t gstr[2] = {128 * 2, (cuuint64_t)ld * 2};
  cuuint32_t box[3] = {64, 1, (cuuint32_t)box_rows};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstr, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    nvc_set_error(ctx, "cuTensorMapEncodeTiled (attention) failed with CUresult " + std::to_string((int)r));
    return NVC_EINVAL;
  }
  return NVC_OK;
}

#include "gemm_epilogue.cuh"
#include "prefill_attn_v2.cuh"

// persistent grid of the second-generation kernel: two CTAs per SM, or one CTA per work item when there are fewer
template <bool kPaged>
cudaError_t launch_v2(nvc_ctx* ctx, int QT, int nh, int kv, int B, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tk,
                      const CUtensorMap& tv, const int32_t* cu_seqlens, float scale_log2e, nvc_bf16* out, const PagedKV& pk) {
  const long long items = (long long)QT * nh * B;
  if (items > 0x7fffffffLL) return cudaErrorInvalidValue;
  const int grid = (int)(items < 2LL * ctx->num_sms ? items : 2LL * ctx->num_sms);
  v2::Sched sc{v2::make_fastdiv(nh * B), v2::make_fastdiv(nh), v2::make_fastdiv(nh / kv), QT, (int)items};
  return nvc_launch(ctx, v2::prefill_attn_v2_kernel<kPaged>, dim3(grid), dim3(v2::kThreads), v2::kSmemBytes, st, tq, tk, tv,
                    cu_seqlens, sc, scale_log2e, out, pk);
}

bool use_v1() {
  static const bool env = getenv("NVC_PREFILL_V1") && atoi(getenv("NVC_PREFILL_V1")) != 0;
  return env || g_nvc_debug[3] == 1;
}

}  // namespace

extern "C" int nvc_prefill_attn_init(nvc_ctx* ctx) {
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(v2::prefill_attn_v2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, v2::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(v2::prefill_attn_v2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, v2::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(prefill_attn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(prefill_attn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  return NVC_OK;
}

extern "C" int nvc_prefill_attn_varlen(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k, int ldk,
                                       const nvc_bf16* v, int ldv, const int32_t* cu_seqlens, int B,
                                       int total_tokens, int max_seqlen, int nh, int kv, int hd, float scale,
                                       nvc_bf16* out, void* stream) {
  NVC_REQUIRE(ctx, ctx && q && k && v && cu_seqlens && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B >= 0 && total_tokens >= 0 && max_seqlen >= 0 && nh > 0 && kv > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, hd == 128, NVC_ESHAPE, "head_dim must be 128");
  NVC_REQUIRE(ctx, nh % kv == 0, NVC_ESHAPE, "nh must be a multiple of kv");
  NVC_REQUIRE(ctx, ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldq >= nh * hd && ldk >= kv * hd && ldv >= kv * hd,
              NVC_ESHAPE, "row strides must cover the heads and be multiples of 8");
  NVC_REQUIRE(ctx, ((uintptr_t)q % 16) == 0 && ((uintptr_t)k % 16) == 0 && ((uintptr_t)v % 16) == 0, NVC_EINVAL,
              "pointers must be 16-byte aligned");
  NVC_REQUIRE(ctx, ctx->encode_tiled != nullptr, NVC_EDEVICE, "cuTensorMapEncodeTiled unavailable");
  if (B == 0 || total_tokens == 0 || max_seqlen == 0) return NVC_OK;
  CUtensorMap tq, tk, tv;
  const bool v1 = use_v1();
  int rc = make_tma_heads(ctx, &tq, q, nh, total_tokens, ldq);
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tk, k, kv, total_tokens, ldk, v1 ? 128 : v2::kKT);
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tv, v, kv, total_tokens, ldv, v1 ? 128 : v2::kKT);
  if (rc) return rc;
  if (!v1) {
    const int QT = (max_seqlen + 127) / 128;
    NVC_CUDA_CHECK(ctx, launch_v2<false>(ctx, QT, nh, kv, B, nvc_stream(stream), tq, tk, tv, cu_seqlens,
                                         scale * 1.4426950408889634f, out, PagedKV{}));
    NVC_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->launches++;
    return NVC_OK;
  }
  // V (B operand, MN-major, 128B swizzle): LBO = stride between the two 64-wide head-dim
  // halves (16 KB), SBO = stride between groups of 8 token rows (1 KB).
  int v_lbo = 16384, v_sbo = 1024;
  if (g_nvc_debug[0] == 1) { v_lbo = 1024; v_sbo = 16384; }
  dim3 grid((max_seqlen + 127) / 128, nh, B);
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, prefill_attn_kernel<false>, grid, dim3(kThreads), kSmemBytes, nvc_stream(stream), tq, tk, tv,
                                 cu_seqlens, nh, nh / kv, scale * 1.4426950408889634f, out, v_lbo, v_sbo, PagedKV{}));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

extern "C" int nvc_prefill_attn_paged(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                      const nvc_bf16* v_cache, const int32_t* block_tables, int bt_stride,
                                      const int32_t* context_lens, const int32_t* cu_seqlens_q, int B,
                                      int total_q_tokens, int max_seqlen_q, int nh, int kv, int hd, int block_size,
                                      int num_blocks, float scale, nvc_bf16* out, void* stream) {
  NVC_REQUIRE(ctx, ctx && q && k_cache && v_cache && block_tables && context_lens && cu_seqlens_q && out, NVC_EINVAL,
              "null pointer");
  NVC_REQUIRE(ctx, B >= 0 && total_q_tokens >= 0 && max_seqlen_q >= 0 && nh > 0 && kv > 0 && bt_stride > 0 &&
                       num_blocks > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, hd == 128, NVC_ESHAPE, "head_dim must be 128");
  NVC_REQUIRE(ctx, nh % kv == 0, NVC_ESHAPE, "nh must be a multiple of kv");
  NVC_REQUIRE(ctx, block_size >= 16 && (block_size >= 128 ? block_size % 128 == 0 : 128 % block_size == 0), NVC_ESHAPE,
              "paged prefill needs block_size in {16, 32, 64} or a multiple of 128");
  NVC_REQUIRE(ctx, ldq % 8 == 0 && ldq >= nh * hd, NVC_ESHAPE, "q row stride must cover the heads and be a multiple of 8");
  NVC_REQUIRE(ctx, ((uintptr_t)q % 16) == 0 && ((uintptr_t)k_cache % 16) == 0 && ((uintptr_t)v_cache % 16) == 0,
              NVC_EINVAL, "pointers must be 16-byte aligned");
  NVC_REQUIRE(ctx, ctx->encode_tiled != nullptr, NVC_EDEVICE, "cuTensorMapEncodeTiled unavailable");
  if (B == 0 || total_q_tokens == 0 || max_seqlen_q == 0) return NVC_OK;
  const bool v1 = use_v1();
  const int tile = v1 ? 128 : v2::kKT;
  const int box_rows = block_size < tile ? block_size : tile;
  CUtensorMap tq, tk, tv;
  int rc = make_tma_heads(ctx, &tq, q, nh, total_q_tokens, ldq);
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tk, k_cache, kv, (int64_t)num_blocks * block_size, kv * hd, box_rows);   // cache = [slots, kv, hd]
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tv, v_cache, kv, (int64_t)num_blocks * block_size, kv * hd, box_rows);
  if (rc) return rc;
  PagedKV pk{block_tables, context_lens, bt_stride, block_size, box_rows};
  dim3 grid((max_seqlen_q + 127) / 128, nh, B);
  if (!v1) {
    const int QT = (max_seqlen_q + 127) / 128;
    NVC_CUDA_CHECK(ctx, launch_v2<true>(ctx, QT, nh, kv, B, nvc_stream(stream), tq, tk, tv, cu_seqlens_q,
                                        scale * 1.4426950408889634f, out, pk));
    NVC_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->launches++;
    return NVC_OK;
  }
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, prefill_attn_kernel<true>, grid, dim3(kThreads), kSmemBytes, nvc_stream(stream), tq, tk, tv,
                                 cu_seqlens_q, nh, nh / kv, scale * 1.4426950408889634f, out, 16384, 1024, pk));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}
