// K4: varlen causal flash attention for prefill on tcgen05 / TMEM.
// Replaces the dense attention of Qwen3Attention::forward (src/models/qwen3.rs:236-277):
// GQA expand (:240-256), QK^T*scale (:258-259), host-built -1e9 mask + H2D (:260-270),
// f32 softmax (:271-273), P*V and head merge (:275-277).  No mask tensor, no materialised
// K/V copies, no [B,nh,T,T] scores.
//
// The kernel is prefill_attn_v2.cuh (persistent grid, 128-row q tiles x 64-token kv tiles, TMA over 3-D tensor maps
// {head_dim, head, token} so q / k / v are read in place from the fused qkv rows, S / P / O in tensor memory); this file
// holds the tensor maps, the launchers and the C ABI.  The first-generation kernel (one CTA per SM, 128-token kv tiles,
// output accumulator in registers: 215-396 TFLOP/s, profiles/r01_s14_prefill_attn_v2.md) was deleted in round 2.
// FLOPs per (sequence, head): 4 * 128 * T^2 / 2 (causal).
//
// kPaged = true (prefix caching): the q rows of sequence b are its LAST q_len
// positions of a context of context_lens[b] tokens; K / V tiles are read from the paged cache
// through block_tables (one TMA box per block, or per 64 tokens when block_size >= 64), after
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
  int box_rows;                               // token rows per K / V TMA box = min(64, block_size)
};

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

}  // namespace

extern "C" int nvc_prefill_attn_init(nvc_ctx* ctx) {
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(v2::prefill_attn_v2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, v2::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(v2::prefill_attn_v2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, v2::kSmemBytes));
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
  int rc = make_tma_heads(ctx, &tq, q, nh, total_tokens, ldq);
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tk, k, kv, total_tokens, ldk, v2::kKT);
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tv, v, kv, total_tokens, ldv, v2::kKT);
  if (rc) return rc;
  const int QT = (max_seqlen + 127) / 128;
  NVC_CUDA_CHECK(ctx, launch_v2<false>(ctx, QT, nh, kv, B, nvc_stream(stream), tq, tk, tv, cu_seqlens,
                                       scale * 1.4426950408889634f, out, PagedKV{}));
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
  const int box_rows = block_size < v2::kKT ? block_size : v2::kKT;
  CUtensorMap tq, tk, tv;
  int rc = make_tma_heads(ctx, &tq, q, nh, total_q_tokens, ldq);
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tk, k_cache, kv, (int64_t)num_blocks * block_size, kv * hd, box_rows);   // cache = [slots, kv, hd]
  if (rc) return rc;
  rc = make_tma_heads(ctx, &tv, v_cache, kv, (int64_t)num_blocks * block_size, kv * hd, box_rows);
  if (rc) return rc;
  PagedKV pk{block_tables, context_lens, bt_stride, block_size, box_rows};
  const int QT = (max_seqlen_q + 127) / 128;
  NVC_CUDA_CHECK(ctx, launch_v2<true>(ctx, QT, nh, kv, B, nvc_stream(stream), tq, tk, tv, cu_seqlens_q,
                                      scale * 1.4426950408889634f, out, pk));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}
