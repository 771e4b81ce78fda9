// Library-internal entry points (not part of the public C ABI).
#pragma once
#include <stddef.h>
#include <stdint.h>
#include "../../include/nvc_b200.h"

extern "C" {
// paged decode with separate block-table row stride and a max-blocks hint for the split choice
int nvc_paged_decode_attn_ex(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                             const nvc_bf16* v_cache, const int32_t* block_tables, int bt_stride, int max_blocks_hint,
                             const int32_t* context_lens, int B, int nh, int kv, int hd, int block_size, float scale,
                             nvc_bf16* out, void* workspace, size_t workspace_bytes, void* stream);
// pairs[b] = {val[b], idx[b]} as 8-byte records
int nvc_pack_argmax_pairs(nvc_ctx* ctx, const float* val, const int32_t* idx, float* pairs, int B, void* stream);
int nvc_gemm_bf16_ex(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_dtype,
                     int M, int Nout, int K, int force_bn, int force_splits, void* stream);
int nvc_gemm_bf16_silu_mul_ex(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W_gu16, nvc_bf16* act, int ldc,
                              int M, int I, int K, int force_bn, int force_splits, void* stream);
int nvc_gemm_bf16_trace(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_dtype,
                        int M, int Nout, int K, int force_bn, int force_splits, void* stream, long long* trace_d);
// set max-dynamic-smem attributes of every kernel instantiation (called once per context)
int nvc_gemm_init(nvc_ctx* ctx);
int nvc_gemm_pair_init(nvc_ctx* ctx);
// CTA-pair (cta_group::2) GEMM for M > 128; force_splits 0 = choose the in-cluster split-K factor
int nvc_gemm_pair(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_f32, int M,
                  int N, int K, int force_splits, int epi, void* stream, long long* trace);
// fp32 C [M <= 128 tokens, N] = X . W^T with the operands swapped in the pair kernel (lm_head at decode batch sizes)
int nvc_gemm_pair_swapped(nvc_ctx* ctx, const nvc_bf16* X, int ldx, const nvc_bf16* W, float* C, int ldc, int M, int N, int K,
                          void* stream, long long* trace);
// K3 + K2 through the pair kernel's persistent epilogue; NVC_ENOTIMPL (nothing launched) when the shape takes another schedule
int nvc_gemm_pair_qk(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, nvc_bf16* C, int ldc, int M, int N, int K,
                     const nvc_bf16* qw, const nvc_bf16* kw, float eps, const float* cos_sin, const int32_t* positions,
                     const int32_t* slot_mapping, nvc_bf16* k_cache, nvc_bf16* v_cache, int nh, int kv, void* stream);
int nvc_prefill_attn_init(nvc_ctx* ctx);
int nvc_paged_decode_init(nvc_ctx* ctx);
}
