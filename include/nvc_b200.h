/*
 * nvc_b200.h -- C ABI of the B200-native (sm_100a) Qwen3 forward hot path that drops in
 * behind nano-vllm-candle's layer / runner API.
 *
 * The reference crate has no FFI of its own: its boundary is a set of Rust traits and
 * inherent methods that call candle tensor ops.  Each entry point
 * below names the reference interface (file:line, relative to the reference crate root)
 * whose body it replaces; INTEGRATION.md shows the Rust `extern "C"` block and the
 * candle CustomOp wrapper a maintainer would add.
 *
 * Conventions
 *   - plain C: pointers + sizes, no torch / candle types.  Every `*_d` / unqualified
 *     tensor pointer is a DEVICE pointer (CUdeviceptr cast to void*), caller-owned,
 *     row-major, contiguous unless a row stride (in ELEMENTS) is given.
 *   - `stream` is a cudaStream_t passed as void*.  Calls only ENQUEUE work on it
 *     (asynchronous w.r.t. the host) unless the name ends in `_sync` or the doc says so.
 *   - return value: NVC_OK (0) on success; negative NVC_E* for argument / shape / state
 *     errors; positive = 1000 + cudaError_t, 2000 + ncclResult_t.  Nothing aborts and
 *     nothing is swallowed (the reference's runner turns errors into EOS tokens,
 *     src/engine/llm_engine.rs:153-175; that policy stays on the Rust side).
 *   - one nvc_ctx per (process, device); a context is used from one host thread at a
 *     time (the reference is single-threaded: Rc<RefCell<..>>, src/engine/llm_engine.rs:17).
 *   - bf16 = IEEE bfloat16 bit pattern in uint16_t storage.
 *   - launch order: every kernel is launched with programmatic stream serialization (PDL) and signals its
 *     dependents at its first instruction, so a kernel of this library may START while the previous kernels of
 *     this library on the stream are still running.  It waits (griddepcontrol.wait) before reading anything
 *     those kernels produce, but reads its STATIC operands earlier: weights, norm vectors, the rope table,
 *     positions / slot_mapping / block_tables / context_lens / cu_seqlens, and - paged decode only - cache rows
 *     other than the slot of each sequence's newest token.  Static operands must therefore be complete when the
 *     call is enqueued in stream order by anything that is not a kernel of this library (memcpy, other
 *     libraries' kernels: those never signal early, so plain stream order is enough) and must never be the
 *     output of a kernel of this library in the same stream.
 *   - no CPU fallback: on a machine without a usable sm_100 device nvc_ctx_create fails.
 */
#ifndef NVC_B200_H
#define NVC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NVC_ABI_VERSION 1

enum {
  NVC_OK = 0,
  NVC_EINVAL = -1,     /* bad argument (null pointer, non-positive size, misalignment) */
  NVC_ESHAPE = -2,     /* unsupported shape (e.g. head_dim != 128, K % 64 != 0)         */
  NVC_ENOTIMPL = -3,
  NVC_ESTATE = -4,     /* wrong call order / missing weights / not initialised          */
  NVC_ENOMEM = -5,     /* workspace / KV pool too small                                 */
  NVC_EDEVICE = -6,    /* no sm_100 device / driver entry point missing                 */
  NVC_ENCCL = -7,      /* NCCL library not loadable                                     */
  NVC_ENVLS = -8,      /* NVSwitch multicast not available (nvc_ctx_nvls_status says why) */
  NVC_CUDA_ERROR_BASE = 1000,
  NVC_NCCL_ERROR_BASE = 2000
};

typedef uint16_t nvc_bf16;
typedef struct nvc_ctx nvc_ctx;
typedef struct nvc_runner nvc_runner;
typedef struct nvc_engine nvc_engine;

/* ------------------------------------------------------------------ context */

int nvc_abi_version(void);
const char* nvc_status_string(int status);
/* Last error message recorded on this context (empty string if none). */
const char* nvc_last_error(const nvc_ctx* ctx);

/* Binds to CUDA device `device`; tp_rank/tp_size mirror TPConfig (src/tp.rs:3-70).
 * Fails with NVC_EDEVICE unless the device is compute capability 10.x. */
int nvc_ctx_create(int device, int tp_rank, int tp_size, nvc_ctx** out);
int nvc_ctx_destroy(nvc_ctx* ctx);
int nvc_ctx_device(const nvc_ctx* ctx);
int nvc_ctx_tp_rank(const nvc_ctx* ctx);
int nvc_ctx_tp_size(const nvc_ctx* ctx);
int nvc_ctx_num_sms(const nvc_ctx* ctx);
/* Kernels launched through this context so far (bench.py reports the per-step difference). */
uint64_t nvc_ctx_launch_count(const nvc_ctx* ctx);

/* Tensor-parallel communicator (absent in the reference: src/tp.rs has no collectives,
 * src/layers/linear.rs:184-198 does no all-reduce).  `unique_id` is the 128-byte
 * ncclUniqueId produced by nvc_nccl_unique_id on rank 0 and shipped to the other ranks
 * by the caller.  Blocking (rendezvous). */
int nvc_nccl_unique_id(void* unique_id_128B);
int nvc_ctx_init_nccl(nvc_ctx* ctx, const void* unique_id_128B);
/* In-place sum over the tp ranks of `count` bf16 / f32 elements (K9). */
int nvc_allreduce_bf16(nvc_ctx* ctx, nvc_bf16* buf, size_t count, void* stream);
int nvc_allreduce_f32(nvc_ctx* ctx, float* buf, size_t count, void* stream);
/* In-place all-gather: every rank contributes `count_per_rank` bytes at offset rank*count. */
int nvc_allgather_bytes(nvc_ctx* ctx, void* buf, size_t bytes_per_rank, void* stream);

/* Peer-memory all-reduce over NVLink / NVSwitch (one-shot, for decode-sized messages), fused with
 * the residual add + RMSNorm that follow every row-parallel linear.  Setup: every rank calls
 * nvc_ctx_p2p_export (allocates [flags | slot 0 | slot 1] and returns a 64-byte CUDA IPC handle),
 * the caller all-gathers the handles, every rank calls nvc_ctx_p2p_import with the tp_size x 64
 * bytes.  The row-parallel GEMM writes its partial [N, H] bf16 result into nvc_ctx_p2p_slot(s)
 * (alternate s between consecutive all-reduces), then
 *   nvc_p2p_allreduce_rmsnorm: sum = sum over ranks (fp32, rank order) of slot s;
 *     w != NULL: s = sum + residual_in; residual_out = s; out = bf16(rmsnorm(s) * w)  (K9 + K1)
 *     w == NULL: out = bf16(sum).
 * tp_size in {2, 4, 8}; N * H * 2 <= slot_bytes. */
int nvc_ctx_p2p_export(nvc_ctx* ctx, size_t slot_bytes, void* handle_64B);
int nvc_ctx_p2p_import(nvc_ctx* ctx, const void* handles_tp_x_64B);
int nvc_ctx_p2p_ready(const nvc_ctx* ctx);
size_t nvc_ctx_p2p_slot_bytes(const nvc_ctx* ctx);
void* nvc_ctx_p2p_slot(nvc_ctx* ctx, int slot);
int nvc_p2p_allreduce_rmsnorm(nvc_ctx* ctx, int slot, const float* residual_in, float* residual_out,
                              const nvc_bf16* w, float eps, nvc_bf16* out, int N, int H, void* stream);

/* In-switch (NVLS) all-reduce over the NVSwitch multicast object of the TP group, fused with the residual add +
 * RMSNorm that follow every row-parallel linear (same contract as nvc_p2p_allreduce_rmsnorm).
 *   nvc_ctx_nvls_init: COLLECTIVE - every rank of the group calls it with the same sizes, after
 *     nvc_ctx_init_nccl (whose unique id names the group; `group_name` may be NULL to use that name).  Allocates
 *     one buffer per rank [flags | slot | 3 LL buffers], binds it into one multicast object (cuMulticastCreate on
 *     rank 0; the POSIX descriptor travels to the peers over an abstract unix socket of the library's own) and maps
 *     it at a unicast and a multicast address.  Returns NVC_ENVLS on every rank alike when any rank cannot do its
 *     part (no multicast support, creation refused); nvc_ctx_nvls_status then says why and the caller stays on the
 *     peer-memory / NCCL path.
 *   nvc_ctx_nvls_slot: this rank's slot (device pointer, slot_bytes): where the row-parallel GEMM writes its
 *     partial [N, H] bf16 result when the two-shot kernel is to reduce it.
 *   nvc_nvls_allreduce_rmsnorm: x_partial = this rank's partial sums.  mode 0 chooses by size; 1 = LL one-shot
 *     (rank r multimem.st's its rows into every rank's LL buffer, readers spin on the data; x_partial anywhere;
 *     tp * N * H * 2 <= ll_buf_bytes); 2 = two-shot in the switch (multimem.ld_reduce of 1/tp of the slot,
 *     multimem.st of the sums in place, two flag barriers; x_partial must be nvc_ctx_nvls_slot); 3 = two-shot with
 *     the sums broadcast into an LL buffer (one barrier; N * H * 2 <= ll_buf_bytes).  Results are bit-identical on
 *     all ranks in every mode.  tp_size in {2, 4, 8}; H % 64 == 0, H <= 8192. */
int nvc_ctx_nvls_init(nvc_ctx* ctx, size_t slot_bytes, size_t ll_buf_bytes, const char* group_name);
int nvc_ctx_nvls_ready(const nvc_ctx* ctx);
const char* nvc_ctx_nvls_status(const nvc_ctx* ctx);
size_t nvc_ctx_nvls_slot_bytes(const nvc_ctx* ctx);
size_t nvc_ctx_nvls_ll_bytes(const nvc_ctx* ctx);
void* nvc_ctx_nvls_slot(nvc_ctx* ctx);
int nvc_nvls_allreduce_rmsnorm(nvc_ctx* ctx, const nvc_bf16* x_partial, const float* residual_in, float* residual_out,
                               const nvc_bf16* w, float eps, nvc_bf16* out, int N, int H, int mode, void* stream);
/* End of a TP step (vocab-parallel lm_head + per-shard argmax / sample): all-gather of the ranks' (best value,
 * token) candidates through the multicast LL buffers and merge - largest value, ties -> the larger token id
 * (llm_engine.rs:135-142 "last maximum") - in one kernel; out_tok [B] is identical on every rank.  out_val may be
 * NULL.  Replaces nvc_pack_argmax_pairs + nvc_allgather_bytes + the host-side merge. */
int nvc_nvls_argmax_merge(nvc_ctx* ctx, const float* best_val, const int32_t* best_tok, int32_t* out_tok, float* out_val,
                          int B, void* stream);

/* ------------------------------------------------------------------ kernels */

/* K0. Replaces Tensor::embedding in Qwen3Model::forward (src/models/qwen3.rs:465-468).
 * out[n, :] = table[ids[n], :].  ids must lie in [0, V) (validated by the caller on the host). */
int nvc_embed_gather(nvc_ctx* ctx, const int32_t* ids, const nvc_bf16* table, nvc_bf16* out,
                     int N, int H, int V, void* stream);

/* K1. Replaces RMSNorm::forward (src/layers/layernorm.rs:44-60), fused:
 *   s = f32(x) + residual_in (if residual_in != NULL, else s = f32(x));
 *   residual_out = s (if residual_out != NULL; may alias residual_in);
 *   out = bf16( s * rsqrt(mean(s^2) + eps) * w ).
 * x, out: bf16 [N, H]; residuals: f32 [N, H] (the reference keeps the residual stream
 * in f32, layernorm.rs:49-50); w: bf16 [H].  H % 8 == 0, H <= 16384. */
int nvc_fused_add_rmsnorm(nvc_ctx* ctx, const nvc_bf16* x, const float* residual_in,
                          float* residual_out, const nvc_bf16* w, float eps, nvc_bf16* out,
                          int N, int H, void* stream);

/* K3. Replaces candle_nn::Linear::forward inside every *ParallelLinear::forward
 * (src/layers/linear.rs:33-38,72-77,99-104,171-175,184-198) and compute_logits
 * (src/models/qwen3.rs:542-550):  C[M, Nout] = A[M, K] * W[Nout, K]^T  (W is the
 * reference's [out, in] layout), bf16 inputs, fp32 accumulation on tcgen05 / TMEM.
 * lda / ldc: row strides in elements.  K % 64 == 0, Nout % 16 == 0, lda % 8 == 0.
 * out_dtype: NVC_DT_BF16 or NVC_DT_F32. */
enum { NVC_DT_BF16 = 0, NVC_DT_F32 = 1 };
int nvc_gemm_bf16(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                  int out_dtype, int M, int Nout, int K, void* stream);

/* K3 + K6 fused.  Replaces Qwen3MLP::forward's gate_up projection and SiluAndMul
 * (src/models/qwen3.rs:323-327, src/layers/activation.rs:13-18) in one kernel:
 *   act[M, I] = silu(A Wgate^T) * (A Wup^T),  silu(x) = x / (1 + exp(-x)) on the fp32 accumulators.
 * W_gu16: bf16 [2I, K], the reference's fused [gate; up] rows (qwen3.rs:310) re-ordered in groups of
 * 16 so that both halves of a feature meet in one epilogue thread: rows [32g, 32g+16) = gate rows
 * [16g, 16g+16), rows [32g+16, 32g+32) = up rows [16g, 16g+16).  I % 16 == 0, K % 64 == 0. */
int nvc_gemm_bf16_silu_mul(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W_gu16, nvc_bf16* act,
                           int ldc, int M, int I, int K, void* stream);

/* Host-side rope table: replaces RotaryEmbedding::build_cos_sin
 * (src/layers/rotary_embedding.rs:56-80), which the reference recomputes per call.
 * out: f32 [max_pos, head_dim]; row p = [cos(p*f_0..f_{hd/2-1}), sin(p*f_0..)], with
 * f_j = 1/theta^(2j/hd), all in f32 exactly as the reference (glibc powf/cosf/sinf on the
 * host, then one H2D copy).  Synchronous. */
int nvc_rope_table_init_sync(nvc_ctx* ctx, float theta, int head_dim, int max_pos, float* out);

/* K2. Replaces the split / q_norm / k_norm / RotaryEmbedding::apply block of
 * Qwen3Attention::forward (src/models/qwen3.rs:206-234, src/layers/rotary_embedding.rs:82-107)
 * and adds the KV-cache scatter (N3).  In place on qkv bf16 [N, (nh+2kv)*hd]:
 *   q_h <- rope(rmsnorm(q_h) * qw, positions[n]);  k_h <- rope(rmsnorm(k_h) * kw, positions[n]);
 *   if slot_mapping != NULL and slot_mapping[n] >= 0:
 *       k_cache[slot, h, :] = k_h;  v_cache[slot, h, :] = v_h      (bit-exact copies)
 * k_cache / v_cache: bf16 [num_slots, kv, hd] = [num_blocks, block_size, kv, hd].
 * qw == kw == NULL skips the norm (rope only = RotaryEmbedding::apply by itself).
 * hd must be 128. */
int nvc_qk_norm_rope_store(nvc_ctx* ctx, nvc_bf16* qkv, const nvc_bf16* qw, const nvc_bf16* kw,
                           float eps, const float* cos_sin, const int32_t* positions,
                           const int32_t* slot_mapping, nvc_bf16* k_cache, nvc_bf16* v_cache,
                           int N, int nh, int kv, int hd, void* stream);

/* K3 + K2: the fused QKV projection with K2 as its epilogue (Qwen3Attention::forward up to the attention call,
 * src/models/qwen3.rs:205-234).  qkv [M, (nh + 2 kv) * hd] contiguous = A W_qkv^T with q / k heads normalised and rotated
 * and the k / v rows scattered to the caches at slot_mapping[row] (may be NULL: no scatter) - exactly what nvc_gemm_bf16
 * followed by nvc_qk_norm_rope_store produce, bit for bit.  Prefill-sized M runs as ONE kernel (the pair GEMM's persistent
 * epilogue parks each head in shared memory and applies K2 there: the separate pass over the QKV buffer disappears);
 * other shapes run the two kernels. */
int nvc_qkv_proj_norm_rope(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W_qkv, nvc_bf16* qkv, int M, int K,
                           const nvc_bf16* q_norm_w, const nvc_bf16* k_norm_w, float eps, const float* cos_sin,
                           const int32_t* positions, const int32_t* slot_mapping, nvc_bf16* k_cache, nvc_bf16* v_cache,
                           int nh, int kv, int hd, void* stream);

/* K2'. Stand-alone slot-mapping scatter (N3): cache[slot_mapping[n]] = {k,v}[n]; slot < 0 skips.
 * k, v: bf16 rows of kv*hd elements with row stride ld (elements).  Bit-exact. */
int nvc_store_kvcache(nvc_ctx* ctx, const nvc_bf16* k, const nvc_bf16* v, int ld,
                      nvc_bf16* k_cache, nvc_bf16* v_cache, const int32_t* slot_mapping,
                      int N, int kv, int hd, void* stream);

/* K4. Prefill: replaces the dense causal attention of Qwen3Attention::forward
 * (src/models/qwen3.rs:236-277) with varlen causal flash attention.  q: [N, nh, hd],
 * k, v: [N, kv, hd] (row strides ldq / ldk / ldv in elements, so they may point into the
 * fused qkv buffer); cu_seqlens: i32 [B+1] (device); sequence b owns rows
 * cu_seqlens[b]..cu_seqlens[b+1], positions 0..len (no cached prefix).
 * out: bf16 [N, nh*hd], head-major like the reference (:276-277).  GQA: q head h reads
 * kv head h / (nh/kv) (:236-256).  hd must be 128. */
int nvc_prefill_attn_varlen(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k, int ldk,
                            const nvc_bf16* v, int ldv, const int32_t* cu_seqlens, int B,
                            int total_tokens, int max_seqlen, int nh, int kv, int hd, float scale,
                            nvc_bf16* out, void* stream);

/* K4 (paged). The same kernel for sequences with a cached prefix (prefix caching, SURVEY.md 8f.1;
 * upstream nano-vllm passes block_table to flash_attn_varlen_func for this): q rows of sequence b
 * are the LAST q_len_b = cu_seqlens_q[b+1] - cu_seqlens_q[b] positions of a context of
 * context_lens[b] tokens whose K / V are read from the paged cache through block_tables
 * [B, bt_stride] (this step's k / v must already be stored there).  Causal: q row i sees keys
 * 0 .. context_lens[b] - q_len_b + i.  Cache rows past the context may hold anything (NaN included).
 * block_size in {16, 32, 64} or a multiple of 128. */
int nvc_prefill_attn_paged(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                           const nvc_bf16* v_cache, const int32_t* block_tables, int bt_stride,
                           const int32_t* context_lens, const int32_t* cu_seqlens_q, int B,
                           int total_q_tokens, int max_seqlen_q, int nh, int kv, int hd, int block_size,
                           int num_blocks, float scale, nvc_bf16* out, void* stream);

/* K5. Decode: paged attention over block_tables (N2) with GQA head grouping.
 * q: [B, nh, hd] row stride ldq; caches [num_blocks, block_size, kv, hd];
 * block_tables: i32 [B, max_blocks] (-1 padded); context_lens: i32 [B] (tokens in the cache
 * INCLUDING the current one).  out: bf16 [B, nh*hd].
 * workspace: device scratch of at least nvc_paged_decode_workspace_bytes(...) bytes. */
size_t nvc_paged_decode_workspace_bytes(int B, int nh, int hd, int num_sms);
int nvc_paged_decode_attn(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                          const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                          const int32_t* context_lens, int B, int nh, int kv, int hd, int block_size,
                          float scale, nvc_bf16* out, void* workspace, size_t workspace_bytes,
                          void* stream);

/* K5, allocation-free variant.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * kernels and results as nvc_paged_decode_attn; the KV-split partials go to a static device array inside the
 * library (nvc_paged_decode_static_scratch_bytes() = 4 MiB per GPU, held until exit) instead of a caller workspace.
 * Returns NVC_ENOMEM (nothing launched, message in nvc_last_error) when the shape's partials exceed that array -
 * the library's own split choice needs at most ~2.5 MB.  Calls no cudaMalloc / cudaFree.  All calls on a device
 * share the array: calls that overlap on the GPU (different streams) race; nothing in it has to be cleared between
 * calls (every partial read is written by the same call). */
size_t nvc_paged_decode_static_scratch_bytes(void);
int nvc_paged_decode_attn_static(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                 const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                                 const int32_t* context_lens, int B, int nh, int kv, int hd, int block_size,
                                 float scale, nvc_bf16* out, void* stream);

/* K6. Replaces SiluAndMul::forward (src/layers/activation.rs:13-18):
 * out[n, i] = silu(x[n, i]) * x[n, I + i]. */
int nvc_silu_and_mul(nvc_ctx* ctx, const nvc_bf16* x, nvc_bf16* out, int N, int I, void* stream);

/* K7. Row gather: out[b, :] = x[rows[b], :] (last token of each sequence before the final
 * norm / lm_head; the reference computes logits for all positions, src/models/qwen3.rs:548). */
int nvc_gather_rows(nvc_ctx* ctx, const nvc_bf16* x, const float* residual, const int32_t* rows,
                    nvc_bf16* x_out, float* residual_out, int B, int H, void* stream);

/* K8. Replaces Qwen3ModelRunner::argmax (src/engine/llm_engine.rs:135-142) on device:
 * idx[b] = LAST index of the maximum of logits[b, :] (Iterator::max_by tie rule);
 * val[b] = that maximum.  idx is offset by `index_offset` (TP vocab shard). */
int nvc_argmax_last(nvc_ctx* ctx, const float* logits, int ld, int32_t* idx, float* val,
                    int B, int V, int index_offset, void* stream);

/* K8s. Replaces Qwen3ModelRunner::sample_token (src/engine/llm_engine.rs:97-133) on device.  The
 * reference draws from softmax((l - max) / max(T, 1e-6)) with an unseeded rand::rng() (:128); this is
 * the Gumbel-max form of the same categorical distribution with counter-based, reproducible noise:
 *   idx[b] = argmax_i( logits[b, i] / max(T_b, 1e-6) + g ),  g = -log(-log(u)),
 *   u = (2 * (x >> 9) + 1) * 2^-24 (exact in fp32, never 0 or 1),  x = word (gi % 4) of Philox4x32-10(key = seed,
 *   counter = (gi / 4, streams[b] lo, streams[b] hi, 0)),  gi = index_offset + i  (global vocab index),
 * ties -> last index.  val[b] = the winning perturbed score, so tensor-parallel vocab shards merge
 * like the greedy (max, idx) pairs.  temperatures: f32 [B]; streams: u64 [B] (one value per
 * (sequence, step)); V, ld, index_offset multiples of 4. */
int nvc_sample_gumbel(nvc_ctx* ctx, const float* logits, int ld, const float* temperatures,
                      const uint64_t* streams, uint64_t seed, int32_t* idx, float* val, int B, int V,
                      int index_offset, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NVC_B200_H */
