/*
 * nvc_b200_engine.h -- C ABI of the host side above the kernels: the Qwen3 step executor
 * ("runner") and the engine book-keeping (Sequence / BlockManager / Scheduler / LLMEngine).
 *
 * The reference's host side is Rust (no toolchain on the build box), so it is written in
 * C++ behind this C ABI, mirroring the reference's types one to one:
 *   nvc_runner  <-> Qwen3ModelRunner + Qwen3ForCausalLM   src/engine/llm_engine.rs:35-189,
 *                                                         src/models/qwen3.rs:402-551
 *   nvc_engine  <-> LLMEngine<R> + Scheduler + BlockManager + Sequence
 *                                                         src/engine/llm_engine.rs:209-326,
 *                                                         src/engine/scheduler.rs:58-250,
 *                                                         src/engine/block_manager.rs:31-99,
 *                                                         src/engine/sequence.rs:16-97
 * All pointers in THIS header are HOST pointers unless named *_d.
 */
#ifndef NVC_B200_ENGINE_H
#define NVC_B200_ENGINE_H

#include "nvc_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Qwen3Config (src/models/qwen3.rs:19-34). */
typedef struct nvc_qwen3_config {
  int32_t vocab_size;
  int32_t hidden_size;
  int32_t head_dim;
  int32_t num_hidden_layers;
  int32_t num_attention_heads;
  int32_t num_key_value_heads;
  int32_t intermediate_size;
  int32_t max_position_embeddings;
  float rms_norm_eps;
  float rope_theta;
  int32_t bos_token_id;
  int32_t eos_token_id;
} nvc_qwen3_config;

typedef struct nvc_runner_options {
  int32_t num_kvcache_blocks;      /* per-rank KV pool size in blocks (SchedulerConfig.num_kvcache_blocks) */
  int32_t kvcache_block_size;      /* tokens per block (256 in the reference, sequence.rs:35)               */
  int32_t max_num_seqs;            /* largest decode batch                                                   */
  int32_t max_num_batched_tokens;  /* largest prefill step in tokens                                         */
  int32_t max_model_len;           /* longest sequence (rope table / block table width)                      */
  int32_t use_cuda_graph;          /* capture decode steps per batch size                                    */
  int32_t allreduce_f32;           /* TP: reduce fp32 partials instead of bf16 (default 0)                   */
  int32_t reserved[4];
} nvc_runner_options;

/* ------------------------------------------------------------------ runner */

/* Allocates weights (this rank's shard), KV pool, activations.  Replaces
 * Qwen3ModelRunner::from_hf_dir's model construction (src/engine/llm_engine.rs:43-58); file
 * parsing stays with the caller, tensors are handed over with nvc_runner_load_weight. */
int nvc_runner_create(nvc_ctx* ctx, const nvc_qwen3_config* cfg, const nvc_runner_options* opt, nvc_runner** out);
int nvc_runner_destroy(nvc_runner* r);

/* One HF tensor (names / shapes of src/models/qwen3.rs:146-187,301-314,351-363,431-447,525-528),
 * FULL (unsharded) bf16 data in host (or device) memory, row-major [rows, cols] (cols = 1 for vectors).
 * The runner fuses q/k/v (:171) and gate/up (:310) and slices its tensor-parallel shard:
 * q/k/v/gate/up rows and lm_head rows by head / column group, o_proj / down_proj input
 * columns (the reference's tp_dim 0 for RowParallelLinear, linear.rs:217, is wrong for this). */
int nvc_runner_load_weight(nvc_runner* r, const char* hf_name, const nvc_bf16* data, int64_t rows, int64_t cols);
/* Verifies every tensor arrived, builds the rope table.  Must be called before any step. */
int nvc_runner_finalize(nvc_runner* r);

/* ---- native HF checkpoint loading (host side; error text in nvc_hf_last_error, thread-local) ---- */
const char* nvc_hf_last_error(void);
/* Non-fatal notes of the last nvc_runner_load_hf_dir on this thread (tensors the model does not use were skipped). */
const char* nvc_hf_last_warning(void);
/* Qwen3Config::from_hf_dir (src/models/qwen3.rs:77-101): <model_dir>/config.json, same required
 * fields, head_dim defaults to hidden_size / num_attention_heads, rope_theta to 1e6.
 * *tie_word_embeddings (optional) = that flag of config.json (absent = 0). */
int nvc_qwen3_config_from_hf_dir(const char* model_dir, nvc_qwen3_config* out, int* tie_word_embeddings);
/* Qwen3ForCausalLM::from_hf_dir (src/models/qwen3.rs:515-536) for an existing runner: reads
 * model.safetensors, or every shard named by model.safetensors.index.json, through mmap; BF16 is
 * handed over as is, F16 / F32 are rounded to bf16 (nearest even) on the host; every tensor goes
 * through nvc_runner_load_weight; a checkpoint without lm_head.weight uses embed_tokens (tied
 * embeddings; the reference fails there, :525-528); ends with nvc_runner_finalize.  Like the reference's
 * VarBuilder it takes only the tensors the model names: others are skipped (nvc_hf_last_warning); shard names of
 * the index must be plain file names inside model_dir; overlapping data_offsets are refused. */
int nvc_runner_load_hf_dir(nvc_runner* r, const char* model_dir);
/* safetensors inspection (what the loader sees), also used by the CPU tests. */
typedef struct nvc_safetensors nvc_safetensors;
enum { NVC_ST_BF16 = 0, NVC_ST_F16 = 1, NVC_ST_F32 = 2, NVC_ST_OTHER = 3 };
int nvc_safetensors_open(const char* path, nvc_safetensors** out);
int nvc_safetensors_close(nvc_safetensors* st);
int nvc_safetensors_num_tensors(const nvc_safetensors* st);
/* shape4: up to 4 dims, padded with 1. */
int nvc_safetensors_tensor_info(const nvc_safetensors* st, int i, const char** name, int* dtype, int* ndim,
                                int64_t* shape4);
/* Converts tensor `name` to bf16 into the HOST buffer `out` (capacity in elements). */
int nvc_safetensors_read_bf16(const nvc_safetensors* st, const char* name, nvc_bf16* out, int64_t capacity_elems);

/* One prefill step over B sequences / N new tokens (flat varlen layout replacing build_batch's
 * right-padded [B, max_len], src/engine/llm_engine.rs:60-95).  Synchronous: returns the greedy
 * next token of every sequence in out_tokens[B] (argmax, ties -> last index,
 * src/engine/llm_engine.rs:135-142). */
int nvc_runner_prefill(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                       const int32_t* slot_mapping, int B, int N, int max_seqlen, int32_t* out_tokens);
/* Prefill step whose sequences may have a cached prefix (prefix caching): the N new
 * tokens of sequence b are positions context_lens[b] - q_len_b .. context_lens[b] - 1 and attend
 * to ALL context_lens[b] tokens, read from the paged cache through block_tables [B, max_blocks]
 * (the new tokens' K/V are scattered to slot_mapping first, as in every step). */
int nvc_runner_prefill_paged(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                             const int32_t* slot_mapping, const int32_t* block_tables, int max_blocks,
                             const int32_t* context_lens, int B, int N, int max_seqlen, int32_t* out_tokens);
/* Mixed step (chunked prefill): the first num_decode sequences have exactly one query
 * token (cu_seqlens[b] == b) and attend through the paged DECODE kernel (K5); the other B -
 * num_decode sequences are prompt chunks of any length through the paged PREFILL kernel (K4
 * paged); everything else (embedding, norms, GEMMs, KV store, head) runs once over all N tokens.
 * out_tokens[B]: next token of every sequence from its last query row (the caller ignores it for
 * a chunk that does not end its prompt).  num_decode == B is a decode step without a CUDA graph;
 * num_decode == 0 is nvc_runner_prefill_paged. */
int nvc_runner_step_mixed(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                          const int32_t* slot_mapping, const int32_t* block_tables, int max_blocks,
                          const int32_t* context_lens, int num_decode, int B, int N, int max_seqlen,
                          int32_t* out_tokens);
/* One decode step: one new token per sequence, attention over the paged cache. */
int nvc_runner_decode(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* slot_mapping,
                      const int32_t* block_tables, int max_blocks, const int32_t* context_lens, int B,
                      int32_t* out_tokens);
/* Arms on-device temperature sampling (K8s, nvc_sample_gumbel) for the NEXT prefill / decode step
 * only: temperatures f32 [B] (> 0), streams u64 [B] (one noise stream per (sequence, step)), seed =
 * Philox key.  temperatures == NULL disarms.  Without this call a step is greedy (argmax, ties ->
 * last index).  Replaces the host loop of Qwen3ModelRunner::sample_token (llm_engine.rs:97-133). */
int nvc_runner_set_sampling(nvc_runner* r, const float* temperatures, const uint64_t* streams, int B, uint64_t seed);
/* fp32 logits [B, vocab] of the sequences of the last step (parity checks; gathers TP shards). */
int nvc_runner_last_logits(nvc_runner* r, float* out, int B);
/* Device-resident timing hook for bench.py: replays the last decode step `iters` times on the
 * runner's stream with inputs already in HBM and returns the average step time in ms measured
 * with CUDA events; *launches_per_step = kernels per step. */
int nvc_runner_time_decode(nvc_runner* r, int iters, float* ms_per_step, int* launches_per_step);
/* Times ONE kernel family with the last decode step's inputs, over every layer's own weights /
 * KV pool (working set >> L2), `iters` passes, CUDA events on the runner's stream.
 * *ms_per_launch = average duration of one call; *launches = kernels launched in total. */
enum {
  NVC_OP_PAGED_DECODE = 0, NVC_OP_GEMM_QKV = 1, NVC_OP_GEMM_O = 2, NVC_OP_GEMM_GATE_UP = 3, NVC_OP_GEMM_DOWN = 4,
  NVC_OP_LM_HEAD = 5, NVC_OP_RMSNORM = 6, NVC_OP_QK_NORM_ROPE = 7, NVC_OP_SILU_MUL = 8, NVC_OP_ARGMAX = 9,
  NVC_OP_NOOP = 10, NVC_OP_ALLREDUCE = 11
};
int nvc_runner_time_op(nvc_runner* r, int op, int iters, float* ms_per_launch, int* launches);
/* The options the runner was created with (nvc_engine_create checks its scheduler config against them). */
int nvc_runner_get_options(const nvc_runner* r, nvc_runner_options* out);
/* Host-to-device bytes the last step copied (ids / positions / slots / context lens / block-table rows). */
int64_t nvc_runner_h2d_bytes_last_step(const nvc_runner* r);
int64_t nvc_runner_weight_bytes(const nvc_runner* r);
int64_t nvc_runner_kv_bytes_per_token(const nvc_runner* r);

/* ------------------------------------------------------------------ engine */

/* SchedulerConfig (src/engine/scheduler.rs:10-56). */
typedef struct nvc_scheduler_config {
  int32_t max_num_seqs;
  int32_t max_num_batched_tokens;
  int32_t eos;
  int32_t num_kvcache_blocks;
  int32_t kvcache_block_size;
  /* Beyond the reference; 0 = the reference's behaviour. */
  int32_t enable_prefix_caching;   /* 8f.1: hash-chained dedup of full KV blocks (upstream nano-vllm's
                                      BlockManager, which block_manager.rs:24-29 says it will mirror)     */
  int32_t enable_chunked_prefill;  /* 8f.2: one step = the decode tokens of the running sequences + as many prompt
                                      tokens of waiting sequences as the token budget allows; long prompts are
                                      split over steps; max_num_seqs bounds ALL running sequences and the budget
                                      counts only tokens actually computed (the two scheduler.rs quirks fixed).
                                      Needs max_num_batched_tokens >= max_num_seqs.                             */
  int32_t max_model_len;           /* longest sequence (prompt + completion) the runner holds; a sequence that reaches it
                                      finishes like one that reached max_tokens, a longer prompt is refused by
                                      nvc_engine_add_request.  0 = take the runner's (no limit without a runner).    */
  int32_t reserved[1];
} nvc_scheduler_config;

typedef struct nvc_schedule_info {
  int32_t is_prefill;
  int32_t num_seqs;
  int32_t num_tokens;     /* prefill: new tokens in the step; decode: num_seqs */
  int32_t max_blocks;     /* width of block_tables                             */
  int32_t max_seqlen;
  int32_t num_decode_seqs; /* chunked mode: the first num_decode_seqs sequences have one query token each and
                              go through paged decode attention; without chunked prefill: num_seqs of a decode
                              step, 0 of a prefill step                                                        */
} nvc_schedule_info;

/* StepOutput (src/engine/llm_engine.rs:328-333). */
typedef struct nvc_step_output {
  int32_t is_prefill;
  int32_t num_tokens;     /* reference convention: prefill = sum of seq lens AFTER the append, decode = -num_seqs */
  int32_t num_seqs;
  int32_t num_finished;   /* sequences that finished in this step */
  int32_t num_query_tokens;  /* tokens computed in this step (sum of q_len)                                    */
  int32_t num_sampled;       /* sequences that received a token (chunked mode: not the unfinished chunks)      */
  int32_t num_decode_seqs;   /* leading one-token sequences of the step (= num_seqs of a decode step)          */
} nvc_step_output;

/* runner may be NULL: the engine then behaves like DummyModelRunner (next = last_token + 1,
 * src/engine/llm_engine.rs:20-33).  With a runner the scheduler config must agree with the runner's options:
 * the block size must be equal and num_kvcache_blocks must not exceed the runner's pool (NVC_EINVAL otherwise: a
 * mismatch would read the wrong KV rows); num_kvcache_blocks == 0 and max_model_len == 0 take the runner's values;
 * max_num_seqs / max_num_batched_tokens / max_model_len are clamped to what the runner can execute.
 * When a runner step fails, nvc_engine_step returns its status AFTER rolling the batch back: its sequences go back
 * to the waiting queue as if preempted and the prefix-cache entries of their blocks are dropped (their K / V were
 * never written), so the caller may retry or continue. */
int nvc_engine_create(const nvc_scheduler_config* cfg, nvc_runner* runner, nvc_engine** out);
int nvc_engine_destroy(nvc_engine* e);
const char* nvc_engine_last_error(const nvc_engine* e);
/* LLMEngine::add_request (:228-237).  Returns the new seq_id (>= 0) or a negative status (NVC_EINVAL for an
 * empty prompt, max_tokens < 0, or a temperature the reference's SamplingParams refuses). */
int64_t nvc_engine_add_request(nvc_engine* e, const int32_t* token_ids, int n, double temperature,
                               int max_tokens, int ignore_eos);
/* Scheduler::schedule (:106-111) + step-input build (N1/N2).  Fails with NVC_ESTATE when idle. */
int nvc_engine_schedule(nvc_engine* e, nvc_schedule_info* info);
/* Views of the inputs built by the last nvc_engine_schedule (valid until the next one). */
int nvc_engine_step_inputs(const nvc_engine* e, const int32_t** ids, const int32_t** positions,
                           const int32_t** cu_seqlens, const int32_t** slot_mapping,
                           const int32_t** block_tables, const int32_t** context_lens, const int64_t** seq_ids);
/* Chunked mode: emit[b] != 0 when sequence b of the last schedule ends its pending tokens in this
 * step and takes the sampled token (always 1 without chunked prefill). */
int nvc_engine_step_emit_mask(const nvc_engine* e, const int32_t** emit);
/* Scheduler::post_process (:213-249) for the batch of the last schedule. */
int nvc_engine_post_process(nvc_engine* e, const int32_t* token_ids, int n);
/* LLMEngine::step (:239-264): schedule -> runner -> post_process. */
int nvc_engine_step(nvc_engine* e, nvc_step_output* out);
int nvc_engine_is_finished(const nvc_engine* e);
/* enabled != 0: every nvc_engine_step samples with each sequence's SamplingParams.temperature
 * (sampling_params.rs:6) on device, noise stream = (seq_id << 32 | completion tokens so far), so
 * a sequence's tokens do not depend on batching, preemption or tensor-parallel size.  enabled == 0
 * (default): greedy argmax, the mode the parity tests and the benchmark use. */
int nvc_engine_set_sampling(nvc_engine* e, int enabled, uint64_t seed);
/* Scheduler::preempt on the running sequence `seq_id` (:200-211). */
int nvc_engine_preempt(nvc_engine* e, int64_t seq_id);
/* Finished outputs (GenerationOutput, :192-207): completion tokens of seq_id. */
int nvc_engine_num_finished(const nvc_engine* e);
int nvc_engine_get_output(const nvc_engine* e, int64_t seq_id, int32_t* tokens, int capacity, int* n_tokens);
/* Introspection for the contract tests. */
int nvc_engine_seq_block_table(const nvc_engine* e, int64_t seq_id, int32_t* blocks, int capacity, int* n_blocks);
int nvc_engine_num_free_blocks(const nvc_engine* e);
/* Prefix cache counters in blocks: full blocks looked up / found by BlockManager::allocate. */
int nvc_engine_prefix_cache_stats(const nvc_engine* e, int64_t* hit_blocks, int64_t* queried_blocks);
int nvc_engine_num_waiting(const nvc_engine* e);
int nvc_engine_num_running(const nvc_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* NVC_B200_ENGINE_H */
