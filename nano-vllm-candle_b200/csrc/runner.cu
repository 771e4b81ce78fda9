// Qwen3 step executor: owns this rank's weight shard, the per-rank paged KV pool, the
// activation buffers and the decode CUDA graphs, and runs one engine step from flat step
// inputs.  Replaces Qwen3ModelRunner::run (src/engine/llm_engine.rs:146-189) and the model
// forward it drives (src/models/qwen3.rs:202-281,323-327,374-399,458-499,542-550).
//
// Per layer (prefill and decode differ only in the attention kernel):
//   fused_add_rmsnorm -> qkv GEMM -> qk_norm_rope_store (one kernel from 8 192 rows up: nvc_qkv_proj_norm_rope)
//   -> attention -> o_proj GEMM [-> all-reduce]
//   -> fused_add_rmsnorm -> gate_up GEMM (+ SiLU-mul epilogue) -> down GEMM [-> all-reduce]
// then: gather last rows -> final fused_add_rmsnorm -> lm_head GEMM (fp32) -> argmax.
#include "common.cuh"
#include "internal.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <vector>

namespace {

struct Layer {
  nvc_bf16 *w_qkv = nullptr, *w_o = nullptr, *q_norm = nullptr, *k_norm = nullptr;
  nvc_bf16 *w_gu = nullptr, *w_down = nullptr, *ln1 = nullptr, *ln2 = nullptr;
  nvc_bf16 *k_cache = nullptr, *v_cache = nullptr;
};

struct ArgPair { float val; int32_t idx; };

}  // namespace

struct nvc_runner {
  nvc_ctx* ctx = nullptr;
  nvc_qwen3_config cfg{};
  nvc_runner_options opt{};
  int tp = 1, rank = 0;
  int nh_l = 0, kv_l = 0, I_l = 0, V_l = 0, qkv_l = 0, max_blocks_cap = 0;
  bool finalized = false;
  cudaStream_t stream = nullptr;
  std::vector<Layer> layers;
  nvc_bf16 *embed = nullptr, *final_norm = nullptr, *lm_head = nullptr;
  float* rope = nullptr;
  std::set<std::string> loaded;
  int64_t weight_bytes = 0;
  std::vector<void*> allocs;
  // activations
  nvc_bf16 *x = nullptr, *xn = nullptr, *qkv = nullptr, *attn = nullptr, *act = nullptr;
  float* residual = nullptr;
  nvc_bf16 *x_last = nullptr, *xn_last = nullptr;
  float *res_last = nullptr, *logits = nullptr;
  void* decode_ws = nullptr;
  size_t decode_ws_bytes = 0;
  // step inputs on device + pinned staging
  int32_t* d_cu2 = nullptr;     // mixed step: cu_seqlens of the prompt chunks, relative to their first token
  int32_t *d_ids = nullptr, *d_pos = nullptr, *d_cu = nullptr, *d_slots = nullptr, *d_bt = nullptr, *d_ctx = nullptr,
          *d_last_rows = nullptr, *d_tok = nullptr;
  float* d_tok_val = nullptr;
  ArgPair* d_pairs = nullptr;
  int32_t* h_stage = nullptr;   // pinned
  size_t h_stage_ints = 0;
  size_t in_pitch = 0;             // ints between d_ids / d_pos / d_slots / d_ctx (one block: a decode step uploads all four in ONE 2-D copy)
  std::vector<int32_t> bt_mirror;  // the padded block table the device holds (decode steps re-upload it only when it changed)
  int bt_mirror_B = -1;            // rows of bt_mirror that are valid; -1: unknown (a prefill / mixed step wrote d_bt)
  ArgPair* h_pairs = nullptr;   // pinned
  int32_t* h_tok = nullptr;     // pinned
  int last_B = 0;
  // on-device temperature sampling (K8s), armed per step by nvc_runner_set_sampling
  float* d_temps = nullptr;
  uint64_t* d_streams = nullptr;
  float* h_temps = nullptr;        // pinned
  uint64_t* h_streams = nullptr;   // pinned
  bool sampling_armed = false;
  int last_sampling_B = 0;
  uint64_t sampling_seed = 0;
  // decode graphs keyed by (B, 2 * blocks hint + sampling)
  std::map<std::pair<int, int>, cudaGraphExec_t> graphs;
  int launches_per_decode = 0;
  int64_t last_h2d_bytes = 0;      // host-to-device bytes of the last step's inputs
  bool x_nvls = false;             // r->x is the context's NVLS slot: all-reduces go through the switch
  int last_decode_hint = 0;
  int last_decode_key2 = 0;
};

namespace {

#define R_CUDA(expr) NVC_CUDA_CHECK(r->ctx, expr)

template <typename T>
int dev_alloc(nvc_runner* r, T** p, size_t count) {
  void* q = nullptr;
  R_CUDA(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
  r->allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return NVC_OK;
}


// The transformer body over N tokens already embedded in r->x.  decode: attention over the
// paged cache; prefill: varlen causal flash over the tokens of this step.
// Under TP the partial sums of o_proj / down_proj are reduced either by NCCL (in place, then the
// fused add + RMSNorm kernel) or, for decode-sized messages when the NVLink peer buffers are set
// up, by ONE kernel that reads the peers' partials and applies residual add + RMSNorm
// (p2p_allreduce.cu).  In that mode the norm that opens layer l+1 is produced at the end of layer
// l, and the final norm by the last layer (*final_norm_done).
// paged_prefill with num_decode > 0 is a mixed step (chunked prefill): the first num_decode
// sequences are single tokens (rows 0 .. num_decode-1) and go through the paged decode kernel,
// the other sequences through the paged prefill kernel with their own cu_seqlens (r->d_cu2).
int run_layers(nvc_runner* r, int N, bool decode, int B, int max_seqlen, int blocks_hint, bool* final_norm_done,
               bool paged_prefill = false, int num_decode = 0) {
  const nvc_qwen3_config& c = r->cfg;
  nvc_ctx* ctx = r->ctx;
  void* st = r->stream;
  const int H = c.hidden_size, hd = c.head_dim, L = c.num_hidden_layers;
  const float scale = 1.0f / sqrtf((float)hd);   // (head_dim as f32).powf(-0.5), qwen3.rs:132
  static const bool p2p_env = !(getenv("NVC_P2P_AR") && atoi(getenv("NVC_P2P_AR")) == 0);
  // NVLS (in-switch) all-reduce + residual + norm for every step kind; else the peer-load kernel for decode-sized
  // messages; else NCCL + the norm kernel
  const bool nvls = r->tp > 1 && r->x_nvls && nvc_ctx_nvls_ready(ctx);
  const bool p2p = !nvls && r->tp > 1 && decode && p2p_env && nvc_ctx_p2p_ready(ctx) &&
                   (size_t)N * H * 2 <= nvc_ctx_p2p_slot_bytes(ctx);
  nvc_bf16* slot0 = p2p ? reinterpret_cast<nvc_bf16*>(nvc_ctx_p2p_slot(ctx, 0)) : nullptr;
  nvc_bf16* slot1 = p2p ? reinterpret_cast<nvc_bf16*>(nvc_ctx_p2p_slot(ctx, 1)) : nullptr;
  *final_norm_done = false;
  bool xn_ready = false;
  int rc;
  for (int li = 0; li < L; ++li) {
    Layer& Ly = r->layers[li];
    if (!xn_ready) {
      // input_layernorm with fused residual add; layer 0 seeds the residual stream (qwen3.rs:376-387)
      rc = nvc_fused_add_rmsnorm(ctx, r->x, li == 0 ? nullptr : r->residual, r->residual, Ly.ln1, c.rms_norm_eps, r->xn, N, H, st);
      if (rc) return rc;
    }
    xn_ready = false;
    // QKV projection + q/k-norm + RoPE + KV scatter: one kernel at prefill sizes, GEMM + K2 otherwise (same bits)
    rc = nvc_qkv_proj_norm_rope(ctx, r->xn, H, Ly.w_qkv, r->qkv, N, H, Ly.q_norm, Ly.k_norm, c.rms_norm_eps, r->rope, r->d_pos,
                                r->d_slots, Ly.k_cache, Ly.v_cache, r->nh_l, r->kv_l, hd, st);
    if (rc) return rc;
    if (decode) {
      rc = nvc_paged_decode_attn_ex(ctx, r->qkv, r->qkv_l, Ly.k_cache, Ly.v_cache, r->d_bt, r->max_blocks_cap, blocks_hint,
                                    r->d_ctx, B, r->nh_l, r->kv_l, hd, r->opt.kvcache_block_size, scale, r->attn,
                                    r->decode_ws, r->decode_ws_bytes, st);
    } else if (paged_prefill) {
      // cached prefix: K / V of the whole context come from the paged cache (this step's were just stored)
      rc = NVC_OK;
      if (num_decode > 0)
        rc = nvc_paged_decode_attn_ex(ctx, r->qkv, r->qkv_l, Ly.k_cache, Ly.v_cache, r->d_bt, r->max_blocks_cap, blocks_hint,
                                      r->d_ctx, num_decode, r->nh_l, r->kv_l, hd, r->opt.kvcache_block_size, scale, r->attn,
                                      r->decode_ws, r->decode_ws_bytes, st);
      if (rc == NVC_OK && B > num_decode)
        rc = nvc_prefill_attn_paged(ctx, r->qkv + (size_t)num_decode * r->qkv_l, r->qkv_l, Ly.k_cache, Ly.v_cache,
                                    r->d_bt + (size_t)num_decode * r->max_blocks_cap, r->max_blocks_cap, r->d_ctx + num_decode,
                                    num_decode ? r->d_cu2 : r->d_cu, B - num_decode, N - num_decode, max_seqlen, r->nh_l,
                                    r->kv_l, hd, r->opt.kvcache_block_size, r->opt.num_kvcache_blocks, scale,
                                    r->attn + (size_t)num_decode * r->nh_l * hd, st);
    } else {
      rc = nvc_prefill_attn_varlen(ctx, r->qkv, r->qkv_l, r->qkv + (size_t)r->nh_l * hd, r->qkv_l,
                                   r->qkv + (size_t)(r->nh_l + r->kv_l) * hd, r->qkv_l, r->d_cu, B, N, max_seqlen,
                                   r->nh_l, r->kv_l, hd, scale, r->attn, st);
    }
    if (rc) return rc;
    // o_proj (row-parallel) -> [all-reduce, N4] -> post_attention_layernorm with residual add
    rc = nvc_gemm_bf16(ctx, r->attn, r->nh_l * hd, Ly.w_o, p2p ? slot0 : r->x, H, NVC_DT_BF16, N, H, r->nh_l * hd, st);
    if (rc) return rc;
    if (nvls) {
      rc = nvc_nvls_allreduce_rmsnorm(ctx, r->x, r->residual, r->residual, Ly.ln2, c.rms_norm_eps, r->xn, N, H, 0, st);
    } else if (p2p) {
      rc = nvc_p2p_allreduce_rmsnorm(ctx, 0, r->residual, r->residual, Ly.ln2, c.rms_norm_eps, r->xn, N, H, st);
    } else {
      if (r->tp > 1) { rc = nvc_allreduce_bf16(ctx, r->x, (size_t)N * H, st); if (rc) return rc; }
      rc = nvc_fused_add_rmsnorm(ctx, r->x, r->residual, r->residual, Ly.ln2, c.rms_norm_eps, r->xn, N, H, st);
    }
    if (rc) return rc;
    // gate_up projection with SiluAndMul fused into the epilogue (w_gu rows interleaved in groups of 16 at load)
    rc = nvc_gemm_bf16_silu_mul(ctx, r->xn, H, Ly.w_gu, r->act, r->I_l, N, r->I_l, H, st);
    if (rc) return rc;
    // down_proj (row-parallel) -> [all-reduce, N4]
    rc = nvc_gemm_bf16(ctx, r->act, r->I_l, Ly.w_down, p2p ? slot1 : r->x, H, NVC_DT_BF16, N, H, r->I_l, st);
    if (rc) return rc;
    if (nvls) {
      // the norm that opens layer l+1 (or the final norm) comes out of the same kernel; a prefill step keeps all
      // N normed rows in xn and run_head gathers the last-token rows from there
      const bool last = li + 1 == L;
      rc = nvc_nvls_allreduce_rmsnorm(ctx, r->x, r->residual, r->residual, last ? r->final_norm : r->layers[li + 1].ln1,
                                      c.rms_norm_eps, last && decode ? r->xn_last : r->xn, N, H, 0, st);
      if (rc) return rc;
      xn_ready = !last;
      *final_norm_done = last;
    } else if (p2p) {
      const bool last = li + 1 == L;
      rc = nvc_p2p_allreduce_rmsnorm(ctx, 1, r->residual, r->residual, last ? r->final_norm : r->layers[li + 1].ln1,
                                     c.rms_norm_eps, last ? r->xn_last : r->xn, N, H, st);
      if (rc) return rc;
      xn_ready = !last;
      *final_norm_done = last;
    } else if (r->tp > 1) {
      rc = nvc_allreduce_bf16(ctx, r->x, (size_t)N * H, st);
      if (rc) return rc;
    }
  }
  return NVC_OK;
}

// last-token rows -> final norm -> lm_head (fp32 logits) -> argmax (vocab shard offset)
int run_head(nvc_runner* r, int B, bool decode, bool final_norm_done) {
  const nvc_qwen3_config& c = r->cfg;
  nvc_ctx* ctx = r->ctx;
  void* st = r->stream;
  const int H = c.hidden_size;
  int rc;
  const nvc_bf16* xs = r->x;
  const float* rs = r->residual;
  if (!decode && final_norm_done) {           // prefill through the fused all-reduce: xn holds the normed rows
    rc = nvc_gather_rows(ctx, r->xn, nullptr, r->d_last_rows, r->xn_last, nullptr, B, H, st);
    if (rc) return rc;
  } else if (!decode) {
    rc = nvc_gather_rows(ctx, r->x, r->residual, r->d_last_rows, r->x_last, r->res_last, B, H, st);
    if (rc) return rc;
    xs = r->x_last;
    rs = r->res_last;
  }
  if (!final_norm_done) {
    rc = nvc_fused_add_rmsnorm(ctx, xs, rs, nullptr, r->final_norm, c.rms_norm_eps, r->xn_last, B, H, st);   // qwen3.rs:497
    if (rc) return rc;
  }
  rc = nvc_gemm_bf16(ctx, r->xn_last, H, r->lm_head, r->logits, r->V_l, NVC_DT_F32, B, r->V_l, H, st);     // qwen3.rs:548
  if (rc) return rc;
  if (r->sampling_armed)   // llm_engine.rs:97-133, seeded and on device
    rc = nvc_sample_gumbel(ctx, r->logits, r->V_l, r->d_temps, r->d_streams, r->sampling_seed, r->d_tok, r->d_tok_val, B,
                           r->V_l, r->rank * r->V_l, st);
  else
    rc = nvc_argmax_last(ctx, r->logits, r->V_l, r->d_tok, r->d_tok_val, B, r->V_l, r->rank * r->V_l, st);    // llm_engine.rs:135-142
  if (rc || r->tp == 1) return rc;
  // TP: every rank holds (max, idx) of its vocab shard.  Through the switch: one kernel gathers and merges them
  // (d_tok then holds the winning tokens on every rank); else gather the 8-byte pairs with NCCL, merged on the host
  if (r->x_nvls && nvc_ctx_nvls_ready(ctx) && (size_t)B * r->tp * 16 <= nvc_ctx_nvls_ll_bytes(ctx))
    return nvc_nvls_argmax_merge(ctx, r->d_tok_val, r->d_tok, r->d_tok, nullptr, B, st);
  rc = nvc_pack_argmax_pairs(ctx, r->d_tok_val, r->d_tok, reinterpret_cast<float*>(r->d_pairs + (size_t)r->rank * B), B, st);
  if (rc) return rc;
  return nvc_allgather_bytes(ctx, r->d_pairs, (size_t)B * sizeof(ArgPair), st);
}

bool merged_on_device(const nvc_runner* r, int B) {
  return r->tp == 1 || (r->x_nvls && nvc_ctx_nvls_ready(r->ctx) && (size_t)B * r->tp * 16 <= nvc_ctx_nvls_ll_bytes(r->ctx));
}

// device -> pinned host copy of the step's result (capturable: the decode graph ends with it)
int enqueue_token_readback(nvc_runner* r, int B) {
  if (merged_on_device(r, B)) {
    R_CUDA(cudaMemcpyAsync(r->h_tok, r->d_tok, B * sizeof(int32_t), cudaMemcpyDeviceToHost, r->stream));
  } else {
    R_CUDA(cudaMemcpyAsync(r->h_pairs, r->d_pairs, (size_t)r->tp * B * sizeof(ArgPair), cudaMemcpyDeviceToHost, r->stream));
  }
  return NVC_OK;
}

// after the stream has drained
int collect_tokens(nvc_runner* r, int B, int32_t* out_tokens) {
  if (merged_on_device(r, B)) {
    memcpy(out_tokens, r->h_tok, B * sizeof(int32_t));
    return NVC_OK;
  }
  // TP without the switch: pick the largest value over the gathered shard maxima, ties -> the larger index
  // (= the reference's "last maximum" over the full row).
  for (int b = 0; b < B; ++b) {
    ArgPair best = r->h_pairs[b];
    for (int k = 1; k < r->tp; ++k) {
      const ArgPair p = r->h_pairs[(size_t)k * B + b];
      if (p.val > best.val || (p.val == best.val && p.idx > best.idx)) best = p;
    }
    out_tokens[b] = best.idx;
  }
  return NVC_OK;
}

int finish_tokens(nvc_runner* r, int B, int32_t* out_tokens) {
  int rc = enqueue_token_readback(r, B);
  if (rc) return rc;
  R_CUDA(cudaStreamSynchronize(r->stream));
  return collect_tokens(r, B, out_tokens);
}

int decode_body(nvc_runner* r, int B, int blocks_hint) {
  int rc = nvc_embed_gather(r->ctx, r->d_ids, r->embed, r->x, B, r->cfg.hidden_size, r->cfg.vocab_size, r->stream);
  if (rc) return rc;
  bool fnd = false;
  rc = run_layers(r, B, true, B, 1, blocks_hint, &fnd);
  if (rc) return rc;
  return run_head(r, B, true, fnd);
}

int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

// H2D of the armed step's temperatures / noise streams (no-op for a greedy step)
int stage_sampling(nvc_runner* r, int B) {
  if (!r->sampling_armed) return NVC_OK;
  NVC_REQUIRE(r->ctx, r->last_sampling_B == B, NVC_ESTATE, "nvc_runner_set_sampling was armed for a different batch size");
  R_CUDA(cudaMemcpyAsync(r->d_temps, r->h_temps, B * sizeof(float), cudaMemcpyHostToDevice, r->stream));
  R_CUDA(cudaMemcpyAsync(r->d_streams, r->h_streams, B * sizeof(uint64_t), cudaMemcpyHostToDevice, r->stream));
  return NVC_OK;
}

}  // namespace

extern "C" {

int nvc_runner_create(nvc_ctx* ctx, const nvc_qwen3_config* cfg, const nvc_runner_options* opt, nvc_runner** out) {
  NVC_REQUIRE(ctx, ctx && cfg && opt && out, NVC_EINVAL, "null pointer");
  *out = nullptr;
  const int tp = ctx->tp_size;
  NVC_REQUIRE(ctx, cfg->head_dim == 128, NVC_ESHAPE, "head_dim must be 128");
  NVC_REQUIRE(ctx, cfg->num_attention_heads % cfg->num_key_value_heads == 0, NVC_ESHAPE, "nh % kv != 0");
  NVC_REQUIRE(ctx, cfg->num_key_value_heads % tp == 0 && cfg->num_attention_heads % tp == 0, NVC_ESHAPE,
              "heads must divide by tp_size");
  NVC_REQUIRE(ctx, cfg->intermediate_size % (tp * 64) == 0, NVC_ESHAPE, "intermediate_size must divide by tp_size*64");
  NVC_REQUIRE(ctx, cfg->vocab_size % (tp * 16) == 0, NVC_ESHAPE, "vocab_size must divide by tp_size*16");
  NVC_REQUIRE(ctx, cfg->hidden_size % 64 == 0, NVC_ESHAPE, "hidden_size must be a multiple of 64");
  // the paged prefill kernel (prefix caching, mixed steps) reads a 64-token kv tile as 64 / block_size boxes or as one
  // box of a block that is a multiple of 128 tokens: apply its rule here, not at the first cached-prefix step
  NVC_REQUIRE(ctx, opt->kvcache_block_size == 16 || opt->kvcache_block_size == 32 || opt->kvcache_block_size == 64 ||
                       (opt->kvcache_block_size > 0 && opt->kvcache_block_size % 128 == 0), NVC_ESHAPE,
              "kvcache_block_size must be 16, 32, 64 or a multiple of 128");
  NVC_REQUIRE(ctx, opt->allreduce_f32 == 0, NVC_ENOTIMPL, "allreduce_f32 is not implemented");
  NVC_REQUIRE(ctx, opt->num_kvcache_blocks > 0 && opt->max_num_seqs > 0 && opt->max_num_batched_tokens > 0 &&
                       opt->max_model_len > 0, NVC_EINVAL, "bad runner options");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  nvc_runner* r = new nvc_runner();
  r->ctx = ctx;
  r->cfg = *cfg;
  r->opt = *opt;
  r->tp = tp;
  r->rank = ctx->tp_rank;
  r->nh_l = cfg->num_attention_heads / tp;
  r->kv_l = cfg->num_key_value_heads / tp;
  r->I_l = cfg->intermediate_size / tp;
  r->V_l = cfg->vocab_size / tp;
  r->qkv_l = (r->nh_l + 2 * r->kv_l) * cfg->head_dim;
  r->max_blocks_cap = (opt->max_model_len + opt->kvcache_block_size - 1) / opt->kvcache_block_size;
  *out = r;   // from here on the caller destroys it, also on failure
  int rc;
  R_CUDA(cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking));
  const int H = cfg->hidden_size, hd = cfg->head_dim;
  const size_t T = (size_t)std::max(opt->max_num_batched_tokens, opt->max_num_seqs);
  const size_t Bm = (size_t)opt->max_num_seqs;
  r->layers.resize(cfg->num_hidden_layers);
  const size_t cache_elems = (size_t)opt->num_kvcache_blocks * opt->kvcache_block_size * r->kv_l * hd;
  for (auto& L : r->layers) {
    if ((rc = dev_alloc(r, &L.w_qkv, (size_t)r->qkv_l * H))) return rc;
    if ((rc = dev_alloc(r, &L.w_o, (size_t)H * r->nh_l * hd))) return rc;
    if ((rc = dev_alloc(r, &L.q_norm, hd))) return rc;
    if ((rc = dev_alloc(r, &L.k_norm, hd))) return rc;
    if ((rc = dev_alloc(r, &L.w_gu, (size_t)2 * r->I_l * H))) return rc;
    if ((rc = dev_alloc(r, &L.w_down, (size_t)H * r->I_l))) return rc;
    if ((rc = dev_alloc(r, &L.ln1, H))) return rc;
    if ((rc = dev_alloc(r, &L.ln2, H))) return rc;
    if ((rc = dev_alloc(r, &L.k_cache, cache_elems))) return rc;
    if ((rc = dev_alloc(r, &L.v_cache, cache_elems))) return rc;
    r->weight_bytes += 2 * ((int64_t)r->qkv_l * H + (int64_t)H * r->nh_l * hd + 2 * hd + 3 * (int64_t)r->I_l * H + 2 * H);
  }
  if ((rc = dev_alloc(r, &r->embed, (size_t)cfg->vocab_size * H))) return rc;
  if ((rc = dev_alloc(r, &r->final_norm, H))) return rc;
  if ((rc = dev_alloc(r, &r->lm_head, (size_t)r->V_l * H))) return rc;
  r->weight_bytes += 2 * ((int64_t)cfg->vocab_size * H + H + (int64_t)r->V_l * H);
  if ((rc = dev_alloc(r, &r->rope, (size_t)opt->max_model_len * hd))) return rc;
  // under TP the row-parallel GEMMs write their partial sums where the switch can reduce them: x is the context's
  // NVLS slot when the multicast set-up succeeds (collective: every rank takes the same branch)
  static const bool nvls_env = !(getenv("NVC_NVLS") && atoi(getenv("NVC_NVLS")) == 0);
  if (tp > 1 && nvls_env && !ctx->tp_group_name.empty()) {
    const size_t slot = (T * H * 2 + 255) / 256 * 256;
    const size_t ll = (size_t)4 << 20;
    if (!(nvc_ctx_nvls_ready(ctx) && nvc_ctx_nvls_slot_bytes(ctx) >= slot && nvc_ctx_nvls_ll_bytes(ctx) >= ll))
      (void)nvc_ctx_nvls_init(ctx, slot, ll, nullptr);       // on refusal: nvc_ctx_nvls_status, and the paths below
    if (nvc_ctx_nvls_ready(ctx)) {
      r->x = reinterpret_cast<nvc_bf16*>(nvc_ctx_nvls_slot(ctx));
      r->x_nvls = true;
    }
  }
  if (!r->x_nvls && (rc = dev_alloc(r, &r->x, T * H))) return rc;
  if ((rc = dev_alloc(r, &r->xn, T * H))) return rc;
  if ((rc = dev_alloc(r, &r->residual, T * H))) return rc;
  if ((rc = dev_alloc(r, &r->qkv, T * r->qkv_l))) return rc;
  if ((rc = dev_alloc(r, &r->attn, T * r->nh_l * hd))) return rc;
  if ((rc = dev_alloc(r, &r->act, T * r->I_l))) return rc;
  if ((rc = dev_alloc(r, &r->x_last, Bm * H))) return rc;
  if ((rc = dev_alloc(r, &r->xn_last, Bm * H))) return rc;
  if ((rc = dev_alloc(r, &r->res_last, Bm * H))) return rc;
  if ((rc = dev_alloc(r, &r->logits, Bm * r->V_l))) return rc;
  r->decode_ws_bytes = 0;
  for (int b = 1; b <= opt->max_num_seqs; ++b)
    r->decode_ws_bytes = std::max(r->decode_ws_bytes, nvc_paged_decode_workspace_bytes(b, r->nh_l, hd, ctx->num_sms));
  if ((rc = dev_alloc(r, reinterpret_cast<uint8_t**>(&r->decode_ws), r->decode_ws_bytes))) return rc;
  {
    r->in_pitch = (std::max(T, Bm) + 3) / 4 * 4;
    int32_t* blk = nullptr;
    if ((rc = dev_alloc(r, &blk, 4 * r->in_pitch))) return rc;
    r->d_ids = blk;
    r->d_pos = blk + r->in_pitch;
    r->d_slots = blk + 2 * r->in_pitch;
    r->d_ctx = blk + 3 * r->in_pitch;
  }
  if ((rc = dev_alloc(r, &r->d_cu, Bm + 1))) return rc;
  if ((rc = dev_alloc(r, &r->d_cu2, Bm + 1))) return rc;
  if ((rc = dev_alloc(r, &r->d_bt, Bm * r->max_blocks_cap))) return rc;
  if ((rc = dev_alloc(r, &r->d_last_rows, Bm))) return rc;
  if ((rc = dev_alloc(r, &r->d_tok, Bm))) return rc;
  if ((rc = dev_alloc(r, &r->d_tok_val, Bm))) return rc;
  if ((rc = dev_alloc(r, &r->d_pairs, Bm * tp))) return rc;
  if ((rc = dev_alloc(r, &r->d_temps, Bm))) return rc;
  if ((rc = dev_alloc(r, &r->d_streams, Bm))) return rc;
  R_CUDA(cudaMallocHost(&r->h_temps, Bm * sizeof(float)));
  R_CUDA(cudaMallocHost(&r->h_streams, Bm * sizeof(uint64_t)));
  r->h_stage_ints = 3 * T + 2 * (Bm + 1) + Bm * r->max_blocks_cap + 2 * Bm;
  R_CUDA(cudaMallocHost(&r->h_stage, r->h_stage_ints * sizeof(int32_t)));
  r->bt_mirror.assign(Bm * r->max_blocks_cap, -1);
  R_CUDA(cudaMallocHost(&r->h_pairs, Bm * tp * sizeof(ArgPair)));
  R_CUDA(cudaMallocHost(&r->h_tok, Bm * sizeof(int32_t)));
  return NVC_OK;
}

int nvc_runner_destroy(nvc_runner* r) {
  if (!r) return NVC_EINVAL;
  cudaSetDevice(r->ctx->device);
  if (r->stream) cudaStreamSynchronize(r->stream);
  for (auto& g : r->graphs) cudaGraphExecDestroy(g.second);
  for (void* p : r->allocs) cudaFree(p);
  if (r->h_stage) cudaFreeHost(r->h_stage);
  if (r->h_pairs) cudaFreeHost(r->h_pairs);
  if (r->h_tok) cudaFreeHost(r->h_tok);
  if (r->h_temps) cudaFreeHost(r->h_temps);
  if (r->h_streams) cudaFreeHost(r->h_streams);
  if (r->stream) cudaStreamDestroy(r->stream);
  delete r;
  return NVC_OK;
}

int64_t nvc_runner_weight_bytes(const nvc_runner* r) { return r ? r->weight_bytes : 0; }
int64_t nvc_runner_h2d_bytes_last_step(const nvc_runner* r) { return r ? r->last_h2d_bytes : 0; }
int nvc_runner_get_options(const nvc_runner* r, nvc_runner_options* out) {
  if (!r || !out) return NVC_EINVAL;
  *out = r->opt;
  return NVC_OK;
}
int64_t nvc_runner_kv_bytes_per_token(const nvc_runner* r) {
  return r ? (int64_t)2 * r->cfg.num_hidden_layers * r->kv_l * r->cfg.head_dim * 2 : 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// weights
// ------------------------------------------------------------------------------------------
namespace {

int copy_rows(nvc_runner* r, nvc_bf16* dst, const nvc_bf16* src, int64_t row0, int64_t nrows, int64_t cols) {
  R_CUDA(cudaMemcpy(dst, src + row0 * cols, (size_t)nrows * cols * 2, cudaMemcpyDefault));   // host or device source
  return NVC_OK;
}
// dst[rows, ncols] = src[rows, col0 : col0+ncols] of a [rows, cols] matrix
int copy_cols(nvc_runner* r, nvc_bf16* dst, const nvc_bf16* src, int64_t rows, int64_t cols, int64_t col0, int64_t ncols) {
  R_CUDA(cudaMemcpy2D(dst, (size_t)ncols * 2, src + col0, (size_t)cols * 2, (size_t)ncols * 2, (size_t)rows,
                      cudaMemcpyDefault));
  return NVC_OK;
}

// dst rows [32g, 32g+16) = src rows [16g, 16g+16): groups of 16 rows, one group skipped between them
int copy_rows_grouped(nvc_runner* r, nvc_bf16* dst, const nvc_bf16* src, int64_t nrows, int64_t cols) {
  R_CUDA(cudaMemcpy2D(dst, (size_t)32 * cols * 2, src, (size_t)16 * cols * 2, (size_t)16 * cols * 2, (size_t)(nrows / 16),
                      cudaMemcpyDefault));
  return NVC_OK;
}

}  // namespace

extern "C" int nvc_runner_load_weight(nvc_runner* r, const char* hf_name, const nvc_bf16* data, int64_t rows,
                                      int64_t cols) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  NVC_REQUIRE(ctx, hf_name && data, NVC_EINVAL, "null pointer");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  const nvc_qwen3_config& c = r->cfg;
  const int64_t H = c.hidden_size, hd = c.head_dim, V = c.vocab_size, I = c.intermediate_size;
  const int64_t nh = c.num_attention_heads, kv = c.num_key_value_heads;
  const std::string name(hf_name);
  auto shape_is = [&](int64_t rr, int64_t cc) { return rows == rr && cols == cc; };
  auto bad_shape = [&]() {
    nvc_set_error(ctx, "nvc_runner_load_weight: unexpected shape [" + std::to_string(rows) + "," +
                           std::to_string(cols) + "] for " + name);
    return NVC_ESHAPE;
  };
  int rc = NVC_OK;
  if (name == "model.embed_tokens.weight") {
    if (!shape_is(V, H)) return bad_shape();
    rc = copy_rows(r, r->embed, data, 0, V, H);
  } else if (name == "model.norm.weight") {
    if (!shape_is(H, 1)) return bad_shape();
    rc = copy_rows(r, r->final_norm, data, 0, H, 1);
  } else if (name == "lm_head.weight") {
    if (!shape_is(V, H)) return bad_shape();
    rc = copy_rows(r, r->lm_head, data, (int64_t)r->rank * r->V_l, r->V_l, H);   // vocab (column-parallel) shard
  } else if (name.rfind("model.layers.", 0) == 0) {
    const size_t p0 = strlen("model.layers.");
    const size_t p1 = name.find('.', p0);
    NVC_REQUIRE(ctx, p1 != std::string::npos, NVC_EINVAL, "malformed tensor name " + name);
    const int li = atoi(name.substr(p0, p1 - p0).c_str());
    NVC_REQUIRE(ctx, li >= 0 && li < c.num_hidden_layers, NVC_EINVAL, "layer index out of range in " + name);
    Layer& L = r->layers[li];
    const std::string sub = name.substr(p1 + 1);
    const int64_t qrows = (int64_t)r->nh_l * hd, kvrows = (int64_t)r->kv_l * hd;
    if (sub == "self_attn.q_proj.weight") {            // fused rows [q; k; v], qwen3.rs:171
      if (!shape_is(nh * hd, H)) return bad_shape();
      rc = copy_rows(r, L.w_qkv, data, r->rank * qrows, qrows, H);
    } else if (sub == "self_attn.k_proj.weight") {
      if (!shape_is(kv * hd, H)) return bad_shape();
      rc = copy_rows(r, L.w_qkv + qrows * H, data, r->rank * kvrows, kvrows, H);
    } else if (sub == "self_attn.v_proj.weight") {
      if (!shape_is(kv * hd, H)) return bad_shape();
      rc = copy_rows(r, L.w_qkv + (qrows + kvrows) * H, data, r->rank * kvrows, kvrows, H);
    } else if (sub == "self_attn.o_proj.weight") {     // row-parallel: shard the INPUT dim
      if (!shape_is(H, nh * hd)) return bad_shape();
      rc = copy_cols(r, L.w_o, data, H, nh * hd, r->rank * qrows, qrows);
    } else if (sub == "self_attn.q_norm.weight") {
      if (!shape_is(hd, 1)) return bad_shape();
      rc = copy_rows(r, L.q_norm, data, 0, hd, 1);
    } else if (sub == "self_attn.k_norm.weight") {
      if (!shape_is(hd, 1)) return bad_shape();
      rc = copy_rows(r, L.k_norm, data, 0, hd, 1);
    } else if (sub == "mlp.gate_proj.weight") {        // fused gate_up (qwen3.rs:310), rows interleaved in groups of 16:
      if (!shape_is(I, H)) return bad_shape();         // [32g, 32g+16) = gate rows, [32g+16, 32g+32) = up rows (nvc_gemm_bf16_silu_mul)
      rc = copy_rows_grouped(r, L.w_gu, data + (int64_t)r->rank * r->I_l * H, r->I_l, H);
    } else if (sub == "mlp.up_proj.weight") {
      if (!shape_is(I, H)) return bad_shape();
      rc = copy_rows_grouped(r, L.w_gu + 16 * H, data + (int64_t)r->rank * r->I_l * H, r->I_l, H);
    } else if (sub == "mlp.down_proj.weight") {
      if (!shape_is(H, I)) return bad_shape();
      rc = copy_cols(r, L.w_down, data, H, I, (int64_t)r->rank * r->I_l, r->I_l);
    } else if (sub == "input_layernorm.weight") {
      if (!shape_is(H, 1)) return bad_shape();
      rc = copy_rows(r, L.ln1, data, 0, H, 1);
    } else if (sub == "post_attention_layernorm.weight") {
      if (!shape_is(H, 1)) return bad_shape();
      rc = copy_rows(r, L.ln2, data, 0, H, 1);
    } else {
      NVC_REQUIRE(ctx, false, NVC_EINVAL, "unknown tensor " + name);
    }
  } else {
    NVC_REQUIRE(ctx, false, NVC_EINVAL, "unknown tensor " + name);
  }
  if (rc == NVC_OK) r->loaded.insert(name);
  return rc;
}

extern "C" int nvc_runner_finalize(nvc_runner* r) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  const size_t want = 3 + 11 * (size_t)r->cfg.num_hidden_layers;
  NVC_REQUIRE(ctx, r->loaded.size() == want, NVC_ESTATE,
              "missing weights: " + std::to_string(r->loaded.size()) + " of " + std::to_string(want) +
                  " tensors loaded (nvc_runner_load_weight takes lm_head.weight explicitly; nvc_runner_load_hf_dir ties it to embed_tokens when absent)");
  int rc = nvc_rope_table_init_sync(ctx, r->cfg.rope_theta, r->cfg.head_dim, r->opt.max_model_len, r->rope);
  if (rc) return rc;
  r->finalized = true;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------------
// steps
// ------------------------------------------------------------------------------------------
namespace {

// block_tables == nullptr: no cached prefix, attention over this step's own k / v (K4);
// otherwise attention over the paged cache through block_tables / context_lens (K4 paged).
int prefill_impl(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                 const int32_t* slot_mapping, const int32_t* block_tables, int max_blocks, const int32_t* context_lens,
                 int B, int N, int max_seqlen, int32_t* out_tokens, int num_decode = 0) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  const bool paged = block_tables != nullptr;
  NVC_REQUIRE(ctx, r->finalized, NVC_ESTATE, "nvc_runner_finalize not called");
  NVC_REQUIRE(ctx, ids && positions && cu_seqlens && slot_mapping && out_tokens, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B > 0 && N > 0 && B <= r->opt.max_num_seqs && N <= r->opt.max_num_batched_tokens, NVC_ENOMEM,
              "prefill step exceeds max_num_seqs / max_num_batched_tokens");
  NVC_REQUIRE(ctx, cu_seqlens[0] == 0 && cu_seqlens[B] == N, NVC_EINVAL, "cu_seqlens does not cover N tokens");
  for (int b = 0; b < B; ++b)   // strictly increasing from 0 to N: every entry is a valid row index below
    NVC_REQUIRE(ctx, cu_seqlens[b + 1] > cu_seqlens[b] && cu_seqlens[b + 1] - cu_seqlens[b] <= max_seqlen, NVC_EINVAL,
                "bad cu_seqlens / max_seqlen");
  const int bs = r->opt.kvcache_block_size;
  const int64_t nslots = (int64_t)r->opt.num_kvcache_blocks * bs;
  for (int i = 0; i < N; ++i) {
    NVC_REQUIRE(ctx, ids[i] >= 0 && ids[i] < r->cfg.vocab_size, NVC_EINVAL, "token id out of range");
    NVC_REQUIRE(ctx, positions[i] >= 0 && positions[i] < r->opt.max_model_len, NVC_EINVAL, "position out of range");
    NVC_REQUIRE(ctx, slot_mapping[i] >= -1 && slot_mapping[i] < nslots, NVC_EINVAL, "slot out of range");
  }
  const int cap = r->max_blocks_cap;
  NVC_REQUIRE(ctx, num_decode >= 0 && num_decode <= B && (paged || num_decode == 0), NVC_EINVAL, "bad num_decode");
  for (int b = 0; b < num_decode; ++b)
    NVC_REQUIRE(ctx, cu_seqlens[b] == b && cu_seqlens[b + 1] == b + 1 && slot_mapping[b] >= 0, NVC_EINVAL,
                "the first num_decode sequences must have exactly one query token and a slot");
  if (paged) {
    NVC_REQUIRE(ctx, context_lens, NVC_EINVAL, "null pointer");
    NVC_REQUIRE(ctx, max_blocks > 0 && max_blocks <= cap, NVC_ENOMEM, "block table wider than max_model_len allows");
    for (int b = 0; b < B; ++b) {
      const int qlen = cu_seqlens[b + 1] - cu_seqlens[b];
      NVC_REQUIRE(ctx, context_lens[b] >= qlen && (int64_t)context_lens[b] <= (int64_t)max_blocks * bs, NVC_EINVAL,
                  "context_len vs q_len / block table");
      const int nb = (context_lens[b] + bs - 1) / bs;
      NVC_REQUIRE(ctx, qlen <= 0 || positions[cu_seqlens[b + 1] - 1] == context_lens[b] - 1, NVC_EINVAL,
                  "the new tokens must be the last positions of the context");
      for (int j = 0; j < nb; ++j)
        NVC_REQUIRE(ctx, block_tables[(size_t)b * max_blocks + j] >= 0 &&
                             block_tables[(size_t)b * max_blocks + j] < r->opt.num_kvcache_blocks, NVC_EINVAL, "block id out of range");
    }
  }
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  int32_t* h = r->h_stage;
  memcpy(h, ids, N * 4);
  memcpy(h + N, positions, N * 4);
  memcpy(h + 2 * N, slot_mapping, N * 4);
  memcpy(h + 3 * N, cu_seqlens, (B + 1) * 4);
  int32_t* last = h + 3 * N + B + 1;
  for (int b = 0; b < B; ++b) last[b] = cu_seqlens[b + 1] - 1;
  cudaStream_t st = r->stream;
  R_CUDA(cudaMemcpyAsync(r->d_ids, h, N * 4, cudaMemcpyHostToDevice, st));
  R_CUDA(cudaMemcpyAsync(r->d_pos, h + N, N * 4, cudaMemcpyHostToDevice, st));
  R_CUDA(cudaMemcpyAsync(r->d_slots, h + 2 * N, N * 4, cudaMemcpyHostToDevice, st));
  R_CUDA(cudaMemcpyAsync(r->d_cu, h + 3 * N, (B + 1) * 4, cudaMemcpyHostToDevice, st));
  R_CUDA(cudaMemcpyAsync(r->d_last_rows, last, B * 4, cudaMemcpyHostToDevice, st));
  r->last_h2d_bytes = (int64_t)N * 12 + (int64_t)(B + 1) * 4 + (int64_t)B * 4 +
                      (paged ? (int64_t)B * 4 + (int64_t)B * r->max_blocks_cap * 4 : 0);
  if (paged) {
    int32_t* hctx = last + B;
    int32_t* hbt = hctx + B;
    memcpy(hctx, context_lens, B * 4);
    for (int b = 0; b < B; ++b) {
      memcpy(hbt + (size_t)b * cap, block_tables + (size_t)b * max_blocks, max_blocks * 4);
      for (int j = max_blocks; j < cap; ++j) hbt[(size_t)b * cap + j] = -1;
    }
    R_CUDA(cudaMemcpyAsync(r->d_ctx, hctx, B * 4, cudaMemcpyHostToDevice, st));
    r->bt_mirror_B = -1;           // the decode path's cached copy no longer describes d_bt
    R_CUDA(cudaMemcpyAsync(r->d_bt, hbt, (size_t)B * cap * 4, cudaMemcpyHostToDevice, st));
    if (num_decode > 0 && num_decode < B) {
      int32_t* hcu2 = hbt + (size_t)B * cap;
      for (int b = num_decode; b <= B; ++b) hcu2[b - num_decode] = cu_seqlens[b] - num_decode;
      R_CUDA(cudaMemcpyAsync(r->d_cu2, hcu2, (B - num_decode + 1) * 4, cudaMemcpyHostToDevice, st));
    }
  }
  int hint = 0, max_seqlen_chunks = max_seqlen;
  if (num_decode > 0) {
    int mb = 1;
    for (int b = 0; b < num_decode; ++b) mb = std::max(mb, (context_lens[b] + bs - 1) / bs);
    hint = std::min(next_pow2(mb), cap);
    max_seqlen_chunks = 1;
    for (int b = num_decode; b < B; ++b) max_seqlen_chunks = std::max(max_seqlen_chunks, cu_seqlens[b + 1] - cu_seqlens[b]);
  }
  int rc = stage_sampling(r, B);
  if (rc) return rc;
  rc = nvc_embed_gather(ctx, r->d_ids, r->embed, r->x, N, r->cfg.hidden_size, r->cfg.vocab_size, st);
  if (rc) return rc;
  bool fnd = false;
  rc = run_layers(r, N, false, B, max_seqlen_chunks, hint, &fnd, paged, num_decode);
  if (rc) return rc;
  rc = run_head(r, B, false, fnd);
  r->sampling_armed = false;
  if (rc) return rc;
  r->last_B = B;
  return finish_tokens(r, B, out_tokens);
}

}  // namespace

extern "C" int nvc_runner_prefill(nvc_runner* r, const int32_t* ids, const int32_t* positions,
                                  const int32_t* cu_seqlens, const int32_t* slot_mapping, int B, int N, int max_seqlen,
                                  int32_t* out_tokens) {
  return prefill_impl(r, ids, positions, cu_seqlens, slot_mapping, nullptr, 0, nullptr, B, N, max_seqlen, out_tokens);
}

extern "C" int nvc_runner_prefill_paged(nvc_runner* r, const int32_t* ids, const int32_t* positions,
                                        const int32_t* cu_seqlens, const int32_t* slot_mapping,
                                        const int32_t* block_tables, int max_blocks, const int32_t* context_lens, int B,
                                        int N, int max_seqlen, int32_t* out_tokens) {
  if (r && !block_tables) {
    nvc_set_error(r->ctx, "nvc_runner_prefill_paged: null block_tables");
    return NVC_EINVAL;
  }
  return prefill_impl(r, ids, positions, cu_seqlens, slot_mapping, block_tables, max_blocks, context_lens, B, N,
                      max_seqlen, out_tokens);
}

extern "C" int nvc_runner_step_mixed(nvc_runner* r, const int32_t* ids, const int32_t* positions,
                                     const int32_t* cu_seqlens, const int32_t* slot_mapping, const int32_t* block_tables,
                                     int max_blocks, const int32_t* context_lens, int num_decode, int B, int N,
                                     int max_seqlen, int32_t* out_tokens) {
  if (r && !block_tables) {
    nvc_set_error(r->ctx, "nvc_runner_step_mixed: null block_tables");
    return NVC_EINVAL;
  }
  return prefill_impl(r, ids, positions, cu_seqlens, slot_mapping, block_tables, max_blocks, context_lens, B, N,
                      max_seqlen, out_tokens, num_decode);
}

extern "C" int nvc_runner_decode(nvc_runner* r, const int32_t* ids, const int32_t* positions,
                                 const int32_t* slot_mapping, const int32_t* block_tables, int max_blocks,
                                 const int32_t* context_lens, int B, int32_t* out_tokens) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  NVC_REQUIRE(ctx, r->finalized, NVC_ESTATE, "nvc_runner_finalize not called");
  NVC_REQUIRE(ctx, ids && positions && slot_mapping && block_tables && context_lens && out_tokens, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B > 0 && B <= r->opt.max_num_seqs, NVC_ENOMEM, "decode batch exceeds max_num_seqs");
  NVC_REQUIRE(ctx, max_blocks > 0 && max_blocks <= r->max_blocks_cap, NVC_ENOMEM, "block table wider than max_model_len allows");
  const int bs = r->opt.kvcache_block_size;
  const int64_t nslots = (int64_t)r->opt.num_kvcache_blocks * bs;
  for (int b = 0; b < B; ++b) {
    NVC_REQUIRE(ctx, ids[b] >= 0 && ids[b] < r->cfg.vocab_size, NVC_EINVAL, "token id out of range");
    NVC_REQUIRE(ctx, positions[b] >= 0 && positions[b] < r->opt.max_model_len, NVC_EINVAL, "position out of range");
    NVC_REQUIRE(ctx, slot_mapping[b] >= 0 && slot_mapping[b] < nslots, NVC_EINVAL, "slot out of range");
    NVC_REQUIRE(ctx, context_lens[b] > 0 && (int64_t)context_lens[b] <= (int64_t)max_blocks * bs, NVC_EINVAL, "context_len vs block table");
    for (int j = 0; j < (context_lens[b] + bs - 1) / bs; ++j)     // the blocks the context needs exist (range: below, on upload)
      NVC_REQUIRE(ctx, block_tables[(size_t)b * max_blocks + j] >= 0, NVC_EINVAL, "block id out of range");
  }
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  int32_t* h = r->h_stage;
  memcpy(h, ids, B * 4);
  memcpy(h + B, positions, B * 4);
  memcpy(h + 2 * B, slot_mapping, B * 4);
  memcpy(h + 3 * B, context_lens, B * 4);
  cudaStream_t st = r->stream;
  // block table: the device copy is kept between decode steps; a row changes only when a sequence crosses a block
  // boundary or the batch changes, so most steps upload (and validate) nothing but the 4 x B scalars above
  int32_t* hbt = h + 4 * B;
  const int cap = r->max_blocks_cap;
  bool bt_changed = r->bt_mirror_B != B;
  for (int b = 0; b < B; ++b) {
    int32_t* row = hbt + (size_t)b * cap;
    memcpy(row, block_tables + (size_t)b * max_blocks, max_blocks * 4);
    for (int j = max_blocks; j < cap; ++j) row[j] = -1;
    bt_changed = bt_changed || memcmp(row, r->bt_mirror.data() + (size_t)b * cap, (size_t)cap * 4) != 0;
  }
  r->last_h2d_bytes = (int64_t)B * 16;
  if (bt_changed) {
    for (int b = 0; b < B; ++b)
      for (int j = 0; j < max_blocks; ++j) {
        const int32_t id = hbt[(size_t)b * cap + j];
        NVC_REQUIRE(ctx, id >= -1 && id < r->opt.num_kvcache_blocks, NVC_EINVAL, "block id out of range");
      }
    R_CUDA(cudaMemcpyAsync(r->d_bt, hbt, (size_t)B * cap * 4, cudaMemcpyHostToDevice, st));
    memcpy(r->bt_mirror.data(), hbt, (size_t)B * cap * 4);
    r->bt_mirror_B = B;
    r->last_h2d_bytes += (int64_t)B * cap * 4;
  }
  const int hint = std::min(next_pow2(max_blocks), cap);
  if (r->sampling_armed)
    NVC_REQUIRE(ctx, r->last_sampling_B == B, NVC_ESTATE, "nvc_runner_set_sampling was armed for a different batch size");
  const int key2 = 2 * hint + (r->sampling_armed ? 1 : 0);
  // the whole step - one strided H2D of the four input rows, the sampling parameters, the kernels, the D2H of the
  // tokens - is ONE graph launch: the host touches pinned memory, launches, and waits
  auto whole_step = [&]() -> int {
    cudaError_t ce = cudaMemcpy2DAsync(r->d_ids, r->in_pitch * 4, h, (size_t)B * 4, (size_t)B * 4, 4, cudaMemcpyHostToDevice, st);
    if (ce != cudaSuccess) return NVC_CUDA_ERROR_BASE + (int)ce;
    int rc2 = stage_sampling(r, B);
    if (rc2) return rc2;
    rc2 = decode_body(r, B, hint);
    if (rc2) return rc2;
    return enqueue_token_readback(r, B);
  };
  int rc = NVC_OK;
  if (r->opt.use_cuda_graph) {
    const auto key = std::make_pair(B, key2);
    auto it = r->graphs.find(key);
    if (it == r->graphs.end()) {
      // capture this (batch size, context bucket) once; pointers and shapes are fixed inside
      const uint64_t l0 = ctx->launches;
      R_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      rc = whole_step();
      cudaGraph_t graph = nullptr;
      cudaError_t ce = cudaStreamEndCapture(st, &graph);
      if (rc) { r->sampling_armed = false; if (graph) cudaGraphDestroy(graph); return rc; }
      R_CUDA(ce);
      cudaGraphExec_t exec = nullptr;
      ce = cudaGraphInstantiate(&exec, graph, 0);
      cudaGraphDestroy(graph);
      R_CUDA(ce);
      r->launches_per_decode = (int)(ctx->launches - l0);
      it = r->graphs.emplace(key, exec).first;
    }
    R_CUDA(cudaGraphLaunch(it->second, st));
  } else {
    const uint64_t l0 = ctx->launches;
    rc = whole_step();
    if (rc) { r->sampling_armed = false; return rc; }
    r->launches_per_decode = (int)(ctx->launches - l0);
  }
  r->sampling_armed = false;
  r->last_B = B;
  r->last_decode_hint = hint;
  r->last_decode_key2 = key2;
  R_CUDA(cudaStreamSynchronize(st));
  return collect_tokens(r, B, out_tokens);
}

extern "C" int nvc_runner_set_sampling(nvc_runner* r, const float* temperatures, const uint64_t* streams, int B,
                                       uint64_t seed) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  if (!temperatures) {   // disarm: the next step is greedy
    r->sampling_armed = false;
    return NVC_OK;
  }
  NVC_REQUIRE(ctx, streams && B > 0 && B <= r->opt.max_num_seqs, NVC_EINVAL, "bad sampling arguments");
  for (int b = 0; b < B; ++b) NVC_REQUIRE(ctx, temperatures[b] > 0.f, NVC_EINVAL, "temperature must be positive");
  if (seed != r->sampling_seed) {
    // the seed is a kernel argument baked into the captured decode graphs of sampling steps
    NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    R_CUDA(cudaStreamSynchronize(r->stream));
    for (auto it = r->graphs.begin(); it != r->graphs.end();) {
      if (it->first.second & 1) { cudaGraphExecDestroy(it->second); it = r->graphs.erase(it); } else { ++it; }
    }
    r->sampling_seed = seed;
  }
  R_CUDA(cudaStreamSynchronize(r->stream));   // the previous step's H2D out of the pinned buffers has completed
  memcpy(r->h_temps, temperatures, B * sizeof(float));
  memcpy(r->h_streams, streams, B * sizeof(uint64_t));
  r->last_sampling_B = B;
  r->sampling_armed = true;
  return NVC_OK;
}

extern "C" int nvc_runner_last_logits(nvc_runner* r, float* out, int B) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  NVC_REQUIRE(ctx, out && B > 0 && B <= r->last_B, NVC_EINVAL, "no logits for that batch");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  R_CUDA(cudaMemcpyAsync(out, r->logits, (size_t)B * r->V_l * sizeof(float), cudaMemcpyDeviceToHost, r->stream));
  R_CUDA(cudaStreamSynchronize(r->stream));
  return NVC_OK;
}

extern "C" int nvc_runner_time_decode(nvc_runner* r, int iters, float* ms_per_step, int* launches_per_step) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  NVC_REQUIRE(ctx, ms_per_step && iters > 0, NVC_EINVAL, "bad arguments");
  NVC_REQUIRE(ctx, r->last_B > 0 && r->last_decode_hint > 0, NVC_ESTATE, "no decode step has run yet");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  cudaEvent_t e0, e1;
  R_CUDA(cudaEventCreate(&e0));
  R_CUDA(cudaEventCreate(&e1));
  cudaStream_t st = r->stream;
  auto it = r->graphs.find(std::make_pair(r->last_B, r->last_decode_key2));
  R_CUDA(cudaStreamSynchronize(st));
  R_CUDA(cudaEventRecord(e0, st));
  int rc = NVC_OK;
  for (int i = 0; i < iters && rc == NVC_OK; ++i) {
    if (it != r->graphs.end()) {
      cudaError_t ce = cudaGraphLaunch(it->second, st);
      if (ce != cudaSuccess) rc = NVC_CUDA_ERROR_BASE + (int)ce;
    } else {
      rc = decode_body(r, r->last_B, r->last_decode_hint);
    }
  }
  cudaEventRecord(e1, st);
  cudaError_t ce = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (ce == cudaSuccess) ce = cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (rc) return rc;
  R_CUDA(ce);
  *ms_per_step = ms / iters;
  if (launches_per_step) *launches_per_step = r->launches_per_decode;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------------
// Per-kernel timing for bench.py's roofline line: runs ONE kernel family over every layer's
// own weights / KV pool (so the working set is the real one, far larger than L2) with the
// inputs of the last decode step, `iters` times, between CUDA events on the runner's stream.
// ------------------------------------------------------------------------------------------
extern "C" int nvc_runner_time_op(nvc_runner* r, int op, int iters, float* ms_per_launch, int* launches) {
  if (!r) return NVC_EINVAL;
  nvc_ctx* ctx = r->ctx;
  NVC_REQUIRE(ctx, ms_per_launch && iters > 0, NVC_EINVAL, "bad arguments");
  NVC_REQUIRE(ctx, r->last_B > 0 && r->last_decode_hint > 0, NVC_ESTATE, "no decode step has run yet");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  const nvc_qwen3_config& c = r->cfg;
  const int B = r->last_B, H = c.hidden_size, hd = c.head_dim;
  const float scale = 1.0f / sqrtf((float)hd);
  void* st = r->stream;
  cudaEvent_t e0, e1;
  R_CUDA(cudaEventCreate(&e0));
  R_CUDA(cudaEventCreate(&e1));
  R_CUDA(cudaStreamSynchronize(r->stream));
  int rc = NVC_OK, n = 0;
  const uint64_t l0 = ctx->launches;
  // one pass over the layers is captured into a CUDA graph and replayed `iters` times, so the
  // number is GPU time (launch gaps as in the real step), not host launch cost
  R_CUDA(cudaStreamBeginCapture(r->stream, cudaStreamCaptureModeThreadLocal));
  for (int it = 0; it < 1 && rc == NVC_OK; ++it) {
    if (op == NVC_OP_LM_HEAD) {
      rc = nvc_gemm_bf16(ctx, r->xn_last, H, r->lm_head, r->logits, r->V_l, NVC_DT_F32, B, r->V_l, H, st);
      ++n;
      continue;
    }
    if (op == NVC_OP_ARGMAX) {
      rc = nvc_argmax_last(ctx, r->logits, r->V_l, r->d_tok, r->d_tok_val, B, r->V_l, r->rank * r->V_l, st);
      ++n;
      continue;
    }
    for (int li = 0; li < c.num_hidden_layers && rc == NVC_OK; ++li, ++n) {
      Layer& L = r->layers[li];
      switch (op) {
        case NVC_OP_PAGED_DECODE:
          rc = nvc_paged_decode_attn_ex(ctx, r->qkv, r->qkv_l, L.k_cache, L.v_cache, r->d_bt, r->max_blocks_cap,
                                        r->last_decode_hint, r->d_ctx, B, r->nh_l, r->kv_l, hd,
                                        r->opt.kvcache_block_size, scale, r->attn, r->decode_ws, r->decode_ws_bytes, st);
          break;
        case NVC_OP_GEMM_QKV:
          rc = nvc_gemm_bf16(ctx, r->xn, H, L.w_qkv, r->qkv, r->qkv_l, NVC_DT_BF16, B, r->qkv_l, H, st);
          break;
        case NVC_OP_GEMM_O:
          rc = nvc_gemm_bf16(ctx, r->attn, r->nh_l * hd, L.w_o, r->x, H, NVC_DT_BF16, B, H, r->nh_l * hd, st);
          break;
        case NVC_OP_GEMM_GATE_UP:
          rc = nvc_gemm_bf16_silu_mul(ctx, r->xn, H, L.w_gu, r->act, r->I_l, B, r->I_l, H, st);
          break;
        case NVC_OP_GEMM_DOWN:
          rc = nvc_gemm_bf16(ctx, r->act, r->I_l, L.w_down, r->x, H, NVC_DT_BF16, B, H, r->I_l, st);
          break;
        case NVC_OP_RMSNORM:
          rc = nvc_fused_add_rmsnorm(ctx, r->x, r->residual, r->residual, L.ln2, c.rms_norm_eps, r->xn, B, H, st);
          break;
        case NVC_OP_QK_NORM_ROPE:
          // slot_mapping = NULL: do not scatter (the qkv buffer no longer holds this step's k/v)
          rc = nvc_qk_norm_rope_store(ctx, r->qkv, L.q_norm, L.k_norm, c.rms_norm_eps, r->rope, r->d_pos, nullptr,
                                      nullptr, nullptr, B, r->nh_l, r->kv_l, hd, st);
          break;
        case NVC_OP_SILU_MUL:
          rc = nvc_silu_and_mul(ctx, r->qkv, r->act, B, r->qkv_l / 2, st);   // stand-alone K6 (not on the step any more)
          break;
        case NVC_OP_NOOP:   // smallest possible kernel: the per-launch floor inside a graph
          rc = nvc_gather_rows(ctx, r->x, nullptr, r->d_last_rows, r->x_last, nullptr, 1, H, st);
          break;
        case NVC_OP_ALLREDUCE:     // what the decode step uses: fused in-switch / peer all-reduce + norm, or NCCL
          if (r->x_nvls && nvc_ctx_nvls_ready(ctx))
            rc = nvc_nvls_allreduce_rmsnorm(ctx, r->x, r->residual, r->residual, L.ln2, c.rms_norm_eps, r->xn, B, H, 0, st);
          else if (nvc_ctx_p2p_ready(ctx) && (size_t)B * H * 2 <= nvc_ctx_p2p_slot_bytes(ctx) &&
              !(getenv("NVC_P2P_AR") && atoi(getenv("NVC_P2P_AR")) == 0))
            rc = nvc_p2p_allreduce_rmsnorm(ctx, li & 1, r->residual, r->residual, L.ln2, c.rms_norm_eps, r->xn, B, H, st);
          else
            rc = nvc_allreduce_bf16(ctx, r->x, (size_t)B * H, st);
          break;
        default:
          rc = NVC_EINVAL;
      }
    }
  }
  cudaGraph_t graph = nullptr;
  cudaError_t ce = cudaStreamEndCapture(r->stream, &graph);
  cudaGraphExec_t exec = nullptr;
  if (ce == cudaSuccess && rc == NVC_OK) ce = cudaGraphInstantiate(&exec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  float ms = 0.f;
  if (ce == cudaSuccess && rc == NVC_OK) {
    ce = cudaGraphLaunch(exec, r->stream);                      // warm-up replay
    if (ce == cudaSuccess) ce = cudaEventRecord(e0, r->stream);
    for (int it = 0; it < iters && ce == cudaSuccess; ++it) ce = cudaGraphLaunch(exec, r->stream);
    if (ce == cudaSuccess) ce = cudaEventRecord(e1, r->stream);
    if (ce == cudaSuccess) ce = cudaEventSynchronize(e1);
    if (ce == cudaSuccess) ce = cudaEventElapsedTime(&ms, e0, e1);
  }
  if (exec) cudaGraphExecDestroy(exec);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (rc) return rc;
  R_CUDA(ce);
  const int kernels = (int)(ctx->launches - l0) * iters;
  n *= iters;
  *ms_per_launch = n ? ms / n : 0.f;
  if (launches) *launches = kernels;
  return NVC_OK;
}
