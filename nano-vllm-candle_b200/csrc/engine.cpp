// Host-side engine book-keeping behind the C ABI of include/nvc_b200_engine.h:
// Sequence / BlockManager / Scheduler / LLMEngine of the reference, in C++ because the
// reference's own host language (Rust) has no toolchain here.
//
//   Sequence      src/engine/sequence.rs:16-97
//   BlockManager  src/engine/block_manager.rs:31-99  (the reference's is a stub with identity
//                 block tables and always-true capacity; this one keeps the same methods and
//                 makes them real: FIFO free list, disjoint block ids, truthful can_allocate /
//                 can_append, may_append grows the table when len % block_size == 1)
//   Scheduler     src/engine/scheduler.rs:58-250     (same control flow, statement by statement)
//   LLMEngine     src/engine/llm_engine.rs:209-326
// plus the integer contract the north_star adds (N1 slot_mapping, N2 block_tables), written
// only in terms of the Sequence helpers the reference exposes (sequence.rs:73-91).
#include <algorithm>
#include <cstdint>
#include <deque>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/nvc_b200_engine.h"

namespace {

enum class Status { Waiting, Running, Finished };

struct Sequence {
  int64_t seq_id;
  int block_size;
  std::vector<int32_t> block_table;
  Status status = Status::Waiting;
  std::vector<int32_t> token_ids;
  int32_t last_token = 0;
  size_t num_tokens = 0;
  size_t num_prompt_tokens = 0;
  size_t num_cached_tokens = 0;
  size_t num_computed_tokens = 0;   // chunked prefill: tokens whose K / V are in the cache
  double temperature = 1.0;
  size_t max_tokens = 64;
  bool ignore_eos = false;

  size_t len() const { return num_tokens; }
  bool is_finished() const { return status == Status::Finished; }
  size_t num_completion_tokens() const { return num_tokens - num_prompt_tokens; }
  size_t num_cached_blocks() const { return num_cached_tokens / block_size; }
  size_t num_blocks() const { return (num_tokens + block_size - 1) / block_size; }
  size_t last_block_num_tokens() const { return num_tokens - (num_blocks() - 1) * block_size; }
  void append_token(int32_t t) {
    token_ids.push_back(t);
    last_token = t;
    num_tokens += 1;
  }
};
using SeqPtr = std::shared_ptr<Sequence>;

// One KV block's book-keeping for prefix caching (upstream nano-vllm's Block: ref_count, hash,
// token_ids).  Without prefix caching only ref_count is used.
struct Block {
  int ref_count = 0;
  bool has_hash = false;
  bool ready = false;   // chunked prefill: the block's K / V have been computed (a hit needs it)
  uint64_t hash = 0;
  std::vector<int32_t> token_ids;
};

// 64-bit chained block hash: FNV-1a over (prefix hash, token ids) + a splitmix finaliser.  The
// reference reserves `ahash` for this (Cargo.toml:19-22), whose values are not portable; only
// equality of hashes is observable, and a hit is confirmed by comparing the token ids.
uint64_t chain_hash(const int32_t* toks, size_t n, bool has_prefix, uint64_t prefix) {
  uint64_t h = 0xcbf29ce484222325ull;
  auto eat = [&h](uint64_t v, int bytes) {
    for (int i = 0; i < bytes; ++i) {
      h ^= (v >> (8 * i)) & 0xffu;
      h *= 0x100000001b3ull;
    }
  };
  if (has_prefix) eat(prefix, 8);
  for (size_t i = 0; i < n; ++i) eat((uint32_t)toks[i], 4);
  h ^= h >> 30; h *= 0xbf58476d1ce4e5b9ull;
  h ^= h >> 27; h *= 0x94d049bb133111ebull;
  h ^= h >> 31;
  return h;
}

// FIFO of free block ids with O(1) removal from the middle (a cached block that is free is
// revived in place): doubly linked through prev/next arrays.
struct FreeList {
  std::vector<int32_t> prev, next;
  std::vector<uint8_t> in;
  int32_t head = -1, tail = -1;
  size_t count = 0;
  explicit FreeList(int n) : prev(n, -1), next(n, -1), in(n, 0) {}
  size_t size() const { return count; }
  int32_t front() const { return head; }
  void push_back(int32_t b) {
    prev[b] = tail; next[b] = -1;
    if (tail >= 0) next[tail] = b; else head = b;
    tail = b; in[b] = 1; ++count;
  }
  void remove(int32_t b) {
    if (prev[b] >= 0) next[prev[b]] = next[b]; else head = next[b];
    if (next[b] >= 0) prev[next[b]] = prev[b]; else tail = prev[b];
    prev[b] = next[b] = -1; in[b] = 0; --count;
  }
};

// BlockManager (block_manager.rs:31-99).  The reference's is an admitted stub (:24-29); this one
// keeps its method set and follows upstream nano-vllm's allocator, which the reference says its
// API mirrors: FIFO free list, ref-counted blocks and - when `prefix_caching` is on -
// hash-chained dedup of FULL blocks.  One rule is added to upstream's: a block is taken from the
// cache only if at least one token of the sequence is left to compute ((i+1)*block_size < len),
// otherwise a fully cached prompt would schedule a prefill step with no query tokens.
struct BlockManager {
  int num_blocks_;
  int block_size_;
  bool prefix_caching;
  bool chunked = false;   // chunked prefill: blocks are registered at allocation but computed over several steps
  std::vector<Block> blocks;
  FreeList free_;
  std::unordered_map<uint64_t, int32_t> hash_to_block;
  int64_t cache_hits = 0, cache_queries = 0;   // in blocks

  BlockManager(int num_blocks, int block_size, bool prefix)
      : num_blocks_(num_blocks), block_size_(block_size), prefix_caching(prefix), blocks(num_blocks), free_(num_blocks) {
    for (int i = 0; i < num_blocks; ++i) free_.push_back(i);
  }
  int block_size() const { return block_size_; }
  int num_blocks() const { return num_blocks_; }

  Block& take(int32_t id) {            // upstream _allocate_block
    Block& b = blocks[id];
    b.ref_count = 1;
    b.has_hash = false;
    b.ready = false;
    b.token_ids.clear();
    free_.remove(id);
    return b;
  }
  void set_hash(int32_t id, uint64_t h, const int32_t* toks, size_t n) {
    Block& b = blocks[id];
    b.has_hash = true;
    b.hash = h;
    b.token_ids.assign(toks, toks + n);
    hash_to_block[h] = id;
  }

  bool can_allocate(const Sequence& s) const { return free_.size() >= s.num_blocks(); }
  void allocate(Sequence& s) {
    s.block_table.clear();
    s.num_cached_tokens = 0;
    const size_t bs = (size_t)block_size_, nb = s.num_blocks();
    bool has_h = false, miss = !prefix_caching;
    uint64_t h = 0;
    for (size_t i = 0; i < nb; ++i) {
      const int32_t* toks = s.token_ids.data() + i * bs;
      const size_t n = std::min(bs, s.num_tokens - i * bs);
      int32_t id = -1;
      if (prefix_caching) {
        if (n == bs) { h = chain_hash(toks, n, has_h, h); has_h = true; } else { has_h = false; }
        if (has_h) {
          ++cache_queries;
          auto it = hash_to_block.find(h);
          if (it != hash_to_block.end()) id = it->second;
        }
        if (id < 0 || !blocks[id].has_hash || blocks[id].hash != h || blocks[id].token_ids.size() != n ||
            !std::equal(toks, toks + n, blocks[id].token_ids.begin()) || (i + 1) * bs >= s.num_tokens ||
            (chunked && !blocks[id].ready))
          miss = true;
      }
      if (miss) {
        id = free_.front();
        take(id);
      } else {
        ++cache_hits;
        s.num_cached_tokens += bs;
        if (free_.in[id]) {
          // a free block revived from the cache still holds its K / V: it stays ready (take() clears the flag,
          // and mark_computed never revisits a cached block, so the next request with this prefix would miss)
          const bool was_ready = blocks[id].ready;
          take(id);
          blocks[id].ready = was_ready;
        } else {
          blocks[id].ref_count += 1;
        }
      }
      if (prefix_caching && has_h) set_hash(id, h, toks, n);
      s.block_table.push_back(id);
    }
  }
  void deallocate(Sequence& s) {
    for (auto it = s.block_table.rbegin(); it != s.block_table.rend(); ++it) {
      Block& b = blocks[*it];
      if (--b.ref_count == 0) free_.push_back(*it);   // keeps its hash: still a cache candidate until re-taken
    }
    s.block_table.clear();
    s.num_cached_tokens = 0;
  }
  // A step that failed in the runner never wrote the K / V of its sequences: their blocks must not
  // serve prefix-cache hits.  (A block shared with a sequence outside the batch only loses its cache entry.)
  void forget_hashes(const Sequence& s) {
    for (int32_t id : s.block_table) {
      Block& b = blocks[id];
      if (b.has_hash) {
        auto it = hash_to_block.find(b.hash);
        if (it != hash_to_block.end() && it->second == id) hash_to_block.erase(it);
      }
      b.has_hash = false;
      b.ready = false;
      b.token_ids.clear();
    }
  }
  bool can_append(const Sequence& s) const {
    // a block is missing only when the newest token opens one (len % block_size == 1) and the table
    // does not hold it yet (it does when that token is still part of the prompt: chunked prefill)
    return free_.size() >= (size_t)((s.len() % block_size_ == 1 && s.block_table.size() < s.num_blocks()) ? 1 : 0);
  }
  void may_append(Sequence& s) {
    const size_t bs = (size_t)block_size_;
    if (s.len() % bs == 1 && s.block_table.size() < s.num_blocks()) {
      const int32_t id = free_.front();
      take(id);
      s.block_table.push_back(id);
    } else if (prefix_caching && s.len() % bs == 0 && !s.block_table.empty()) {
      // the last block becomes full with this step's token: register it (upstream may_append)
      const int32_t id = s.block_table.back();
      if (!blocks[id].has_hash) {
        const size_t nb = s.block_table.size();
        const bool has_prefix = nb > 1 && blocks[s.block_table[nb - 2]].has_hash;
        if (nb == 1 || has_prefix) {
          const int32_t* toks = s.token_ids.data() + (nb - 1) * bs;
          set_hash(id, chain_hash(toks, bs, has_prefix, has_prefix ? blocks[s.block_table[nb - 2]].hash : 0), toks, bs);
        }
      }
    }
  }
};

struct Scheduler {
  size_t max_num_seqs;
  size_t max_num_batched_tokens;
  BlockManager block_manager;
  std::deque<SeqPtr> waiting, running;
  int32_t eos;
  size_t max_model_len = 0;   // longest sequence the runner can hold (0 = unknown: no limit enforced here)

  explicit Scheduler(const nvc_scheduler_config& c)
      : max_num_seqs(c.max_num_seqs), max_num_batched_tokens(c.max_num_batched_tokens),
        block_manager(c.num_kvcache_blocks, c.kvcache_block_size, c.enable_prefix_caching != 0), eos(c.eos) {
    block_manager.chunked = c.enable_chunked_prefill != 0;
  }

  bool is_finished() const { return waiting.empty() && running.empty(); }
  void add(const SeqPtr& s) { waiting.push_back(s); }

  // scheduler.rs:113-157
  bool try_schedule_prefill(std::vector<SeqPtr>& out) {
    size_t num_seqs = 0, num_batched = 0;
    while (!waiting.empty() && num_seqs < max_num_seqs) {
      SeqPtr seq = waiting.front();
      const bool ok = num_batched + seq->len() <= max_num_batched_tokens && block_manager.can_allocate(*seq);
      if (!ok) break;
      num_seqs += 1;
      block_manager.allocate(*seq);
      num_batched += seq->len() - seq->num_cached_tokens;
      seq->status = Status::Running;
      waiting.pop_front();
      running.push_back(seq);
      out.push_back(seq);
    }
    return !out.empty();
  }

  // scheduler.rs:188-198
  bool ensure_can_append(const SeqPtr& seq) {
    while (!block_manager.can_append(*seq)) {
      if (!running.empty()) {
        SeqPtr victim = running.back();
        running.pop_back();
        preempt(victim);
      } else {
        preempt(seq);
        return false;
      }
    }
    return true;
  }

  // scheduler.rs:159-186.  Returns false where the reference would hit its assert.
  bool schedule_decode(std::vector<SeqPtr>& out) {
    size_t num_seqs = 0;
    while (!running.empty() && num_seqs < max_num_seqs) {
      SeqPtr seq = running.front();
      running.pop_front();
      if (ensure_can_append(seq)) {
        num_seqs += 1;
        block_manager.may_append(*seq);
        out.push_back(seq);
      }
    }
    if (out.empty()) return false;
    for (auto it = out.rbegin(); it != out.rend(); ++it) running.push_front(*it);
    return true;
  }

  // scheduler.rs:200-211
  void preempt(const SeqPtr& seq) {
    seq->status = Status::Waiting;
    block_manager.deallocate(*seq);
    seq->num_computed_tokens = 0;
    waiting.push_front(seq);
  }

  // Chunked prefill (no counterpart in the reference, whose schedule() is either a
  // whole-prompt prefill step or a decode step, scheduler.rs:106-111).  One step =
  //   1. every running sequence, in queue order as schedule_decode visits them: a sequence with one
  //      pending token (decode, or the last token of its prompt) appends exactly as in
  //      scheduler.rs:159-198 (victims are preempted from the back of the queue); a sequence in the
  //      middle of its prompt reserves one token of the budget;
  //   2. the rest of the token budget goes to those prompt chunks in order, then
  //   3. to waiting sequences while running.size() < max_num_seqs (the bound holds over ALL running
  //      sequences, not per round) - unless this step preempted, which means the pool is full.
  // The batch lists the one-token sequences first (they take the paged decode kernel).
  // q_lens[i] = tokens of out[i] computed in this step; it emits a token iff q_lens[i] == pending.
  bool schedule_chunked(std::vector<SeqPtr>& out, std::vector<int32_t>& q_lens, int* num_decode) {
    std::vector<SeqPtr> ones, chunks, visited;
    size_t budget = max_num_batched_tokens;
    bool preempted = false;
    while (!running.empty() && visited.size() < max_num_seqs && budget > 0) {
      SeqPtr seq = running.front();
      running.pop_front();
      const size_t pending = seq->len() - seq->num_computed_tokens;
      if (pending == 1) {
        const size_t before = waiting.size();
        const bool ok = ensure_can_append(seq);
        preempted |= waiting.size() != before;
        if (!ok) continue;
        block_manager.may_append(*seq);
        ones.push_back(seq);
      } else {
        chunks.push_back(seq);
      }
      budget -= 1;
      visited.push_back(seq);
    }
    for (auto it = visited.rbegin(); it != visited.rend(); ++it) running.push_front(*it);
    std::vector<int32_t> chunk_q;
    for (auto& seq : chunks) {
      const size_t pending = seq->len() - seq->num_computed_tokens;
      const size_t extra = std::min(pending - 1, budget);
      chunk_q.push_back((int32_t)(1 + extra));
      budget -= extra;
    }
    while (!preempted && !waiting.empty() && running.size() < max_num_seqs && budget > 0) {
      SeqPtr seq = waiting.front();
      if (!block_manager.can_allocate(*seq)) break;
      block_manager.allocate(*seq);
      seq->num_computed_tokens = seq->num_cached_tokens;
      const size_t q = std::min(seq->len() - seq->num_computed_tokens, budget);
      budget -= q;
      seq->status = Status::Running;
      waiting.pop_front();
      running.push_back(seq);
      chunks.push_back(seq);
      chunk_q.push_back((int32_t)q);
    }
    out = ones;
    q_lens.assign(ones.size(), 1);
    out.insert(out.end(), chunks.begin(), chunks.end());
    q_lens.insert(q_lens.end(), chunk_q.begin(), chunk_q.end());
    *num_decode = (int)ones.size();
    return !out.empty();
  }

  // After a chunked step: the K / V of q more tokens of `seq` are in the cache; blocks that became
  // complete may now serve prefix-cache hits.
  void mark_computed(const SeqPtr& seq, size_t q) {
    const size_t bs = (size_t)block_manager.block_size();
    const size_t b0 = seq->num_computed_tokens / bs;
    seq->num_computed_tokens += q;
    for (size_t i = b0; i < seq->num_computed_tokens / bs && i < seq->block_table.size(); ++i)
      block_manager.blocks[seq->block_table[i]].ready = true;
  }

  // scheduler.rs:219-249; returns true when the sequence finished
  bool process_single(const SeqPtr& seq, int32_t tok) {
    seq->append_token(tok);
    const bool is_eos = tok == eos;
    // the appended token sits at position len - 1; the runner holds positions < max_model_len, so a sequence that
    // has grown past it can not take another decode step: it ends here like one that reached max_tokens
    const bool reached_max = seq->num_completion_tokens() >= seq->max_tokens ||
                             (max_model_len > 0 && seq->len() > max_model_len);
    const bool finished = (!seq->ignore_eos && is_eos) || reached_max;
    if (finished) {
      seq->status = Status::Finished;
      block_manager.deallocate(*seq);
      const int64_t id = seq->seq_id;
      running.erase(std::remove_if(running.begin(), running.end(),
                                   [id](const SeqPtr& s) { return s->seq_id == id; }),
                    running.end());
    }
    return finished;
  }
};

}  // namespace

struct nvc_engine {
  Scheduler scheduler;
  nvc_runner* runner;
  int block_size;
  int64_t next_seq_id = 0;                     // sequence.rs:5 SEQ_COUNTER (per engine here)
  std::map<int64_t, SeqPtr> all;               // every sequence ever added (outputs / introspection)
  int num_finished = 0;
  std::string err;
  // last scheduled batch and its step inputs
  std::vector<SeqPtr> batch;
  bool batch_is_prefill = false;
  bool batch_valid = false;
  bool chunked = false;                        // chunked prefill / mixed steps
  int num_decode = 0;                          // leading one-token sequences of the batch
  std::vector<int32_t> q_lens, emit;           // per sequence of the batch
  std::vector<int32_t> ids, positions, cu_seqlens, slot_mapping, block_tables, context_lens, out_tokens;
  std::vector<int64_t> seq_ids;
  int max_blocks = 0, max_seqlen = 0;
  bool sampling = false;                       // on-device temperature sampling
  uint64_t sampling_seed = 0;
  std::vector<float> temps;
  std::vector<uint64_t> streams;

  nvc_engine(const nvc_scheduler_config& c, nvc_runner* r)
      : scheduler(c), runner(r), block_size(c.kvcache_block_size), chunked(c.enable_chunked_prefill != 0) {}
};

namespace {

// N1 (prefill): block i in num_cached_blocks..num_blocks -> slots block_table[i]*BS + [0, BS)
// (last block: [0, last_block_num_tokens)).
void prefill_slots(const Sequence& s, std::vector<int32_t>& out) {
  const int bs = s.block_size;
  for (size_t i = s.num_cached_blocks(); i < s.num_blocks(); ++i) {
    const int32_t start = s.block_table[i] * bs;
    const size_t n = (i != s.num_blocks() - 1) ? (size_t)bs : s.last_block_num_tokens();
    for (size_t j = 0; j < n; ++j) out.push_back(start + (int32_t)j);
  }
}

void build_inputs(nvc_engine* e) {
  e->ids.clear(); e->positions.clear(); e->cu_seqlens.clear(); e->slot_mapping.clear();
  e->block_tables.clear(); e->context_lens.clear(); e->seq_ids.clear();
  e->max_blocks = 0; e->max_seqlen = 0;
  for (auto& s : e->batch) e->seq_ids.push_back(s->seq_id);
  if (e->chunked) {
    // tokens [num_computed, num_computed + q) of every sequence; slot of token t = N1's rule per token:
    // block_table[t / BS] * BS + t % BS
    e->cu_seqlens.push_back(0);
    e->emit.clear();
    for (size_t b = 0; b < e->batch.size(); ++b) {
      const Sequence& s = *e->batch[b];
      const size_t start = s.num_computed_tokens, q = (size_t)e->q_lens[b];
      for (size_t t = start; t < start + q; ++t) {
        e->ids.push_back(s.token_ids[t]);
        e->positions.push_back((int32_t)t);
        e->slot_mapping.push_back(s.block_table[t / s.block_size] * s.block_size + (int32_t)(t % s.block_size));
      }
      e->cu_seqlens.push_back(e->cu_seqlens.back() + (int32_t)q);
      e->max_seqlen = std::max(e->max_seqlen, (int)q);
      e->context_lens.push_back((int32_t)(start + q));
      e->emit.push_back(start + q == s.len() ? 1 : 0);
    }
  } else if (e->batch_is_prefill) {
    e->cu_seqlens.push_back(0);
    for (auto& sp : e->batch) {
      const Sequence& s = *sp;
      const size_t start = s.num_cached_tokens;
      for (size_t t = start; t < s.len(); ++t) {
        e->ids.push_back(s.token_ids[t]);
        e->positions.push_back((int32_t)t);
      }
      e->cu_seqlens.push_back(e->cu_seqlens.back() + (int32_t)(s.len() - start));
      e->max_seqlen = std::max(e->max_seqlen, (int)(s.len() - start));
      prefill_slots(s, e->slot_mapping);
      e->context_lens.push_back((int32_t)s.len());
    }
  } else {
    for (auto& sp : e->batch) {
      const Sequence& s = *sp;
      e->ids.push_back(s.last_token);
      e->positions.push_back((int32_t)s.len() - 1);
      // N1 (decode): block_table[last]*BS + last_block_num_tokens - 1
      e->slot_mapping.push_back(s.block_table.back() * s.block_size + (int32_t)s.last_block_num_tokens() - 1);
      e->context_lens.push_back((int32_t)s.len());
    }
    e->max_seqlen = 1;
  }
  if (!e->chunked) {
    e->emit.assign(e->batch.size(), 1);
    e->num_decode = e->batch_is_prefill ? 0 : (int)e->batch.size();
  }
  // N2: [B, max_blocks] padded with -1
  for (auto& sp : e->batch) e->max_blocks = std::max(e->max_blocks, (int)sp->block_table.size());
  e->block_tables.assign((size_t)e->batch.size() * e->max_blocks, -1);
  for (size_t b = 0; b < e->batch.size(); ++b)
    std::copy(e->batch[b]->block_table.begin(), e->batch[b]->block_table.end(),
              e->block_tables.begin() + b * e->max_blocks);
}

int do_schedule(nvc_engine* e) {
  e->batch.clear();
  e->batch_valid = false;
  if (e->scheduler.is_finished()) {
    e->err = "schedule: no waiting or running sequences";
    return NVC_ESTATE;
  }
  if (e->chunked) {
    if (!e->scheduler.schedule_chunked(e->batch, e->q_lens, &e->num_decode)) {
      e->err = "no sequence could be scheduled (KV pool exhausted)";
      return NVC_ENOMEM;
    }
    e->batch_is_prefill = e->num_decode < (int)e->batch.size();
  } else if (e->scheduler.try_schedule_prefill(e->batch)) {
    e->batch_is_prefill = true;
  } else {
    e->batch.clear();
    if (!e->scheduler.schedule_decode(e->batch)) {
      e->err = "decode stage should schedule at least one sequence (KV pool exhausted)";   // scheduler.rs:175-178
      return NVC_ENOMEM;
    }
    e->batch_is_prefill = false;
  }
  build_inputs(e);
  e->batch_valid = true;
  return NVC_OK;
}

// per-sequence temperature + noise stream (seq_id, completion tokens so far) for the step about to run
int arm_sampling(nvc_engine* e) {
  e->temps.clear();
  e->streams.clear();
  for (auto& s : e->batch) {
    e->temps.push_back((float)s->temperature);
    e->streams.push_back(((uint64_t)(uint32_t)s->seq_id << 32) | (uint32_t)s->num_completion_tokens());
  }
  return nvc_runner_set_sampling(e->runner, e->temps.data(), e->streams.data(), (int)e->batch.size(), e->sampling_seed);
}

void rollback_failed_step(nvc_engine* e) {
  auto& run = e->scheduler.running;
  for (auto it = e->batch.rbegin(); it != e->batch.rend(); ++it) {   // reverse: preempt() pushes to the FRONT of waiting
    const SeqPtr& seq = *it;
    auto pos = std::find(run.begin(), run.end(), seq);
    if (pos == run.end()) continue;
    run.erase(pos);
    e->scheduler.block_manager.forget_hashes(*seq);
    e->scheduler.preempt(seq);
  }
  e->batch.clear();
  e->batch_valid = false;
}

int do_post_process(nvc_engine* e, const int32_t* toks, int n, int* finished_out) {
  if (!e->batch_valid || n != (int)e->batch.size()) {
    e->err = "post_process: no scheduled batch of that size";
    return NVC_ESTATE;
  }
  int fin = 0;
  for (int i = 0; i < n; ++i) {
    if (e->chunked) e->scheduler.mark_computed(e->batch[i], (size_t)e->q_lens[i]);
    if (e->emit[i]) fin += e->scheduler.process_single(e->batch[i], toks[i]) ? 1 : 0;
  }
  e->num_finished += fin;
  e->batch_valid = false;
  if (finished_out) *finished_out = fin;
  return NVC_OK;
}

}  // namespace

extern "C" {

int nvc_engine_create(const nvc_scheduler_config* cfg, nvc_runner* runner, nvc_engine** out) {
  if (!cfg || !out) return NVC_EINVAL;
  // block size 1 is refused: the append rule `len % block_size == 1` (block_manager.rs / scheduler.rs:159-198) never
  // fires there, so decode would write past the block table (the reference fixes 256, sequence.rs:35; the runner
  // wants a multiple of 16)
  if (cfg->max_num_seqs <= 0 || cfg->max_num_batched_tokens <= 0 || cfg->kvcache_block_size < 2 ||
      cfg->num_kvcache_blocks < 0)
    return NVC_EINVAL;
  if (cfg->enable_chunked_prefill && cfg->max_num_batched_tokens < cfg->max_num_seqs) return NVC_EINVAL;
  if (cfg->max_model_len < 0) return NVC_EINVAL;
  nvc_scheduler_config c = *cfg;
  if (runner != nullptr) {
    // The engine's integer contract (slot = block * block_size + offset, block ids < num_kvcache_blocks) is only
    // meaningful against the pool the runner owns: a different block size would pass the runner's range checks and
    // read the wrong KV rows.  Sizes the runner cannot execute are clamped instead of failing at the first big step.
    nvc_runner_options ro;
    if (nvc_runner_get_options(runner, &ro) != NVC_OK) return NVC_EINVAL;
    if (c.kvcache_block_size != ro.kvcache_block_size || c.num_kvcache_blocks > ro.num_kvcache_blocks) return NVC_EINVAL;
    if (c.num_kvcache_blocks == 0) c.num_kvcache_blocks = ro.num_kvcache_blocks;
    c.max_num_seqs = std::min(c.max_num_seqs, ro.max_num_seqs);
    c.max_num_batched_tokens = std::min(c.max_num_batched_tokens, ro.max_num_batched_tokens);
    c.max_model_len = c.max_model_len == 0 ? ro.max_model_len : std::min(c.max_model_len, ro.max_model_len);
    if (c.enable_chunked_prefill && c.max_num_batched_tokens < c.max_num_seqs) return NVC_EINVAL;
  }
  *out = new nvc_engine(c, runner);
  (*out)->scheduler.max_model_len = (size_t)c.max_model_len;
  return NVC_OK;
}

int nvc_engine_destroy(nvc_engine* e) {
  if (!e) return NVC_EINVAL;
  delete e;
  return NVC_OK;
}

const char* nvc_engine_last_error(const nvc_engine* e) { return e ? e->err.c_str() : ""; }

int64_t nvc_engine_add_request(nvc_engine* e, const int32_t* token_ids, int n, double temperature, int max_tokens,
                               int ignore_eos) {
  if (!e) return NVC_EINVAL;
  // Sequence::new (sequence.rs:31-47) takes an empty prompt and usize max_tokens; a zero-length sequence has no
  // query token for the model to run on, so the C ABI refuses it (and a negative max_tokens) up front.
  if (n <= 0 || !token_ids || max_tokens < 0) {
    e->err = n <= 0 || !token_ids ? "empty prompt" : "max_tokens must be >= 0";
    return NVC_EINVAL;
  }
  if (e->scheduler.max_model_len > 0 && (size_t)n > e->scheduler.max_model_len) {
    e->err = "prompt of " + std::to_string(n) + " tokens exceeds max_model_len " + std::to_string(e->scheduler.max_model_len);
    return NVC_EINVAL;
  }
  if (!(temperature > 1e-10)) {   // sampling_params.rs:20,30
    e->err = "greedy sampling is not permitted (temperature must be > 1e-10)";
    return NVC_EINVAL;
  }
  auto s = std::make_shared<Sequence>();
  s->seq_id = e->next_seq_id++;
  s->block_size = e->block_size;
  s->token_ids.assign(token_ids, token_ids + n);
  s->last_token = token_ids[n - 1];
  s->num_tokens = s->num_prompt_tokens = (size_t)n;
  s->temperature = temperature;
  s->max_tokens = (size_t)max_tokens;
  s->ignore_eos = ignore_eos != 0;
  e->all[s->seq_id] = s;
  e->scheduler.add(s);
  return s->seq_id;
}

int nvc_engine_schedule(nvc_engine* e, nvc_schedule_info* info) {
  if (!e || !info) return NVC_EINVAL;
  int rc = do_schedule(e);
  if (rc) return rc;
  info->is_prefill = e->batch_is_prefill;
  info->num_seqs = (int)e->batch.size();
  info->num_tokens = (int)e->ids.size();
  info->max_blocks = e->max_blocks;
  info->max_seqlen = e->max_seqlen;
  info->num_decode_seqs = e->num_decode;
  return NVC_OK;
}

int nvc_engine_step_inputs(const nvc_engine* e, const int32_t** ids, const int32_t** positions,
                           const int32_t** cu_seqlens, const int32_t** slot_mapping, const int32_t** block_tables,
                           const int32_t** context_lens, const int64_t** seq_ids) {
  if (!e) return NVC_EINVAL;
  if (!e->batch_valid) return NVC_ESTATE;
  if (ids) *ids = e->ids.data();
  if (positions) *positions = e->positions.data();
  if (cu_seqlens) *cu_seqlens = e->cu_seqlens.data();
  if (slot_mapping) *slot_mapping = e->slot_mapping.data();
  if (block_tables) *block_tables = e->block_tables.data();
  if (context_lens) *context_lens = e->context_lens.data();
  if (seq_ids) *seq_ids = e->seq_ids.data();
  return NVC_OK;
}

int nvc_engine_step_emit_mask(const nvc_engine* e, const int32_t** emit) {
  if (!e || !emit) return NVC_EINVAL;
  if (!e->batch_valid) return NVC_ESTATE;
  *emit = e->emit.data();
  return NVC_OK;
}

int nvc_engine_post_process(nvc_engine* e, const int32_t* token_ids, int n) {
  if (!e || !token_ids) return NVC_EINVAL;
  return do_post_process(e, token_ids, n, nullptr);
}

int nvc_engine_step(nvc_engine* e, nvc_step_output* out) {
  if (!e || !out) return NVC_EINVAL;
  int rc = do_schedule(e);
  if (rc) return rc;
  const int B = (int)e->batch.size();
  e->out_tokens.assign(B, 0);
  if (e->runner == nullptr) {
    for (int i = 0; i < B; ++i) e->out_tokens[i] = e->batch[i]->last_token + 1;   // DummyModelRunner
  } else if (e->sampling && (rc = arm_sampling(e)) != NVC_OK) {
    e->err = "nvc_runner_set_sampling failed";
    return rc;
  } else if (e->chunked && e->num_decode == B) {
    // nothing but one-token sequences: the decode step proper (captured CUDA graph, fused peer all-reduce)
    rc = nvc_runner_decode(e->runner, e->ids.data(), e->positions.data(), e->slot_mapping.data(),
                           e->block_tables.data(), e->max_blocks, e->context_lens.data(), B, e->out_tokens.data());
  } else if (e->chunked) {
    rc = nvc_runner_step_mixed(e->runner, e->ids.data(), e->positions.data(), e->cu_seqlens.data(), e->slot_mapping.data(),
                               e->block_tables.data(), e->max_blocks, e->context_lens.data(), e->num_decode, B,
                               (int)e->ids.size(), e->max_seqlen, e->out_tokens.data());
  } else if (e->batch_is_prefill) {
    bool cached_prefix = false;   // some sequence skips a cached prefix: attention must read it from the paged cache
    for (auto& s : e->batch) cached_prefix |= s->num_cached_tokens > 0;
    if (cached_prefix)
      rc = nvc_runner_prefill_paged(e->runner, e->ids.data(), e->positions.data(), e->cu_seqlens.data(),
                                    e->slot_mapping.data(), e->block_tables.data(), e->max_blocks, e->context_lens.data(),
                                    B, (int)e->ids.size(), e->max_seqlen, e->out_tokens.data());
    else
      rc = nvc_runner_prefill(e->runner, e->ids.data(), e->positions.data(), e->cu_seqlens.data(),
                              e->slot_mapping.data(), B, (int)e->ids.size(), e->max_seqlen, e->out_tokens.data());
  } else {
    rc = nvc_runner_decode(e->runner, e->ids.data(), e->positions.data(), e->slot_mapping.data(),
                           e->block_tables.data(), e->max_blocks, e->context_lens.data(), B, e->out_tokens.data());
  }
  if (rc) {
    // errors propagate; the reference would emit eos for every sequence (llm_engine.rs:153-175).  The scheduler had
    // already moved the batch to `running`, allocated and hash-registered its blocks: undo that, or a caller who
    // continues would decode (and take prefix-cache hits) over K / V that was never written.
    rollback_failed_step(e);
    e->err = "runner step failed (the batch was put back on the waiting queue)";
    return rc;
  }
  const bool is_prefill = e->batch_is_prefill;
  std::vector<SeqPtr> seqs = e->batch;
  out->num_query_tokens = (int)e->ids.size();
  out->num_decode_seqs = e->num_decode;
  out->num_sampled = 0;
  for (int32_t m : e->emit) out->num_sampled += m ? 1 : 0;
  int fin = 0;
  rc = do_post_process(e, e->out_tokens.data(), B, &fin);
  if (rc) return rc;
  out->is_prefill = is_prefill;
  out->num_seqs = B;
  out->num_finished = fin;
  if (e->chunked) {
    out->num_tokens = is_prefill ? out->num_query_tokens : -B;   // a mixed step reports the tokens it computed
  } else if (is_prefill) {
    int tot = 0;
    for (auto& s : seqs) tot += (int)s->len();   // llm_engine.rs:253-254 (counted after the append)
    out->num_tokens = tot;
  } else {
    out->num_tokens = -B;
  }
  return NVC_OK;
}

int nvc_engine_is_finished(const nvc_engine* e) { return e ? (e->scheduler.is_finished() ? 1 : 0) : 1; }

int nvc_engine_set_sampling(nvc_engine* e, int enabled, uint64_t seed) {
  if (!e) return NVC_EINVAL;
  e->sampling = enabled != 0;
  e->sampling_seed = seed;
  return NVC_OK;
}

int nvc_engine_preempt(nvc_engine* e, int64_t seq_id) {
  if (!e) return NVC_EINVAL;
  auto& run = e->scheduler.running;
  for (auto it = run.begin(); it != run.end(); ++it) {
    if ((*it)->seq_id == seq_id) {
      SeqPtr s = *it;
      run.erase(it);
      e->scheduler.preempt(s);
      e->batch_valid = false;
      return NVC_OK;
    }
  }
  e->err = "preempt: sequence is not running";
  return NVC_ESTATE;
}

int nvc_engine_num_finished(const nvc_engine* e) { return e ? e->num_finished : 0; }

int nvc_engine_get_output(const nvc_engine* e, int64_t seq_id, int32_t* tokens, int capacity, int* n_tokens) {
  if (!e || !n_tokens) return NVC_EINVAL;
  auto it = e->all.find(seq_id);
  if (it == e->all.end()) return NVC_EINVAL;
  const Sequence& s = *it->second;
  const int n = (int)s.num_completion_tokens();
  *n_tokens = n;
  if (tokens) {
    if (capacity < n) return NVC_ENOMEM;
    std::copy(s.token_ids.begin() + s.num_prompt_tokens, s.token_ids.end(), tokens);
  }
  return s.is_finished() ? NVC_OK : NVC_ESTATE;
}

int nvc_engine_seq_block_table(const nvc_engine* e, int64_t seq_id, int32_t* blocks, int capacity, int* n_blocks) {
  if (!e || !n_blocks) return NVC_EINVAL;
  auto it = e->all.find(seq_id);
  if (it == e->all.end()) return NVC_EINVAL;
  const auto& bt = it->second->block_table;
  *n_blocks = (int)bt.size();
  if (blocks) {
    if (capacity < (int)bt.size()) return NVC_ENOMEM;
    std::copy(bt.begin(), bt.end(), blocks);
  }
  return NVC_OK;
}

int nvc_engine_num_free_blocks(const nvc_engine* e) { return e ? (int)e->scheduler.block_manager.free_.size() : 0; }
int nvc_engine_prefix_cache_stats(const nvc_engine* e, int64_t* hit_blocks, int64_t* queried_blocks) {
  if (!e) return NVC_EINVAL;
  if (hit_blocks) *hit_blocks = e->scheduler.block_manager.cache_hits;
  if (queried_blocks) *queried_blocks = e->scheduler.block_manager.cache_queries;
  return NVC_OK;
}
int nvc_engine_num_waiting(const nvc_engine* e) { return e ? (int)e->scheduler.waiting.size() : 0; }
int nvc_engine_num_running(const nvc_engine* e) { return e ? (int)e->scheduler.running.size() : 0; }

}  // extern "C"
