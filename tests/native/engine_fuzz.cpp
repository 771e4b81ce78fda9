// Native fuzz of the host side under AddressSanitizer / UBSan (built and run by tests/test_native_asan.py, CPU only).
//
// Links csrc/engine.cpp and csrc/hf_loader.cpp against a STUB model runner defined here, so that the
// engine -> runner call path (which the Python CPU tests cannot reach: a real runner needs a device) is exercised:
// the stub checks every step's inputs against the integer contract of SURVEY 8 N1-N3 (slot_mapping, block_tables,
// context_lens, cu_seqlens) exactly as csrc/runner.cu validates them, and answers next = last token + 1 like
// DummyModelRunner (llm_engine.rs:20-33).  It also keeps a model of the paged KV cache (g_kv below).  Random valid and
// invalid call sequences; any contract violation, sanitizer report or wrong output aborts.  test_native_asan.py also
// builds it against deliberately broken copies of engine.cpp and requires a failure.
#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <set>
#include <string>
#include <vector>

#include "../../include/nvc_b200_engine.h"

struct nvc_runner {
  int num_blocks, bs, max_seqs, budget, vocab;
  long prefill = 0, paged = 0, mixed = 0, decode = 0;
};

static std::string g_trace;   // configuration + calls of the current engine, printed when a check fails

#define REQUIRE(c)                                                                                       \
  do {                                                                                                   \
    if (!(c)) {                                                                                          \
      fprintf(stderr, "contract violated at %s:%d: %s\ntrace: %s\n", __FILE__, __LINE__, #c, g_trace.c_str()); \
      abort();                                                                                           \
    }                                                                                                    \
  } while (0)

// Model of the paged KV cache: what (token, position) each slot holds.  Every step writes its new tokens and then
// requires that the whole context of every sequence, read through its block table, is there at the right positions -
// this is what catches a prefix-cache hit on a block that is not computed yet, a block freed while still in use, or
// two sequences writing the same block.
static std::vector<std::pair<int32_t, int32_t>> g_kv;
// Failure injection: when set, the next runner step returns an error BEFORE it touches the cache model (a step that
// failed in the runner wrote nothing).  nvc_engine_step must then roll its batch back (sequences re-queued, the
// prefix-cache entries of their blocks dropped): otherwise a later step reads context that was never written and
// the cache model below goes red.
static bool g_fail_next = false;
static long g_injected = 0;
#define MAYBE_FAIL()            \
  do {                          \
    if (g_fail_next) {          \
      g_fail_next = false;      \
      ++g_injected;             \
      return NVC_EINVAL;        \
    }                           \
  } while (0)

static void check_step(nvc_runner* r, const int32_t* ids, const int32_t* pos, const int32_t* cu, const int32_t* slots,
                       const int32_t* bt, int max_blocks, const int32_t* ctx, int num_decode, int B, int N, int max_seqlen,
                       int32_t* out) {
  REQUIRE(r && ids && pos && cu && slots && out);
  REQUIRE(B > 0 && N > 0 && B <= r->max_seqs && N <= r->budget);
  REQUIRE(cu[0] == 0 && cu[B] == N);
  std::set<int32_t> used;
  for (int b = 0; b < B; ++b) {
    const int q = cu[b + 1] - cu[b];
    REQUIRE(q > 0 && q <= max_seqlen);
    if (b < num_decode) REQUIRE(q == 1);
    for (int i = cu[b]; i < cu[b + 1]; ++i) {
      REQUIRE(ids[i] >= 0 && ids[i] < r->vocab);
      REQUIRE(i == cu[b] || pos[i] == pos[i - 1] + 1);
      REQUIRE(slots[i] >= -1 && slots[i] < (int64_t)r->num_blocks * r->bs);
      if (slots[i] >= 0) REQUIRE(used.insert(slots[i]).second);      // no two tokens of a step share a slot
    }
    if (bt) {
      REQUIRE(ctx && max_blocks > 0);
      REQUIRE(ctx[b] >= q && pos[cu[b + 1] - 1] == ctx[b] - 1);
      const int nb = (ctx[b] + r->bs - 1) / r->bs;
      REQUIRE(nb <= max_blocks);
      for (int j = 0; j < nb; ++j) REQUIRE(bt[(size_t)b * max_blocks + j] >= 0 && bt[(size_t)b * max_blocks + j] < r->num_blocks);
      for (int i = cu[b]; i < cu[b + 1]; ++i)                        // N1: slot = block_table[pos / bs] * bs + pos % bs
        if (slots[i] >= 0) REQUIRE(slots[i] == bt[(size_t)b * max_blocks + pos[i] / r->bs] * r->bs + pos[i] % r->bs);
    } else {
      REQUIRE(pos[cu[b]] == 0);                                      // whole prompt from position 0
    }
    out[b] = (ids[cu[b + 1] - 1] + 1) % r->vocab;
  }
  for (int i = 0; i < N; ++i)
    if (slots[i] >= 0) g_kv[(size_t)slots[i]] = {ids[i], pos[i]};
  if (bt)
    for (int b = 0; b < B; ++b)
      for (int p = 0; p < ctx[b]; ++p)
        REQUIRE(g_kv[(size_t)bt[(size_t)b * max_blocks + p / r->bs] * r->bs + p % r->bs].second == p);
}

extern "C" {
int nvc_runner_prefill(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                       const int32_t* slot_mapping, int B, int N, int max_seqlen, int32_t* out_tokens) {
  MAYBE_FAIL();
  ++r->prefill;
  check_step(r, ids, positions, cu_seqlens, slot_mapping, nullptr, 0, nullptr, 0, B, N, max_seqlen, out_tokens);
  return 0;
}
int nvc_runner_prefill_paged(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                             const int32_t* slot_mapping, const int32_t* block_tables, int max_blocks,
                             const int32_t* context_lens, int B, int N, int max_seqlen, int32_t* out_tokens) {
  MAYBE_FAIL();
  ++r->paged;
  REQUIRE(block_tables);
  check_step(r, ids, positions, cu_seqlens, slot_mapping, block_tables, max_blocks, context_lens, 0, B, N, max_seqlen, out_tokens);
  return 0;
}
int nvc_runner_step_mixed(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* cu_seqlens,
                          const int32_t* slot_mapping, const int32_t* block_tables, int max_blocks,
                          const int32_t* context_lens, int num_decode, int B, int N, int max_seqlen, int32_t* out_tokens) {
  MAYBE_FAIL();
  ++r->mixed;
  REQUIRE(block_tables && num_decode >= 0 && num_decode <= B);
  check_step(r, ids, positions, cu_seqlens, slot_mapping, block_tables, max_blocks, context_lens, num_decode, B, N, max_seqlen, out_tokens);
  return 0;
}
int nvc_runner_decode(nvc_runner* r, const int32_t* ids, const int32_t* positions, const int32_t* slot_mapping,
                      const int32_t* block_tables, int max_blocks, const int32_t* context_lens, int B, int32_t* out_tokens) {
  MAYBE_FAIL();
  ++r->decode;
  REQUIRE(block_tables && B > 0 && B <= r->max_seqs);
  std::vector<int32_t> cu(B + 1);
  for (int b = 0; b <= B; ++b) cu[b] = b;
  for (int b = 0; b < B; ++b) REQUIRE(slot_mapping[b] >= 0);
  nvc_runner wide = *r;
  wide.budget = r->max_seqs;                                         // a decode step is bounded by max_num_seqs only
  check_step(&wide, ids, positions, cu.data(), slot_mapping, block_tables, max_blocks, context_lens, B, B, B, 1, out_tokens);
  return 0;
}
int nvc_runner_set_sampling(nvc_runner*, const float* t, const uint64_t* s, int B, uint64_t) {
  REQUIRE(t && s && B > 0);
  for (int b = 0; b < B; ++b) REQUIRE(t[b] > 1e-10f);
  return 0;
}
int nvc_runner_get_options(const nvc_runner* r, nvc_runner_options* o) {
  memset(o, 0, sizeof(*o));
  o->num_kvcache_blocks = r->num_blocks;
  o->kvcache_block_size = r->bs;
  o->max_num_seqs = r->max_seqs;
  o->max_num_batched_tokens = r->budget;
  o->max_model_len = 1 << 30;
  return 0;
}
int nvc_runner_load_weight(nvc_runner*, const char*, const nvc_bf16*, int64_t, int64_t) { return 0; }
int nvc_runner_finalize(nvc_runner*) { return 0; }
}

// Phase 2: the HF loader's JSON / safetensors readers on mutated files (argv[3] = scratch directory).
static long fuzz_loader(std::mt19937& rng, const std::string& dir, int rounds) {
  const std::string header =
      "{\"__metadata__\":{\"format\":\"pt\",\"note\":\"q \\\" \\u00e9 \\ud83d\\ude00\"},"
      "\"a\":{\"dtype\":\"BF16\",\"shape\":[3,8],\"data_offsets\":[0,48]},"
      "\"b.c\":{\"dtype\":\"F32\",\"shape\":[17],\"data_offsets\":[48,116]},"
      "\"h\":{\"dtype\":\"F16\",\"shape\":[2,2,1],\"data_offsets\":[116,124]},"
      "\"e\":{\"dtype\":\"I64\",\"shape\":[5],\"data_offsets\":[124,164]}}";
  const std::string config =
      "{\n \"vocab_size\": 2048, \"hidden_size\": 256, \"num_hidden_layers\": 3, \"num_attention_heads\": 4,\n"
      " \"num_key_value_heads\": 2, \"intermediate_size\": 512, \"max_position_embeddings\": 4096, \"rms_norm_eps\": 1e-06,\n"
      " \"hidden_act\": \"silu\", \"bos_token_id\": 1, \"eos_token_id\": 2, \"head_dim\": 64, \"rope_theta\": 1000000.0,\n"
      " \"tie_word_embeddings\": true, \"nested\": {\"a\": [1, 2.5e-3, null, true, \"s\\u00e9\"]}}\n";
  static const char* numbers[] = {"0", "-1", "99999999999999999999999", "18446744073709551615", "9223372036854775807", "1e9",
                                  "4294967296", "2147483648", "0.5", "1e400", "-1e400", "01", "1.", "1e", "NaN", "7"};
  auto mutate = [&](std::string t) {
    switch (rng() % 4) {
      case 0: for (int k = 1 + rng() % 5; k > 0; --k) t[rng() % t.size()] = (char)(rng() % 256); break;
      case 1: t.resize(rng() % t.size()); break;
      case 2: {
        std::vector<size_t> starts;
        for (size_t i = 0; i < t.size(); ++i)
          if (isdigit((unsigned char)t[i]) && (i == 0 || !isdigit((unsigned char)t[i - 1]))) starts.push_back(i);
        const size_t b = starts[rng() % starts.size()];
        size_t e2 = b;
        while (e2 < t.size() && isdigit((unsigned char)t[e2])) ++e2;
        t.replace(b, e2 - b, numbers[rng() % 16]);
        break;
      }
      default: for (int k = 1 + rng() % 3; k > 0; --k) t[rng() % t.size()] = "{}[]\",:\\ \0tfn"[rng() % 13]; break;
    }
    return t;
  };
  auto write_file = [](const std::string& path, const std::string& bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    REQUIRE(f != nullptr);
    fwrite(bytes.data(), 1, bytes.size(), f);
    fclose(f);
  };
  long accepted = 0;
  const std::string st_path = dir + "/m.safetensors";
  for (int it = 0; it < rounds; ++it) {
    std::string h = it % 7 == 0 ? header : mutate(header);
    uint64_t hlen = h.size();
    if (it % 11 == 10) hlen = std::vector<uint64_t>{0, 1, h.size() - 1, h.size() + 1, 1ull << 31, 1ull << 63, ~0ull, rng()}[rng() % 8];
    std::string blob((const char*)&hlen, 8);
    blob += h;
    const size_t data_bytes = rng() % 7 == 0 ? rng() % 164 : 164;
    blob.append(data_bytes, '\x3c');
    const bool pristine = it % 7 == 0 && hlen == h.size() && data_bytes == 164;
    write_file(st_path, blob);
    g_trace = "safetensors header: " + h;
    nvc_safetensors* st = nullptr;
    if (nvc_safetensors_open(st_path.c_str(), &st) != 0) {
      REQUIRE(st == nullptr && nvc_hf_last_error()[0] != 0 && !pristine);
    } else {
      ++accepted;
      for (int i = 0; i < nvc_safetensors_num_tensors(st); ++i) {
        const char* name;
        int dt, nd;
        int64_t shape[4];
        REQUIRE(nvc_safetensors_tensor_info(st, i, &name, &dt, &nd, shape) == 0);
        std::vector<nvc_bf16> buf(4096);
        const int rc = nvc_safetensors_read_bf16(st, name, buf.data(), 4096);   // <= 164 data bytes: always fits
        REQUIRE(rc == 0 || dt == NVC_ST_OTHER);
      }
      REQUIRE(nvc_safetensors_close(st) == 0);
    }
    const std::string c = it % 7 == 0 ? config : mutate(config);
    write_file(dir + "/config.json", c);
    g_trace = "config.json: " + c;
    nvc_qwen3_config q;
    int tie = -1;
    if (nvc_qwen3_config_from_hf_dir(dir.c_str(), &q, &tie) == 0) {
      ++accepted;
      REQUIRE(q.vocab_size >= 0 && q.hidden_size >= 0 && q.num_hidden_layers >= 0 && q.num_attention_heads > 0);
      REQUIRE(q.num_key_value_heads >= 0 && q.intermediate_size >= 0 && q.head_dim >= 0 && (tie == 0 || tie == 1));
    } else {
      REQUIRE(nvc_hf_last_error()[0] != 0 && it % 7 != 0);
    }
  }
  return accepted;
}

int main(int argc, char** argv) {
  const unsigned seed = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
  const int engines = argc > 2 ? atoi(argv[2]) : 400;
  std::mt19937 rng(seed);
  if (argc > 3) printf("loader: %ld mutated files accepted and read back\n", fuzz_loader(rng, argv[3], 10 * engines));
  auto pick = [&](std::initializer_list<int> v) { return *(v.begin() + rng() % v.size()); };
  long created = 0, steps_ok = 0, finished = 0, checked = 0;
  nvc_runner totals{};
  for (int it = 0; it < engines; ++it) {
    const bool wild = it % 4 == 0;
    nvc_scheduler_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.max_num_seqs = wild ? pick({-1, 0, 1, 2, 4, 8}) : pick({1, 2, 4, 8});
    cfg.max_num_batched_tokens = wild ? pick({-5, 0, 1, 7, 16, 64, 256}) : pick({8, 16, 40, 64, 256});
    cfg.eos = pick({-1, 0, 7});
    cfg.num_kvcache_blocks = wild ? pick({-1, 0, 1, 2, 8, 16}) : pick({2, 3, 8, 16, 40});
    cfg.kvcache_block_size = wild ? pick({-16, 0, 1, 3, 16, 256}) : pick({2, 3, 16, 16, 256});
    cfg.enable_prefix_caching = rng() % 2;
    cfg.enable_chunked_prefill = rng() % 2;
    nvc_runner runner{cfg.num_kvcache_blocks, cfg.kvcache_block_size, cfg.max_num_seqs, cfg.max_num_batched_tokens, 64};
    const bool with_runner = rng() % 4 != 0;
    g_trace = "cfg seqs " + std::to_string(cfg.max_num_seqs) + " budget " + std::to_string(cfg.max_num_batched_tokens) +
              " eos " + std::to_string(cfg.eos) + " blocks " + std::to_string(cfg.num_kvcache_blocks) + " bs " +
              std::to_string(cfg.kvcache_block_size) + " prefix " + std::to_string(cfg.enable_prefix_caching) + " chunked " +
              std::to_string(cfg.enable_chunked_prefill) + " runner " + std::to_string(with_runner) + " |";
    nvc_engine* e = nullptr;
    if (nvc_engine_create(&cfg, with_runner ? &runner : nullptr, &e) != 0) {
      REQUIRE(e == nullptr);
      continue;
    }
    ++created;
    g_kv.assign((size_t)cfg.num_kvcache_blocks * cfg.kvcache_block_size, {-1, -1});
    if (rng() % 3 == 0) nvc_engine_set_sampling(e, 1, rng());
    struct Req { int64_t id; std::vector<int32_t> prompt; int max_tokens; bool ignore_eos; };
    std::vector<Req> reqs;
    bool clean = true;                                               // no out-of-order post_process / preemption noise
    nvc_schedule_info info;
    bool scheduled = false;
    const int nops = 1 + rng() % 200;
    for (int o = 0; o < nops; ++o) {
      const int op = rng() % 10;
      if (op <= 2) {
        const int n = pick({-1, 0, 1, 3, 15, 16, 17, 33, 40, 300});
        std::vector<int32_t> toks((size_t)std::max(n, 1));
        for (auto& t : toks) t = rng() % 50;
        if (rng() % 2)                                               // shared prefix: prefix-cache hits (8f.1)
          for (size_t i = 0, m = rng() % (toks.size() + 1); i < m; ++i) toks[i] = (int32_t)(i * 7 % 50);
        const double temp = pick({0, 0, 0, 1, 2, 3}) == 0 ? 1.0 : (rng() % 2 ? 0.0 : -1.0);
        const int mt = pick({-1, 0, 1, 5, 30});
        const int ign = rng() % 2;
        const int64_t id = nvc_engine_add_request(e, rng() % 20 ? toks.data() : nullptr, n, temp, mt, ign);
        REQUIRE(id >= 0 || id == NVC_EINVAL);
        if (id >= 0) reqs.push_back({id, toks, mt, ign != 0});
        if (id >= 0) g_trace += " add(n=" + std::to_string(n) + ",max=" + std::to_string(mt) + ",ign=" + std::to_string(ign) + ")";
      } else if (op == 3 || op == 7) {
        // op 7 is a whole step done by hand: schedule, check the inputs, answer last + 1, post_process
        scheduled = nvc_engine_schedule(e, &info) == 0;
        g_trace += scheduled ? (op == 3 ? " sched" : " manual") : " sched!";
        if (scheduled) {
          const int32_t *ids, *pos, *cu, *slots, *bt, *ctx, *emit;
          const int64_t* sid;
          REQUIRE(nvc_engine_step_inputs(e, &ids, &pos, &cu, &slots, &bt, &ctx, &sid) == 0);
          REQUIRE(nvc_engine_step_emit_mask(e, &emit) == 0);
          const int nt = info.is_prefill ? info.num_tokens : info.num_seqs;
          long touched = 0;                                          // ASAN checks the documented extents
          for (int i = 0; i < nt; ++i) touched += ids[i] + pos[i] + slots[i];
          if (info.is_prefill) for (int b = 0; b <= info.num_seqs; ++b) touched += cu[b];
          if (bt && info.max_blocks) for (int i = 0; i < info.num_seqs * info.max_blocks; ++i) touched += bt[i];
          for (int b = 0; b < info.num_seqs; ++b) touched += emit[b] + sid[b];
          checked += touched != -1;
          // a block that receives K / V in this step is written by exactly one sequence of the batch and sits in that
          // sequence's block table (others may hold it as a cached prefix: upstream registers a block's hash at
          // allocation, so a prompt admitted in the same prefill step may already share it - its attention reads it
          // after this step's K / V scatter); block tables never hold a block twice
          std::vector<int> writer((size_t)cfg.num_kvcache_blocks, -1);
          for (int b2 = 0; b2 < info.num_seqs; ++b2) {
            std::vector<int32_t> tbl((size_t)cfg.num_kvcache_blocks + 1);
            int nb = 0;
            REQUIRE(nvc_engine_seq_block_table(e, sid[b2], tbl.data(), (int)tbl.size(), &nb) == 0);
            std::set<int32_t> own(tbl.begin(), tbl.begin() + nb);
            REQUIRE((int)own.size() == nb);
            const int lo = info.is_prefill ? cu[b2] : b2, hi = info.is_prefill ? cu[b2 + 1] : b2 + 1;
            for (int i = lo; i < hi; ++i) {
              if (slots[i] < 0) continue;
              const int32_t blk = slots[i] / cfg.kvcache_block_size;
              REQUIRE(own.count(blk) == 1);
              REQUIRE(writer[(size_t)blk] == -1 || writer[(size_t)blk] == b2);
              writer[(size_t)blk] = b2;
            }
          }
          if (with_runner)                                           // whoever scheduled is assumed to run the model
            for (int i = 0; i < nt; ++i)
              if (slots[i] >= 0) g_kv[(size_t)slots[i]] = {ids[i], pos[i]};
          if (op == 7 && with_runner) {
            // the hand-made step knows the sequences: the cache must hold exactly their tokens
            for (int b = 0; b < info.num_seqs && bt && clean; ++b) {
              const Req* rq = nullptr;
              for (auto& q : reqs) if (q.id == sid[b]) rq = &q;
              REQUIRE(rq != nullptr);
              for (int p = 0; p < ctx[b]; ++p) {
                const int np = (int)rq->prompt.size();
                const int32_t want = p < np ? rq->prompt[(size_t)p] : (rq->prompt.back() + 1 + (p - np)) % 64;
                const auto& cell = g_kv[(size_t)bt[(size_t)b * info.max_blocks + p / cfg.kvcache_block_size] * cfg.kvcache_block_size + p % cfg.kvcache_block_size];
                REQUIRE(cell.first == want && cell.second == p);
              }
            }
          }
          if (op == 7) {
            std::vector<int32_t> next((size_t)info.num_seqs);
            for (int b = 0; b < info.num_seqs; ++b) {
              const int32_t last = info.is_prefill ? ids[cu[b + 1] - 1] : ids[b];
              next[(size_t)b] = with_runner ? (last + 1) % 64 : last + 1;
            }
            REQUIRE(nvc_engine_post_process(e, next.data(), info.num_seqs) == 0);
            scheduled = false;
          }
        }
      } else if (op == 4) {
        const int n = scheduled && rng() % 5 ? info.num_seqs : pick({-1, 0, 1, 5, 100});
        std::vector<int32_t> toks((size_t)std::max(n, 1));
        for (auto& t : toks) t = rng() % 50;
        const int prc = nvc_engine_post_process(e, toks.data(), n);
        if (prc == 0) clean = false;                                 // arbitrary tokens entered a sequence
        g_trace += prc == 0 ? " post" : " post!";
        scheduled = false;
      } else if (op <= 6) {
        nvc_step_output out;
        g_fail_next = with_runner && rng() % 8 == 0;
        g_trace += g_fail_next ? " step(fail)" : " step";
        steps_ok += nvc_engine_step(e, &out) == 0;
        g_fail_next = false;
        scheduled = false;
      } else if (op == 8) {
        const int64_t id = reqs.empty() || rng() % 4 == 0 ? pick({-1, 1000000000}) : reqs[rng() % reqs.size()].id;
        const int prc = nvc_engine_preempt(e, id);
        REQUIRE(prc <= 0);
        if (prc == 0) g_trace += " preempt(" + std::to_string(id) + ")";
      } else {
        const int64_t id = reqs.empty() || rng() % 4 == 0 ? pick({-1, 1000000000}) : reqs[rng() % reqs.size()].id;
        const int cap = pick({0, 1, 4, 64});
        std::vector<int32_t> buf((size_t)std::max(cap, 1));
        int n = 0;
        nvc_engine_get_output(e, id, buf.data(), cap, &n);
        nvc_engine_seq_block_table(e, id, buf.data(), cap, &n);
        REQUIRE(nvc_engine_num_free_blocks(e) >= 0 && nvc_engine_num_waiting(e) >= 0 && nvc_engine_num_running(e) >= 0);
        (void)nvc_engine_last_error(e);
      }
    }
    // drain: with no out-of-order noise every accepted request finishes with prompt[-1] + 1, + 2, ... (mod vocab
    // for the stub, unbounded for the built-in dummy), cut by eos / max_tokens as scheduler.rs:219-249 says
    for (int guard = 0; guard < 20000 && !nvc_engine_is_finished(e); ++guard) {
      nvc_step_output out;
      g_trace += " drain";
      if (nvc_engine_step(e, &out) != 0) break;                      // e.g. the reference-mode stall
    }
    if (nvc_engine_is_finished(e)) {
      REQUIRE(nvc_engine_num_free_blocks(e) == cfg.num_kvcache_blocks);
      for (auto& rq : reqs) {
        std::vector<int32_t> buf(512);
        int n = 0;
        REQUIRE(nvc_engine_get_output(e, rq.id, buf.data(), 512, &n) == 0);
        REQUIRE(n >= 1 && n <= std::max(rq.max_tokens, 1));
        if (!clean) continue;
        for (int i = 0; i < n; ++i) {
          const int32_t want = with_runner ? (rq.prompt.back() + 1 + i) % 64 : rq.prompt.back() + 1 + i;
          REQUIRE(buf[i] == want);
          if (i + 1 < n) REQUIRE(rq.ignore_eos || want != cfg.eos);
        }
        ++finished;
      }
    }
    REQUIRE(nvc_engine_destroy(e) == 0);
    totals.prefill += runner.prefill; totals.paged += runner.paged; totals.mixed += runner.mixed; totals.decode += runner.decode;
  }
  printf("ok engines %ld steps %ld verified outputs %ld views %ld injected %ld runner calls: prefill %ld paged %ld mixed %ld decode %ld\n",
         created, steps_ok, finished, checked, g_injected, totals.prefill, totals.paged, totals.mixed, totals.decode);
  return 0;
}
