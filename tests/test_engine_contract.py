"""CPU tests of the host side (C++ engine behind the C ABI) against the oracle restatement of
the reference's Scheduler / BlockManager / Sequence / LLMEngine, and of the integer contract
(N1 slot_mapping, N2 block_tables): everything bit-exact, step by step."""
import importlib
import os

import numpy as np
import pytest

from oracle import engine_oracle as E

eng = importlib.import_module("nano-vllm-candle_b200.engine")


def _mk(max_num_seqs, max_tokens_batched, nblocks, bs, eos=0, prefix=False):
    cfg = eng.SchedulerConfig().with_max_num_seqs(max_num_seqs).with_max_num_batched_tokens(max_tokens_batched) \
        .with_eos(eos).with_kvcache(nblocks, bs).with_prefix_caching(prefix)
    cpp = eng.LLMEngine(eng.Scheduler(cfg), eng.DummyModelRunner())
    E.Sequence._counter = 0
    ocfg = E.SchedulerConfig(max_num_seqs, max_tokens_batched, eos, nblocks, bs)
    bm = E.BlockManagerPrefix(nblocks, bs) if prefix else E.BlockManager(nblocks, bs)
    orc = E.LLMEngine(E.Scheduler(ocfg, bm), E.DummyModelRunner(), block_size=bs)
    return cpp, orc


def _oracle_schedule(orc):
    seqs, is_prefill = orc.scheduler.schedule()
    if is_prefill:
        ids, pos, cu, slots = E.prepare_prefill(seqs)
        ctx = np.asarray([len(s) for s in seqs], np.int32)
        bt = E.build_block_tables(seqs)
    else:
        ids, pos, slots, ctx, bt = E.prepare_decode(seqs)
        cu = np.empty(0, np.int32)
    return seqs, is_prefill, ids, pos, cu, slots, bt, ctx


def _lockstep(cpp, orc, prompts, sp_cpp, sp_orc, max_steps=100000):
    for p in prompts:
        cpp.add_request(p, sp_cpp)
        orc.add_request(list(p), sp_orc)
    steps = prefills = 0
    while not orc.is_finished():
        assert not cpp.is_finished()
        si = cpp.schedule()
        seqs, is_prefill, ids, pos, cu, slots, bt, ctx = _oracle_schedule(orc)
        assert si.is_prefill == is_prefill
        assert si.seq_ids.tolist() == [s.seq_id for s in seqs]
        for name, got, want in (("ids", si.ids, ids), ("positions", si.positions, pos), ("cu_seqlens", si.cu_seqlens, cu),
                                ("slot_mapping", si.slot_mapping, slots), ("block_tables", si.block_tables, bt),
                                ("context_lens", si.context_lens, ctx)):
            assert got.dtype == np.int32 and np.array_equal(got, want), f"step {steps}: {name} differs"
        toks = orc.model_runner.run(seqs, is_prefill)
        cpp.post_process(toks)
        orc.scheduler.post_process(seqs, toks)
        assert cpp.num_free_blocks() == orc.scheduler.block_manager.num_free_blocks()
        assert cpp.num_waiting() == len(orc.scheduler.waiting) and cpp.num_running() == len(orc.scheduler.running)
        steps += 1
        prefills += is_prefill
        assert steps < max_steps
    assert cpp.is_finished()
    return steps, prefills


@pytest.mark.parametrize("bs,nblocks,max_seqs,seed", [(256, 64, 8, 0), (16, 40, 6, 1), (16, 23, 16, 2), (32, 300, 64, 3)])
def test_engine_matches_oracle_step_by_step(bs, nblocks, max_seqs, seed):
    rng = np.random.default_rng(seed)
    cpp, orc = _mk(max_seqs, 4096, nblocks, bs, eos=7)
    prompts = [rng.integers(10, 1000, size=int(rng.integers(1, 6 * bs))).tolist() for _ in range(24)]
    # prompts that can never fit the pool would stall both engines forever: clip to pool size
    cap = (nblocks // 2) * bs - 40
    prompts = [p[:cap] for p in prompts]
    steps, prefills = _lockstep(cpp, orc, prompts, eng.SamplingParams(1.0, 37, True), E.SamplingParams(1.0, 37, True))
    assert prefills >= 1 and steps > prefills
    outs = orc.generate([], E.SamplingParams())      # nothing left; outputs were collected above
    for sid in range(len(prompts)):
        got, fin = cpp.output_of(sid)
        assert fin and got == [prompts[sid][-1] + 1 + i for i in range(37)]   # DummyModelRunner: last + 1


def test_preemption_under_kv_pressure_is_identical():
    """Small pool: decode must preempt (scheduler.rs:188-211); block ids stay disjoint."""
    bs, nblocks = 16, 12
    cpp, orc = _mk(8, 4096, nblocks, bs)
    prompts = [[5] * 40, [6] * 33, [7] * 47, [8] * 20]
    _lockstep(cpp, orc, prompts, eng.SamplingParams(1.0, 60, True), E.SamplingParams(1.0, 60, True))
    assert cpp.num_free_blocks() == nblocks


def test_block_ids_are_disjoint_across_running_sequences():
    cpp, _ = _mk(4, 4096, 32, 16)
    sp = eng.SamplingParams(1.0, 5, True)
    for n in (17, 16, 1, 40):
        cpp.add_request(list(range(1, n + 1)), sp)
    si = cpp.schedule()
    used = si.block_tables[si.block_tables >= 0].tolist()
    assert len(used) == len(set(used)) == 2 + 1 + 1 + 3
    # N1: slots of a sequence are block*bs + offset, in order
    assert si.slot_mapping[:17].tolist() == [si.block_tables[0, 0] * 16 + i for i in range(16)] + [si.block_tables[0, 1] * 16]


def test_eos_and_max_tokens_finish_rules():
    """scheduler.rs:224-228: eos stops unless ignore_eos; max_tokens always stops."""
    cpp, _ = _mk(4, 4096, 16, 16, eos=11)
    cpp.add_request([9], eng.SamplingParams(1.0, 64, False))      # dummy runner: 10, 11(eos) -> stop
    cpp.add_request([9], eng.SamplingParams(1.0, 4, True))        # ignore eos: 10, 11, 12, 13
    outs = cpp.generate([], eng.SamplingParams())
    assert [o.token_ids for o in outs] == [[10, 11], [10, 11, 12, 13]]
    assert [o.seq_id for o in outs] == [0, 1]


def test_step_output_token_accounting():
    """llm_engine.rs:251-257: prefill counts sum(len) after the append, decode counts -num_seqs."""
    cpp, _ = _mk(4, 4096, 16, 16)
    sp = eng.SamplingParams(1.0, 3, True)
    cpp.add_request([1, 2, 3], sp)
    cpp.add_request([4, 5], sp)
    s = cpp.step()
    assert s.is_prefill and s.num_tokens == 4 + 3
    s = cpp.step()
    assert not s.is_prefill and s.num_tokens == -2


def test_sampling_params_and_config_builders():
    """src/sampling_params.rs:48-77, src/engine/scheduler.rs:10-56."""
    with pytest.raises(ValueError):
        eng.SamplingParams(temperature=0.0)
    sp = eng.SamplingParams().with_temperature(0.5).with_max_tokens(128).with_ignore_eos(True)
    assert (sp.temperature, sp.max_tokens, sp.ignore_eos) == (0.5, 128, True)
    d = eng.SchedulerConfig()
    assert (d.max_num_seqs, d.max_num_batched_tokens, d.eos, d.num_kvcache_blocks, d.kvcache_block_size) == (1, 4096, 0, 0, 256)


def test_kv_pool_exhaustion_is_an_error_not_an_abort():
    cpp, _ = _mk(2, 4096, 1, 16)
    cpp.add_request([1] * 16, eng.SamplingParams(1.0, 8, True))
    cpp.step()                       # prefill fills the only block
    with pytest.raises(eng.NvcError):
        cpp.step()                   # the reference asserts here (scheduler.rs:175-178)


# ----------------------------------------------------------------------------- prefix caching
def _shared_prefix_prompts(rng, bs, n, n_prefixes=3):
    """Prompts that share block-aligned and unaligned prefixes, plus exact duplicates."""
    prefixes = [rng.integers(10, 1000, size=int(rng.integers(1, 5)) * bs + int(rng.integers(0, bs))).tolist()
                for _ in range(n_prefixes)]
    out = []
    for i in range(n):
        p = prefixes[int(rng.integers(0, n_prefixes))]
        kind = int(rng.integers(0, 4))
        if kind == 0:
            out.append(list(p))                                              # exact duplicate
        elif kind == 1:
            out.append(p[:(len(p) // bs) * bs])                              # exactly the full blocks
        else:
            out.append(p + rng.integers(10, 1000, size=int(rng.integers(1, 3 * bs))).tolist())
    return out


@pytest.mark.parametrize("bs,nblocks,max_seqs,seed", [(16, 64, 4, 0), (16, 40, 8, 1), (32, 48, 3, 2), (256, 40, 6, 3)])
def test_prefix_caching_matches_oracle_step_by_step(bs, nblocks, max_seqs, seed):
    """Block ids, num_cached_tokens (through ids / positions / cu_seqlens), slot_mapping and
    block_tables of the C++ prefix-caching allocator against the restatement of upstream
    nano-vllm's BlockManager, including preemption and cache reuse after deallocation."""
    rng = np.random.default_rng(seed)
    cpp, orc = _mk(max_seqs, 8 * bs * max_seqs, nblocks, bs, eos=7, prefix=True)
    prompts = _shared_prefix_prompts(rng, bs, 20)
    steps, prefills = _lockstep(cpp, orc, prompts, eng.SamplingParams(1.0, 2 * bs + 5, True),
                                E.SamplingParams(1.0, 2 * bs + 5, True))
    hit, queried = cpp.prefix_cache_stats()
    bm = orc.scheduler.block_manager
    assert (hit, queried) == (bm.hits, bm.queries) and hit > 0
    assert cpp.num_free_blocks() == nblocks


def test_prefix_caching_skips_cached_blocks_and_keeps_one_token():
    bs = 16
    cpp, _ = _mk(1, 4096, 32, bs, prefix=True)
    sp = eng.SamplingParams(1.0, 1, True)
    a = list(range(100, 100 + 3 * bs))                  # exactly three full blocks
    cpp.add_request(a, sp)
    si = cpp.schedule()
    assert si.cu_seqlens.tolist() == [0, 3 * bs] and si.positions[0] == 0
    first_table = si.block_tables[0].tolist()
    cpp.post_process([5])                                 # finished (max_tokens 1): blocks freed, hashes kept
    cpp.add_request(a, sp)                                # same prompt: blocks 0 and 1 hit, the last one is recomputed
    si = cpp.schedule()
    assert si.is_prefill and si.cu_seqlens.tolist() == [0, bs]
    assert si.positions.tolist() == list(range(2 * bs, 3 * bs))
    assert si.context_lens.tolist() == [3 * bs]
    assert si.block_tables[0, :2].tolist() == first_table[:2]
    assert si.slot_mapping.tolist() == [si.block_tables[0, 2] * bs + i for i in range(bs)]
    cpp.post_process([5])
    cpp.add_request(a + [1, 2, 3], sp)                    # longer prompt: all three full blocks hit
    si = cpp.schedule()
    assert si.cu_seqlens.tolist() == [0, 3] and si.positions.tolist() == [3 * bs, 3 * bs + 1, 3 * bs + 2]
    assert cpp.prefix_cache_stats() == (5, 9)


def test_prefix_caching_shares_blocks_between_running_sequences():
    bs = 16
    cpp, _ = _mk(4, 4096, 32, bs, prefix=True)
    sp = eng.SamplingParams(1.0, 4, True)
    base = list(range(50, 50 + 2 * bs))
    cpp.add_request(base + [1], sp)
    cpp.add_request(base + [2, 3], sp)                    # scheduled in the same step: hits the blocks just registered
    si = cpp.schedule()
    assert si.block_tables[0, :2].tolist() == si.block_tables[1, :2].tolist()
    assert si.block_tables[0, 2] != si.block_tables[1, 2]
    assert si.cu_seqlens.tolist() == [0, 2 * bs + 1, 2 * bs + 3]
    assert cpp.num_free_blocks() == 32 - 4               # two shared + two private
    cpp.post_process([9, 9])
    assert cpp.num_free_blocks() == 32 - 4


def test_prefix_caching_off_is_the_default_and_never_caches():
    cpp, _ = _mk(2, 4096, 16, 16)
    sp = eng.SamplingParams(1.0, 1, True)
    a = list(range(40))
    cpp.add_request(a, sp); cpp.schedule(); cpp.post_process([1])
    cpp.add_request(a, sp)
    si = cpp.schedule()
    assert si.cu_seqlens.tolist() == [0, 40] and cpp.prefix_cache_stats() == (0, 0)


# ----------------------------------------------------------------------------- chunked prefill
def _mk_chunked(max_num_seqs, budget, nblocks, bs, eos=0, prefix=False):
    cfg = eng.SchedulerConfig().with_max_num_seqs(max_num_seqs).with_max_num_batched_tokens(budget) \
        .with_eos(eos).with_kvcache(nblocks, bs).with_prefix_caching(prefix).with_chunked_prefill(True)
    cpp = eng.LLMEngine(eng.Scheduler(cfg), eng.DummyModelRunner())
    E.Sequence._counter = 0
    ocfg = E.SchedulerConfig(max_num_seqs, budget, eos, nblocks, bs)
    bm = E.BlockManagerPrefix(nblocks, bs) if prefix else E.BlockManager(nblocks, bs)
    return cpp, E.SchedulerChunked(ocfg, bm)


def _lockstep_chunked(cpp, sch, prompts, max_tokens, budget, max_num_seqs, bs, nblocks, arrivals=None, max_steps=100000):
    """Drives the C++ engine and the oracle's SchedulerChunked step by step: identical batches and
    step inputs (bit-exact), plus the invariants of the mode.  arrivals[i] = step at which prompt i
    is added (default: all at step 0).  Returns per-step (num_decode, num_chunk_tokens)."""
    arrivals = arrivals or [0] * len(prompts)
    seqs_by_id, added, log = {}, 0, []
    stored = {}                                            # seq_id -> positions whose K / V are in the cache
    steps = 0
    while added < len(prompts) or not sch.is_finished():
        while added < len(prompts) and arrivals[added] <= steps:
            cpp.add_request(prompts[added], eng.SamplingParams(1.0, max_tokens, True))
            s = E.Sequence(list(prompts[added]), E.SamplingParams(1.0, max_tokens, True), bs)
            seqs_by_id[s.seq_id] = s
            sch.add(s)
            added += 1
        if sch.is_finished():
            steps += 1
            continue
        si = cpp.schedule()
        seqs, q_lens, nd = sch.schedule_chunked()
        ids, pos, cu, slots, ctx, bt, emit = E.prepare_mixed(seqs, q_lens)
        assert si.seq_ids.tolist() == [s.seq_id for s in seqs], f"step {steps}: batch differs"
        assert si.num_decode_seqs == nd and si.is_prefill == (nd < len(seqs))
        for name, got, want in (("ids", si.ids, ids), ("positions", si.positions, pos), ("cu_seqlens", si.cu_seqlens, cu),
                                ("slot_mapping", si.slot_mapping, slots), ("block_tables", si.block_tables, bt),
                                ("context_lens", si.context_lens, ctx), ("emit", si.emit, emit)):
            assert got.dtype == np.int32 and np.array_equal(got, want), f"step {steps}: {name} differs"
        # invariants: token budget, sequence bound, one-token sequences first, slots unique within the step
        assert len(ids) <= budget and len(seqs) <= max_num_seqs and len(sch.running) <= max_num_seqs
        assert all(q == 1 for q in q_lens[:nd]) and all(q >= 1 for q in q_lens)
        assert len(set(slots.tolist())) == len(slots) and slots.min() >= 0 and slots.max() < nblocks * bs
        # every position is computed exactly once per residency, in order, behind a fully stored context
        for s, q in zip(seqs, q_lens):
            have = stored.setdefault(s.seq_id, 0)
            start = s.num_computed_tokens
            assert start == max(have, s.num_cached_tokens) or have == 0, f"step {steps}: seq {s.seq_id} skips or repeats tokens"
            stored[s.seq_id] = start + q
        toks = [s.last_token + 1 for s in seqs]            # DummyModelRunner (llm_engine.rs:20-33)
        cpp.post_process(toks)
        before = {s.seq_id for s in sch.waiting}
        sch.post_process_chunked(seqs, q_lens, toks)
        for s in sch.waiting:                               # preempted: its K / V are gone
            stored[s.seq_id] = 0
        assert cpp.num_free_blocks() == sch.block_manager.num_free_blocks()
        assert cpp.num_waiting() == len(sch.waiting) and cpp.num_running() == len(sch.running)
        log.append((nd, len(ids) - nd))
        steps += 1
        assert steps < max_steps
    assert cpp.is_finished()
    for sid, s in seqs_by_id.items():
        got, fin = cpp.output_of(sid)
        assert fin and got == s.completion_token_ids() == [prompts[sid][-1] + 1 + i for i in range(max_tokens)]
    return log


@pytest.mark.parametrize("bs,nblocks,max_seqs,budget,seed", [(16, 64, 4, 48, 0), (16, 30, 6, 20, 1), (32, 64, 8, 100, 2),
                                                             (256, 24, 5, 300, 3), (16, 25, 16, 16, 4)])
def test_chunked_prefill_matches_oracle_step_by_step(bs, nblocks, max_seqs, budget, seed):
    """Mixed steps: batches, ids / positions / cu_seqlens / slot_mapping / block_tables / context_lens
    and the emit mask of the C++ scheduler against the oracle's SchedulerChunked, with staggered
    arrivals, prompts longer than the budget and KV pressure (preemption of half-computed prompts)."""
    rng = np.random.default_rng(seed)
    cpp, sch = _mk_chunked(max_seqs, budget, nblocks, bs, eos=7)
    cap = (nblocks // 2) * bs - 45
    prompts = [rng.integers(10, 1000, size=int(rng.integers(1, 5 * bs))).tolist()[:cap] for _ in range(20)]
    arrivals = sorted(int(rng.integers(0, 40)) for _ in prompts)
    log = _lockstep_chunked(cpp, sch, prompts, 37, budget, max_seqs, bs, nblocks, arrivals)
    assert any(nd > 0 and nc > 0 for nd, nc in log), "no step mixed decode tokens with prompt chunks"
    assert cpp.num_free_blocks() == nblocks


def test_chunked_prefill_splits_a_long_prompt_and_samples_once():
    bs = 16
    cpp, _ = _mk_chunked(2, 10, 16, bs)
    p = list(range(100, 135))                               # 35 tokens, budget 10 -> chunks 10, 10, 10, 5
    cpp.add_request(p, eng.SamplingParams(1.0, 2, True))
    sizes, emits = [], []
    for _ in range(4):
        si = cpp.schedule()
        sizes.append(len(si.ids)); emits.append(si.emit.tolist())
        assert si.positions.tolist() == list(range(si.context_lens[0] - len(si.ids), si.context_lens[0]))
        assert si.slot_mapping.tolist() == [si.block_tables[0, t // bs] * bs + t % bs for t in si.positions.tolist()]
        cpp.post_process([777])
    assert sizes == [10, 10, 10, 5] and emits == [[0], [0], [0], [1]]
    si = cpp.schedule()                                     # now a pure decode step of the sampled token
    assert not si.is_prefill and si.num_decode_seqs == 1 and si.ids.tolist() == [777] and si.positions.tolist() == [35]
    cpp.post_process([778])
    assert cpp.is_finished() and cpp.output_of(0) == ([777, 778], True)


def test_chunked_prefill_decode_tokens_ride_along_with_a_new_prompt():
    """A running sequence keeps decoding while a newly arrived prompt is prefilled in chunks: the
    reference would stall it for a whole-prompt prefill step (scheduler.rs:106-111)."""
    bs = 16
    cpp, _ = _mk_chunked(4, 12, 32, bs)
    sp = eng.SamplingParams(1.0, 8, True)
    cpp.add_request([5, 6, 7], sp)
    si = cpp.schedule(); cpp.post_process([50])
    assert si.cu_seqlens.tolist() == [0, 3] and si.emit.tolist() == [1]
    cpp.add_request(list(range(200, 230)), sp)              # 30 tokens -> 11 + 11 + 8 next to one decode token
    seen = []
    for step in range(3):
        si = cpp.schedule()
        assert si.num_decode_seqs == 1 and si.seq_ids.tolist() == [0, 1]
        seen.append(int(si.cu_seqlens[2] - si.cu_seqlens[1]))
        cpp.post_process([51 + step, 900])
    assert seen == [11, 11, 8]
    si = cpp.schedule()
    assert not si.is_prefill and si.num_decode_seqs == 2 and si.ids.tolist() == [53, 900]


def test_chunked_prefill_bounds_all_running_sequences():
    """max_num_seqs bounds the running set, not one scheduling round (scheduler.rs:118)."""
    cpp, _ = _mk_chunked(2, 64, 32, 16)
    sp = eng.SamplingParams(1.0, 4, True)
    for i in range(5):
        cpp.add_request([10 + i, 20 + i], sp)
    while not cpp.is_finished():
        si = cpp.schedule()
        assert len(si.seq_ids) <= 2 and cpp.num_running() <= 2
        cpp.post_process([1] * len(si.seq_ids))
    ref, _ = _mk(2, 64, 32, 16)                              # the reference rule admits two more every prefill round
    for i in range(5):
        ref.add_request([10 + i, 20 + i], sp)
    ref.schedule(); ref.post_process([1, 1]); ref.schedule()
    assert ref.num_running() == 4


@pytest.mark.parametrize("bs,nblocks,max_seqs,budget,seed", [(16, 64, 4, 40, 0), (16, 40, 8, 24, 1), (32, 48, 3, 50, 2)])
def test_chunked_prefill_with_prefix_caching_matches_oracle(bs, nblocks, max_seqs, budget, seed):
    """Prefix caching under chunked prefill: a block registered at allocation serves hits only once
    its K / V have been computed (it is filled over several steps)."""
    rng = np.random.default_rng(seed)
    cpp, sch = _mk_chunked(max_seqs, budget, nblocks, bs, eos=7, prefix=True)
    prompts = _shared_prefix_prompts(rng, bs, 20)
    arrivals = sorted(int(rng.integers(0, 60)) for _ in prompts)
    _lockstep_chunked(cpp, sch, prompts, 2 * bs + 5, budget, max_seqs, bs, nblocks, arrivals)
    hit, queried = cpp.prefix_cache_stats()
    bm = sch.block_manager
    assert (hit, queried) == (bm.hits, bm.queries) and hit > 0
    assert cpp.num_free_blocks() == nblocks


def test_chunked_prefill_never_hits_a_block_that_is_not_computed_yet():
    bs = 16
    cpp, _ = _mk_chunked(4, 30, 32, bs, prefix=True)
    sp = eng.SamplingParams(1.0, 1, True)
    a = list(range(100, 100 + 3 * bs + 4))
    cpp.add_request(a, sp)
    si = cpp.schedule(); cpp.post_process([0])             # 30 of 52 tokens: block 0 complete, block 1 partly
    cpp.add_request(a, sp)                                  # same prompt arrives: only block 0 may be shared
    si = cpp.schedule()
    assert si.seq_ids.tolist() == [0, 1]
    assert si.block_tables[1, 0] == si.block_tables[0, 0] and si.block_tables[1, 1] != si.block_tables[0, 1]
    assert si.positions[int(si.cu_seqlens[1])] == bs        # the newcomer starts right after the shared block
    assert cpp.prefix_cache_stats()[0] == 1


def test_chunked_prefill_needs_a_budget_of_at_least_max_num_seqs():
    cfg = eng.SchedulerConfig(8, 4, 0, 16, 16, enable_chunked_prefill=True)
    with pytest.raises(eng.NvcError):
        eng.LLMEngine(eng.Scheduler(cfg), eng.DummyModelRunner())


def test_chunked_step_output_counts_computed_and_sampled_tokens():
    cpp, _ = _mk_chunked(4, 6, 16, 16)
    sp = eng.SamplingParams(1.0, 3, True)
    cpp.add_request([1, 2, 3], sp)
    cpp.add_request(list(range(10, 18)), sp)                # 8 tokens: 3 now, 5 in the next step
    s = cpp.step()
    assert s.is_prefill and (s.num_query_tokens, s.num_decode_seqs, s.num_sampled, s.num_tokens) == (6, 0, 1, 6)
    s = cpp.step()
    assert s.is_prefill and (s.num_query_tokens, s.num_decode_seqs, s.num_sampled) == (6, 1, 2)
    s = cpp.step()
    assert not s.is_prefill and (s.num_query_tokens, s.num_decode_seqs, s.num_sampled, s.num_tokens) == (2, 2, 2, -2)


# ----------------------------------------------------------------------------- property-based (hypothesis)
# 40 derandomised examples per property in the regular suite; NVC_FUZZ_EXAMPLES=5000 runs a deep, randomised search.
_FUZZ_EXAMPLES = int(os.environ.get("NVC_FUZZ_EXAMPLES", "40"))


def test_chunked_scheduler_fuzz_against_oracle():
    """Random pool sizes, budgets, prompt lengths (shared prefixes included), arrival times and prefix caching
    on / off: the C++ chunked scheduler and the oracle's SchedulerChunked must agree step by step and keep the
    invariants of _lockstep_chunked (budget, sequence bound, unique slots, every position computed once per
    residency, outputs of the unchunked engine)."""
    hyp = pytest.importorskip("hypothesis")
    st = hyp.strategies

    @hyp.settings(max_examples=_FUZZ_EXAMPLES, deadline=None, derandomize=_FUZZ_EXAMPLES == 40,
                  suppress_health_check=list(hyp.HealthCheck))
    @hyp.given(st.data())
    def run(data):
        bs = data.draw(st.sampled_from([3, 16, 32]))
        max_seqs = data.draw(st.integers(1, 6))
        budget = data.draw(st.integers(max_seqs, 4 * bs))
        nblocks = data.draw(st.integers(8, 40))
        prefix = data.draw(st.booleans())
        nseq = data.draw(st.integers(1, 8))
        cap = max(1, (nblocks // 2) * bs - 30)               # a prompt + its completion always fits the pool
        base = data.draw(st.lists(st.integers(10, 99), min_size=1, max_size=3 * bs))
        prompts = []
        for _ in range(nseq):
            share = data.draw(st.integers(0, len(base)))
            tail = data.draw(st.lists(st.integers(100, 999), min_size=0 if share else 1, max_size=2 * bs))
            prompts.append((base[:share] + tail)[:cap])
        arrivals = sorted(data.draw(st.lists(st.integers(0, 30), min_size=nseq, max_size=nseq)))
        max_tokens = data.draw(st.integers(1, 20))
        cpp, sch = _mk_chunked(max_seqs, budget, nblocks, bs, eos=7, prefix=prefix)
        _lockstep_chunked(cpp, sch, prompts, max_tokens, budget, max_seqs, bs, nblocks, arrivals, max_steps=20000)
        assert cpp.num_free_blocks() == nblocks

    run()


def test_reference_scheduler_fuzz_against_oracle():
    """The reference's scheduling mode (whole-prompt prefill step or decode step, scheduler.rs:106-249) with the
    real allocator, prefix caching on / off, random pools / prompts: C++ engine == oracle, step by step."""
    hyp = pytest.importorskip("hypothesis")
    st = hyp.strategies

    @hyp.settings(max_examples=_FUZZ_EXAMPLES, deadline=None, derandomize=_FUZZ_EXAMPLES == 40,
                  suppress_health_check=list(hyp.HealthCheck))
    @hyp.given(st.data())
    def run(data):
        bs = data.draw(st.sampled_from([3, 16, 32]))
        max_seqs = data.draw(st.integers(1, 6))
        nblocks = data.draw(st.integers(8, 40))
        prefix = data.draw(st.booleans())
        max_tokens = data.draw(st.integers(1, 20))
        cap = max(1, (nblocks // 2) * bs - max_tokens - 5)   # one sequence always fits, so decode can always preempt its way out
        base = data.draw(st.lists(st.integers(10, 99), min_size=1, max_size=3 * bs))
        prompts = []
        for _ in range(data.draw(st.integers(1, 8))):
            share = data.draw(st.integers(0, len(base)))
            tail = data.draw(st.lists(st.integers(100, 999), min_size=0 if share else 1, max_size=2 * bs))
            prompts.append((base[:share] + tail)[:cap])
        # every prompt fits one prefill step - also after a preemption, when the sequence is re-prefilled with its
        # completion so far (a budget below that stalls the reference for good: prefill refuses, decode has nothing)
        budget = max(len(p) for p in prompts) + max_tokens + data.draw(st.integers(0, 4 * bs))
        cpp, orc = _mk(max_seqs, budget, nblocks, bs, eos=7, prefix=prefix)
        _lockstep(cpp, orc, prompts, eng.SamplingParams(1.0, max_tokens, True), E.SamplingParams(1.0, max_tokens, True),
                  max_steps=20000)
        assert cpp.num_free_blocks() == nblocks
        for sid, p in enumerate(prompts):
            got, fin = cpp.output_of(sid)
            assert fin and got == [p[-1] + 1 + i for i in range(max_tokens)]

    run()


def test_preempted_sequence_that_outgrew_the_budget_stalls_the_reference_but_not_chunked_prefill():
    """Found by the fuzz test: with max_num_batched_tokens = 14, a 14-token prompt that is preempted after 19
    generated tokens can never be prefilled again (33 > 14: scheduler.rs:125-130 refuses, then the decode stage has
    nothing to schedule and the reference asserts, :175-178).  The C++ engine reports it as a status; the chunked
    scheduler re-prefills the sequence in chunks and finishes."""
    bs, nblocks, max_tokens = 16, 8, 20
    prompts = [[100], [100], [100], [100] * 14]
    cpp, orc = _mk(4, 14, nblocks, bs, eos=7)
    sp, osp = eng.SamplingParams(1.0, max_tokens, True), E.SamplingParams(1.0, max_tokens, True)
    for p in prompts:
        cpp.add_request(p, sp)
        orc.add_request(list(p), osp)
    with pytest.raises(AssertionError):
        for _ in range(200):
            seqs, is_prefill = orc.scheduler.schedule()
            orc.scheduler.post_process(seqs, orc.model_runner.run(seqs, is_prefill))
    with pytest.raises(eng.NvcError):
        for _ in range(200):
            cpp.step()
    assert cpp.num_waiting() == 1 and cpp.num_running() == 0          # the stalled sequence, like the oracle's
    assert len(orc.scheduler.waiting) == 1 and not orc.scheduler.running
    chunked, sch = _mk_chunked(4, 14, nblocks, bs, eos=7)
    _lockstep_chunked(chunked, sch, prompts, max_tokens, 14, 4, bs, nblocks)
    assert chunked.is_finished() and chunked.output_of(3) == ([101 + i for i in range(max_tokens)], True)


# ----------------------------------------------------------------------------- C ABI misuse
def test_add_request_refuses_what_the_model_cannot_run():
    """Found by the call-sequence fuzz below: Sequence::new (sequence.rs:31-47) accepts an empty prompt, which has
    no query token to run; in chunked mode it made the scheduler read past the prompt.  The C ABI refuses it."""
    import ctypes as C
    for chunked in (False, True):
        cpp = eng.LLMEngine(eng.Scheduler(eng.SchedulerConfig(2, 16, 7, 4, 16, enable_chunked_prefill=chunked)),
                            eng.DummyModelRunner())
        with pytest.raises(eng.NvcError, match="empty prompt"):
            cpp.add_request([], eng.SamplingParams(1.0, 4, True))
        l, h = cpp.l, cpp.h
        one = (C.c_int32 * 1)(5)
        assert l.nvc_engine_add_request(h, None, 0, 1.0, 4, 1) == -1
        assert l.nvc_engine_add_request(h, None, 3, 1.0, 4, 1) == -1
        assert l.nvc_engine_add_request(h, one, 1, 1.0, -1, 1) == -1
        assert l.nvc_engine_add_request(h, one, 1, float("nan"), 4, 1) == -1
        assert cpp.is_finished() and cpp.num_waiting() == 0
        assert cpp.add_request([5], eng.SamplingParams(1.0, 0, True)) == 0       # max_tokens 0: one token, like :225
        while not cpp.is_finished():
            cpp.step()
        assert cpp.output_of(0) == ([6], True)


def test_engine_c_abi_survives_random_call_sequences():
    """Seeded random call sequences straight at the C ABI (valid and invalid scheduler configurations, requests
    with bad sizes / temperatures / NULL prompts, schedule and post_process out of order or with the wrong batch
    size, preemption of unknown ids, tiny output buffers): every call returns a status, views handed out by
    nvc_engine_step_inputs are readable over their whole documented extent, nothing crashes."""
    import ctypes as C
    import random
    sig = importlib.import_module("nano-vllm-candle_b200._engine_sigs")
    l = importlib.import_module("nano-vllm-candle_b200._lib").lib()
    rng = random.Random(7)
    created = steps_ok = 0
    for it in range(500):
        wild = it % 2 == 0                                   # half the engines get arbitrary configurations
        cfg = sig.SchedulerConfigC()
        cfg.max_num_seqs = rng.choice([-1, 0, 1, 2, 4, 8] if wild else [1, 2, 4, 8])
        cfg.max_num_batched_tokens = rng.choice([-5, 0, 1, 7, 16, 64, 256] if wild else [8, 16, 40, 64, 256])
        cfg.eos = rng.choice([-1, 0, 7])
        cfg.num_kvcache_blocks = rng.choice([-1, 0, 1, 2, 8, 16] if wild else [1, 2, 3, 8, 16])
        cfg.kvcache_block_size = rng.choice([-16, 0, 1, 3, 16, 256] if wild else [2, 3, 16, 16, 256])
        cfg.enable_prefix_caching = rng.randrange(2)
        cfg.enable_chunked_prefill = rng.randrange(2)
        e = C.c_void_p()
        if l.nvc_engine_create(C.byref(cfg), None, C.byref(e)) != 0:
            assert not e.value
            continue
        created += 1
        ids, scheduled = [], None
        for _ in range(rng.randrange(1, 100)):
            op = rng.randrange(10)
            if op <= 2:
                n = rng.choice([-1, 0, 1, 3, 15, 16, 17, 33, 40, 300])
                toks = (C.c_int32 * max(n, 1))(*[rng.randrange(0, 50) for _ in range(max(n, 1))])
                r = l.nvc_engine_add_request(e, toks if rng.random() > 0.05 else None, n,
                                             rng.choice([1.0, 1.0, 0.0, -1.0, 1e-12, float("nan")]),
                                             rng.choice([-1, 0, 1, 5, 30]), rng.randrange(2))
                assert r >= 0 or r == -1
                if r >= 0:
                    ids.append(r)
            elif op == 3:
                info = sig.ScheduleInfoC()
                st = l.nvc_engine_schedule(e, C.byref(info))
                scheduled = info if st == 0 else None
                if st == 0:
                    ps = [C.POINTER(C.c_int32)() for _ in range(6)]
                    sid, em = C.POINTER(C.c_int64)(), C.POINTER(C.c_int32)()
                    assert l.nvc_engine_step_inputs(e, *[C.byref(p) for p in ps], C.byref(sid)) == 0
                    assert l.nvc_engine_step_emit_mask(e, C.byref(em)) == 0
                    nt = info.num_tokens if info.is_prefill else info.num_seqs
                    assert nt > 0 and info.num_seqs > 0
                    touched = sum(ps[0][i] + ps[1][i] + ps[3][i] for i in range(nt))
                    touched += sum(ps[2][i] for i in range(info.num_seqs + 1)) if info.is_prefill else 0
                    if info.max_blocks and ps[4]:
                        touched += sum(ps[4][i] for i in range(info.num_seqs * info.max_blocks))
                    touched += sum(em[i] + sid[i] for i in range(info.num_seqs))
            elif op == 4:
                n = scheduled.num_seqs if scheduled is not None and rng.random() < 0.8 else rng.choice([-1, 0, 1, 5, 100])
                toks = (C.c_int32 * max(n, 1))(*[rng.randrange(0, 50) for _ in range(max(n, 1))])
                l.nvc_engine_post_process(e, toks, n)
                scheduled = None
            elif op <= 7:
                out = sig.StepOutputC()
                steps_ok += l.nvc_engine_step(e, C.byref(out)) == 0
                scheduled = None
            elif op == 8:
                assert l.nvc_engine_preempt(e, rng.choice(ids + [-1, 10**9])) <= 0
            else:
                sid = rng.choice(ids + [-1, 10**9])
                cap = rng.choice([0, 1, 4, 64])
                buf, n = (C.c_int32 * max(cap, 1))(), C.c_int(0)
                l.nvc_engine_get_output(e, sid, buf, cap, C.byref(n))
                l.nvc_engine_seq_block_table(e, sid, buf, cap, C.byref(n))
                assert l.nvc_engine_num_free_blocks(e) >= 0 and l.nvc_engine_num_waiting(e) >= 0
                assert l.nvc_engine_num_running(e) >= 0 and l.nvc_engine_is_finished(e) in (0, 1)
                l.nvc_engine_last_error(e)
        assert l.nvc_engine_destroy(e) == 0
    assert created > 200 and steps_ok > 500, (created, steps_ok)



def test_revived_cached_block_stays_ready_for_later_hits():
    """chunked prefill + prefix caching.  A finishes (its blocks go to the free list, hashes kept); B takes
    them back from the cache; C, admitted while B is still running, must hit the same blocks - a revived block keeps
    its `ready` flag (its K / V are still there).  The oracle mirrors the rule; both must count the same hits."""
    bs = 4
    cfg = eng.SchedulerConfig(4, 64, 0, 16, bs, enable_prefix_caching=True, enable_chunked_prefill=True)
    cpp = eng.LLMEngine(eng.Scheduler(cfg), eng.DummyModelRunner())
    prompt = [5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17]          # 3 full blocks + 1 token
    sp = eng.SamplingParams(1.0, 1, True)
    cpp.add_request(prompt, sp)
    while not cpp.is_finished():
        cpp.step()
    assert cpp.prefix_cache_stats() == (0, 3)
    cpp.add_request(prompt, eng.SamplingParams(1.0, 8, True))          # B: revives A's three blocks
    cpp.step()
    assert cpp.prefix_cache_stats() == (3, 6)
    cpp.add_request(prompt, eng.SamplingParams(1.0, 8, True))          # C: admitted while B runs
    cpp.step()
    assert cpp.prefix_cache_stats() == (6, 9), "a revived block must keep serving prefix-cache hits"
    assert cpp.block_table_of(2)[:3] == cpp.block_table_of(1)[:3]
    while not cpp.is_finished():
        cpp.step()


def test_max_model_len_finishes_sequences_and_refuses_long_prompts():
    """the engine knows the longest sequence the runner holds.  A sequence that grows to it finishes like
    one that reached max_tokens (instead of failing a later step with 'position out of range'); a prompt that is
    longer is refused at add_request."""
    cfg = eng.SchedulerConfig(4, 256, -1, 32, 16)
    cfg.max_model_len = 40
    e = eng.LLMEngine(eng.Scheduler(cfg), eng.DummyModelRunner())
    with pytest.raises(eng.NvcError):
        e.add_request(list(range(41)), eng.SamplingParams(1.0, 4, True))
    a = e.add_request(list(range(30)), eng.SamplingParams(1.0, 100, True))      # would run to 130 tokens
    b = e.add_request(list(range(40)), eng.SamplingParams(1.0, 100, True))      # fits exactly: one token from the prefill
    c = e.add_request(list(range(10)), eng.SamplingParams(1.0, 5, True))        # unaffected
    outs = {o.seq_id: o.token_ids for o in e.generate([], eng.SamplingParams(1.0, 1, True))}
    assert len(outs[a]) == 11          # positions 30..39 computed, the token appended after position 39 ends it
    assert len(outs[b]) == 1 and len(outs[c]) == 5
    assert e.num_free_blocks() == 32
