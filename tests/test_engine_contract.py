"""CPU tests of the host side (C++ engine behind the C ABI) against the oracle restatement of
the reference's Scheduler / BlockManager / Sequence / LLMEngine, and of the integer contract
(N1 slot_mapping, N2 block_tables): everything bit-exact, step by step."""
import importlib

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
