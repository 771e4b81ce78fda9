"""CPU tests of the tensor-parallel host logic:
  * TPConfig shard arithmetic (src/tp.rs:59-65,94-98);
  * the sharding rules the runner applies (heads / MLP columns / vocab rows; o_proj and down_proj
    by INPUT columns) reproduce the unsharded result when the ranks' partial outputs are summed -
    N ranks hand-built in one process, the way the reference tests its column-parallel linear
    (src/layers/linear.rs:274-319);
  * world_size-2 gloo run: both ranks drive the replicated C++ scheduler and must build
    bit-identical step inputs (block tables, slot mappings) with no communication."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch

from oracle import engine_oracle as E
from oracle import qwen3_oracle as O

tp_mod = importlib.import_module("nano-vllm-candle_b200.tp")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tp_config_reference_vectors():
    c = tp_mod.TPConfig(4, 2, 0)
    assert c.shard_size(100) == 25 and c.shard_offset(100) == 50          # src/tp.rs:94-98
    assert not tp_mod.TPConfig.single().is_distributed()
    c = tp_mod.TPConfig.single().with_size(4).with_rank(2).with_dim(1)     # src/tp.rs:85-91
    assert (c.size, c.rank, c.dim) == (4, 2, 1) and c.is_distributed()
    assert tp_mod.TPConfig.from_env({"TP_SIZE": "4", "TP_RANK": "9"}).rank == 0   # rank >= size -> 0, src/tp.rs:24-29
    o = E.TPConfig.from_env({"TP_SIZE": "4", "TP_RANK": "3"})
    p = tp_mod.TPConfig.from_env({"TP_SIZE": "4", "TP_RANK": "3"})
    assert (o.size, o.rank) == (p.size, p.rank) == (4, 3)


@pytest.mark.parametrize("tp", [2, 4])
def test_sharded_layer_sums_to_unsharded(tp):
    """One attention + MLP block computed by `tp` hand-built ranks equals the unsharded block."""
    torch.manual_seed(0)
    H, nh, kv, hd, I, T = 128, 8, 4, 16, 256, 5
    wq, wk, wv = torch.randn(nh * hd, H), torch.randn(kv * hd, H), torch.randn(kv * hd, H)
    wo, wg, wu, wd = torch.randn(H, nh * hd), torch.randn(I, H), torch.randn(I, H), torch.randn(H, I)
    x = torch.randn(T, H)

    def block(wq, wk, wv, wo, wg, wu, wd, nh, kv):
        q = O.linear(x, wq).reshape(T, nh, hd).transpose(0, 1)
        k = O.gqa_expand(O.linear(x, wk).reshape(T, kv, hd).transpose(0, 1)[None], nh // kv)[0]
        v = O.gqa_expand(O.linear(x, wv).reshape(T, kv, hd).transpose(0, 1)[None], nh // kv)[0]
        p = torch.softmax(q @ k.transpose(1, 2) / hd ** 0.5, dim=-1)
        attn = O.row_parallel_linear((p @ v).transpose(0, 1).reshape(T, nh * hd), wo, None, 0)
        mlp = O.linear(O.silu_and_mul(O.linear(x, torch.cat([wg, wu]))), wd)
        return attn, mlp

    ref_attn, ref_mlp = block(wq, wk, wv, wo, wg, wu, wd, nh, kv)
    attn, mlp = torch.zeros_like(ref_attn), torch.zeros_like(ref_mlp)
    for r in range(tp):
        c = tp_mod.TPConfig(tp, r, 0)
        s = tp_mod.shard_qwen3_layer(c, dict(q=wq, k=wk, v=wv, o=wo, gate=wg, up=wu, down=wd), nh, kv, hd)
        a, m = block(s["q"], s["k"], s["v"], s["o"], s["gate"], s["up"], s["down"], nh // tp, kv // tp)
        attn += a          # the all-reduce (N4)
        mlp += m
    assert torch.allclose(attn, ref_attn, atol=1e-3, rtol=1e-4)
    assert torch.allclose(mlp, ref_mlp, atol=1e-2, rtol=1e-4)


def test_sharded_lm_head_argmax_merge():
    """Vocab-sharded logits + (max, idx) exchange == argmax over the full row, ties -> last index."""
    torch.manual_seed(1)
    V, H, tp = 64, 16, 4
    w, h = torch.randn(V, H), torch.randn(3, H)
    logits = h @ w.t()
    logits[0, 5] = logits[0, 60] = logits[0].max() + 1      # tie across shards
    pairs = []
    for r in range(tp):
        c = tp_mod.TPConfig(tp, r, 0)
        sl = logits[:, c.shard_offset(V):c.shard_offset(V) + c.shard_size(V)]
        pairs.append([(float(row.max()), O.argmax_last(row.numpy()) + c.shard_offset(V)) for row in sl])
    got = [tp_mod.merge_argmax([pairs[r][b] for r in range(tp)]) for b in range(3)]
    assert got == [O.argmax_last(logits[b].numpy()) for b in range(3)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main_chunked(rank, world, port, q):
    """Chunked prefill under TP: the scheduler is replicated, so every rank must build
    the same mixed batches (ids / positions / cu_seqlens / slots / block tables / emit mask) with
    requests arriving while others are mid-prompt."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import zlib
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    eng = importlib.import_module("nano-vllm-candle_b200.engine")
    cfg = eng.SchedulerConfig(6, 40, 0, 20, 16, enable_prefix_caching=True, enable_chunked_prefill=True)
    e = eng.LLMEngine(eng.Scheduler(cfg), eng.DummyModelRunner())
    rng = np.random.default_rng(11)
    base = rng.integers(1, 500, size=50).tolist()
    pending = [base[:int(rng.integers(20, 50))] + rng.integers(1, 500, size=int(rng.integers(1, 70))).tolist() for _ in range(12)]
    digest, step = [], 0
    while pending or not e.is_finished():
        if pending and step % 3 == 0:
            e.add_request(pending.pop(0), eng.SamplingParams(1.0, 21, True))
        step += 1
        if e.is_finished():
            continue
        si = e.schedule()
        digest.append(zlib.crc32(bytes([si.is_prefill, si.num_decode_seqs]) + si.ids.tobytes() + si.positions.tobytes() +
                                 si.cu_seqlens.tobytes() + si.slot_mapping.tobytes() + si.block_tables.tobytes() +
                                 si.context_lens.tobytes() + si.emit.tobytes()))
        e.post_process(np.asarray([si.ids[c - 1] + 1 for c in si.cu_seqlens[1:]], np.int32))
    t = torch.tensor(digest, dtype=torch.int64)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    q.put((rank, all(torch.equal(o, t) for o in out), len(digest)))
    dist.destroy_process_group()


def _rank_main(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    eng = importlib.import_module("nano-vllm-candle_b200.engine")
    e = eng.LLMEngine(eng.Scheduler(eng.SchedulerConfig(6, 4096, 0, 14, 16)), eng.DummyModelRunner())
    rng = np.random.default_rng(7)                      # same seed on every rank: replicated scheduler
    for _ in range(10):
        e.add_request(rng.integers(1, 500, size=int(rng.integers(1, 60))).tolist(), eng.SamplingParams(1.0, 25, True))
    digest = []
    while not e.is_finished():
        si = e.schedule()
        import zlib
        digest.append(zlib.crc32(bytes([si.is_prefill]) + si.ids.tobytes() + si.positions.tobytes() +
                                 si.slot_mapping.tobytes() + si.block_tables.tobytes() + si.context_lens.tobytes()))
        e.post_process((si.ids[-len(si.seq_ids):] + 1) if not si.is_prefill else
                       np.asarray([si.ids[c - 1] + 1 for c in si.cu_seqlens[1:]], np.int32))
    t = torch.tensor(digest, dtype=torch.int64)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    q.put((rank, all(torch.equal(o, t) for o in out), len(digest)))
    dist.destroy_process_group()


@pytest.mark.parametrize("main", ["reference", "chunked"])
def test_replicated_scheduler_world_size_2_gloo(main):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_rank_main if main == "reference" else _rank_main_chunked, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res) and res[0][2] == res[1][2] > 5
