"""Pins the CPU oracle against every golden vector / known-answer test the reference's own test
suite holds for this path that can run without real Qwen3 weights."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import engine_oracle as E
from oracle import qwen3_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_replicated_linear_exact():
    """src/layers/linear.rs:231-267."""
    x = torch.tensor([[1., 2., 3., 4.], [-1., 0., 1., 2.]])
    w = torch.tensor([[1., 0., -1., 2.], [0., 1., 2., -1.], [2., -1., 0., 1.],
                      [-2., 1., 1., 0.], [1., 1., 1., 1.], [3., 0., -2., 1.]])
    b = torch.tensor([1., -2., 0., 3., -1., 2.])
    want = torch.tensor([[7., 2., 4., 6., 9., 3.], [3., -2., 0., 6., 1., -1.]])
    assert torch.equal(O.linear(x, w, b), want)


def test_column_parallel_two_ranks_exact():
    """src/layers/linear.rs:274-319: two hand-built ranks of a column-parallel linear."""
    x = torch.tensor([[1., 2., 3., 4.], [-1., 0., 1., 2.]])
    w0 = torch.tensor([[1., 0., -1., 2.], [0., 1., 2., -1.], [2., -1., 0., 1.]])
    w1 = torch.tensor([[-2., 1., 1., 0.], [1., 1., 1., 1.], [3., 0., -2., 1.]])
    assert torch.equal(O.linear(x, w0, torch.tensor([1., -2., 0.])), torch.tensor([[7., 2., 4.], [3., -2., 0.]]))
    assert torch.equal(O.linear(x, w1, torch.tensor([3., -1., 2.])), torch.tensor([[6., 9., 3.], [6., 1., -1.]]))


def test_fused_qkv_equals_separate_matmuls():
    """src/layers/linear.rs:325-353."""
    wq = torch.tensor([[1., 0., -1., 2.], [0., 1., 2., -1.]])
    wk = torch.tensor([[2., -1., 0., 1.]])
    wv = torch.tensor([[-2., 1., 1., 0.]])
    x = torch.tensor([[1., 2., 3., 4.]])
    out = O.linear(x, torch.cat([wq, wk, wv], 0))
    assert torch.equal(out[:, :2], x @ wq.t()) and torch.equal(out[:, 2:3], x @ wk.t()) and torch.equal(out[:, 3:], x @ wv.t())


def test_row_parallel_bias_only_on_rank0():
    """src/layers/linear.rs:184-198."""
    x, w, b = torch.ones(1, 2), torch.ones(3, 2), torch.tensor([1., 2., 3.])
    assert torch.equal(O.row_parallel_linear(x, w, b, 0), torch.tensor([[3., 4., 5.]]))
    assert torch.equal(O.row_parallel_linear(x, w, b, 1), torch.tensor([[2., 2., 2.]]))


def test_rmsnorm_simple_golden():
    """src/layers/layernorm.rs:67-89: [1,2,3,4], eps 1e-5 -> 0.36515 ... 1.46059 (+-1e-4)."""
    x = torch.tensor([[1., 2., 3., 4.], [1., 1., 1., 1.]])
    out, res = O.rms_norm(x, torch.ones(1, 4), 1e-5)
    assert res is None and out.shape == (2, 4)
    assert abs(out[0, 0].item() - 0.36515) < 1e-4 and abs(out[0, 3].item() - 1.46059) < 1e-4


def test_rmsnorm_with_residual_golden():
    """src/layers/layernorm.rs:91-108."""
    x = torch.full((1, 3), 0.5)
    out, res = O.rms_norm(x, torch.ones(1, 3), 1e-6, x.clone())
    assert res.tolist() == [[1.0, 1.0, 1.0]] and abs(out[0, 0].item() - 1.0) < 1e-5


def test_rmsnorm_dtype_preserved():
    """src/layers/layernorm.rs:110-121."""
    out, _ = O.rms_norm(torch.randn(1, 2), torch.ones(1, 2), 1e-6)
    assert out.dtype == torch.float32


def test_silu_and_mul_golden():
    """src/layers/activation.rs:26-35: [0,1,-1,2] -> [0, 1.4621172]."""
    out = O.silu_and_mul(torch.tensor([[0., 1., -1., 2.]]))
    assert abs(out[0, 0].item()) < 1e-6 and abs(out[0, 1].item() - 1.4621172) < 1e-5


def test_rope_preserves_norm_small():
    """src/layers/rotary_embedding.rs:115-137 (head 8, base 1e4, T=4): |diff| < 1e-5."""
    g = torch.Generator().manual_seed(0)
    q = 0.01 * torch.randn(1, 2, 4, 8, generator=g)
    k = 0.01 * torch.randn(1, 2, 4, 8, generator=g)
    qr, kr = O.rope_apply(q, k, 10000.0)
    assert (q.pow(2).sum(-1) - qr.pow(2).sum(-1)).abs().sum() < 1e-5
    assert (k.pow(2).sum(-1) - kr.pow(2).sum(-1)).abs().sum() < 1e-5


def test_rope_values_properties():
    """tests/layer_test.rs:441-503 (head 128, theta 1e6): norms preserved < 1e-3; identical inputs
    at positions 0 and 1 rotate differently (sum |diff| > 1)."""
    g = torch.Generator().manual_seed(1)
    q, k = torch.randn(1, 2, 4, 128, generator=g), torch.randn(1, 2, 4, 128, generator=g)
    qr, _ = O.rope_apply(q, k, 1e6)
    assert (q.square().sum(-1) - qr.square().sum(-1)).abs().sum() < 1e-3
    ones = torch.ones(1, 2, 4, 128)
    r, _ = O.rope_apply(ones, ones, 1e6)
    assert (r[0, :, 0] - r[0, :, 1]).abs().sum() > 1.0
    # position 0 is the identity rotation
    assert torch.equal(r[0, :, 0], ones[0, :, 0])


def test_gqa_expand_is_interleaved_not_tiled():
    """tests/debug_layer_test.rs:10-71: [h0,h1] x2 -> [h0,h0,h1,h1] (numpy.repeat), not tile."""
    x = torch.tensor([1., 2., 3., 4., 5., 6.]).reshape(1, 2, 1, 3)
    got = O.gqa_expand(x, 2).flatten().tolist()
    assert got == [1., 2., 3., 1., 2., 3., 4., 5., 6., 4., 5., 6.]


def test_tp_config_vectors():
    """src/tp.rs:76-98."""
    c = E.TPConfig(1, 0, 0)
    assert (c.size, c.rank) == (1, 0) and not c.is_distributed()
    c = E.TPConfig(4, 2, 0)
    assert c.shard_size(100) == 25 and c.shard_offset(100) == 50


def test_sampling_params_vectors():
    """src/sampling_params.rs:52-76."""
    d = E.SamplingParams()
    assert (d.temperature, d.max_tokens, d.ignore_eos) == (1.0, 64, False)
    with pytest.raises(AssertionError):
        E.SamplingParams(temperature=0.0)


def test_block_manager_stub_vectors():
    """src/engine/block_manager.rs:105-118 + the stub's allocate (identity table, :72-77)."""
    m = E.BlockManagerStub()
    assert m.block_size() == 256 and m.num_blocks() == 0
    m = E.BlockManagerStub(16, 128)
    assert m.num_blocks() == 16 and m.block_size() == 128
    s = E.Sequence(list(range(600)), E.SamplingParams())
    m.allocate(s)
    assert s.block_table == [0, 1, 2] and s.num_cached_tokens == 600


def test_sequence_block_helpers():
    """src/engine/sequence.rs:73-91 (block_size 256 hard-coded at :35)."""
    s = E.Sequence(list(range(600)), E.SamplingParams())
    assert s.num_blocks() == 3 and s.last_block_num_tokens() == 88 and s.num_cached_blocks() == 0
    assert s.block(2) == list(range(512, 600)) and len(s.block(0)) == 256
    s.append_token(7)
    assert len(s) == 601 and s.last_token == 7 and s.num_completion_tokens() == 1


def test_dummy_runner_and_argmax_tie_rule():
    """src/engine/llm_engine.rs:29-33 (last_token + 1) and :135-142 (max_by keeps the LAST max)."""
    s = E.Sequence([5, 9], E.SamplingParams())
    assert E.DummyModelRunner().run([s], True) == [10]
    assert O.argmax_last(np.array([1., 3., 2., 3., 0.], np.float32)) == 3


def test_causality_property_on_synthetic_model():
    """tests/layer_test.rs:164-202: logits of token 0 are the same alone or followed by another
    token (mean |diff| < 0.01), here on the synthetic tiny checkpoint."""
    from oracle import synth
    cfg = O.MODEL_CONFIGS["qwen3-tiny"]
    m = O.Qwen3Oracle(cfg, {k: v.float() for k, v in synth.make_weights(cfg, 0, norm_jitter=0.1).items()})
    one = m.compute_logits(m.forward(torch.tensor([[971]])))[0, 0]
    two = m.compute_logits(m.forward(torch.tensor([[971, 11]])))[0, 0]
    assert (one - two).abs().mean() < 0.01 and not torch.isnan(two).any()


def test_oracle_reproduces_committed_goldens():
    """The committed fixture (python -m oracle.make_goldens --model qwen3-tiny ...) is what the
    oracle computes today."""
    from oracle import synth
    from oracle.make_goldens import generate
    g = json.load(open(os.path.join(GOLD, "qwen3-tiny_cfg0.json")))
    cfg = O.MODEL_CONFIGS[g["model"]]
    m = O.Qwen3Oracle(cfg, {k: v.float() for k, v in synth.make_weights(cfg, g["weight_seed"], norm_jitter=g["norm_jitter"]).items()})
    case = g["cases"][0]
    assert synth.make_prompts(1, g["prompt_len"], cfg.vocab_size, case["prompt_seed"])[0] == case["prompt"]
    toks, runner = generate(m, cfg, case["prompt"], 8)
    assert toks == case["tokens_fp32"][:8]
    assert np.allclose(runner.top[0][1], case["top8_step0"]["val"], atol=1e-5)


def test_hf_transformers_crosscheck():
    """Independent implementation (HF Qwen3ForCausalLM, CPU fp32) on the same synthetic checkpoint."""
    from oracle.crosscheck_hf import crosscheck
    err, cos, _ = crosscheck("qwen3-tiny", seed=0, batch=2, seqlen=24)
    assert err < 1e-4 and cos > 0.99999


def test_kv_cached_cpu_variant_equals_full_recompute():
    """the KV-cached CPU variant that bench.py times next to the reference algorithm is the
    same arithmetic - prefill in two chunks + decode steps reproduce the full-recompute logits."""
    import torch
    from oracle import synth
    cfg = O.MODEL_CONFIGS["qwen3-tiny"]
    w = {k: v.float() for k, v in synth.make_weights(cfg, seed=3, norm_jitter=0.1).items()}
    model = O.Qwen3Oracle(cfg, w, "fp32")
    ids = synth.make_prompts(1, 37, cfg.vocab_size, 77)[0]
    full = model.compute_logits(model.forward(torch.tensor([ids])))[0]
    inc = O.Qwen3OracleKVCached(model, 64)
    got = [inc.step(ids[:20]), inc.step(ids[20:30])] + [inc.step([t]) for t in ids[30:]]
    want = [full[19], full[29]] + [full[i] for i in range(30, 37)]
    for g, w_ in zip(got, want):
        assert (g - w_).abs().max().item() < 2e-4
        assert O.argmax_last(g.numpy()) == O.argmax_last(w_.numpy())
