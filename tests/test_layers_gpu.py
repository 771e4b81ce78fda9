"""The reference's own layer tests (src/layers/*.rs #[cfg(test)]), run through the B200 layer
classes that carry the reference's names, plus an Attention prefill -> decode round trip against
the oracle."""
import importlib

import numpy as np
import pytest
import torch

from oracle import qwen3_oracle as O
from test_kernels_gpu import _close

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
L = importlib.import_module("nano-vllm-candle_b200.layers")
tp_mod = importlib.import_module("nano-vllm-candle_b200.tp")


def _pad(t, rows, cols):
    o = torch.zeros(rows, cols)
    o[:t.shape[0], :t.shape[1]] = t
    return o.to(BF16).cuda()


def test_rms_norm_simple(ctx):
    """layernorm.rs:67-89 (rows padded to the 8-element vector width with zeros: the mean is then
    over 8, so compare with the oracle on the same row)."""
    norm = L.RMSNorm(ctx, 8, eps=1e-5)
    x = torch.tensor([[1., 2., 3., 4., 0, 0, 0, 0], [1., 1., 1., 1., 1, 1, 1, 1]])
    out, res = norm.forward(x.to(BF16).cuda())
    assert res is None and out.shape == (2, 8) and out.dtype == BF16       # dtype preserved, :110-121
    ref, _ = O.rms_norm(x, torch.ones(1, 8), 1e-5)
    _close(out, ref, 1e-6, 2 ** -8, "layers rmsnorm simple")
    assert abs(out[1, 0].item() - 1.0) < 1e-2


def test_rms_norm_with_residual(ctx):
    """layernorm.rs:91-108."""
    norm = L.RMSNorm(ctx, 8)
    x = torch.full((1, 8), 0.5, dtype=BF16, device="cuda")
    out, res = norm.forward(x, x.float())
    assert res is not None and torch.equal(res.cpu(), torch.ones(1, 8))
    assert (out.float().cpu() - 1.0).abs().max() < 1e-2


def test_silu_and_mul(ctx):
    """activation.rs:26-35."""
    x = torch.zeros(1, 16)
    x[0, 0], x[0, 1], x[0, 8], x[0, 9] = 0., 1., -1., 2.
    v = L.SiluAndMul(ctx).forward(x.to(BF16).cuda()).float().cpu()
    assert abs(v[0, 0].item()) < 1e-6 and abs(v[0, 1].item() - 1.4621172) < 1e-2


def test_rope_norm_preserved_and_values(ctx):
    """rotary_embedding.rs:115-137 / tests/layer_test.rs:441-503 at head 128, theta 1e6, plus the
    values against the oracle."""
    rope = L.RotaryEmbedding(ctx, 128, 4096, 1e6)
    g = torch.Generator().manual_seed(3)
    q, k = torch.randn(1, 2, 4, 128, generator=g).to(BF16), torch.randn(1, 2, 4, 128, generator=g).to(BF16)
    qr, kr = rope.apply(q.cuda(), k.cuda())
    assert (q.float().square().sum(-1) - qr.float().cpu().square().sum(-1)).abs().sum() < 0.5    # bf16 outputs
    q_ref, k_ref = O.rope_apply(q.float(), k.float(), 1e6)
    _close(qr, q_ref, 1e-5, 2 ** -7, "layers rope q")
    _close(kr, k_ref, 1e-5, 2 ** -7, "layers rope k")
    ones = torch.ones(1, 2, 4, 128, dtype=BF16, device="cuda")
    r, _ = rope.apply(ones, ones)
    assert (r[0, :, 0].float() - r[0, :, 1].float()).abs().sum() > 1.0


def test_replicated_and_column_parallel_linear(ctx):
    """linear.rs:231-319: exact small-integer GEMM; two hand-built column-parallel ranks reproduce
    the halves of the unsharded output."""
    x = torch.tensor([[1., 2., 3., 4.], [-1., 0., 1., 2.]])
    w = torch.tensor([[1., 0., -1., 2.], [0., 1., 2., -1.], [2., -1., 0., 1.],
                      [-2., 1., 1., 0.], [1., 1., 1., 1.], [3., 0., -2., 1.]])
    want = O.linear(x, w)
    xp = _pad(x, 2, 64)
    rep = L.ReplicatedLinear(ctx, 64, 32)
    rep.load_weights(_pad(w, 32, 64))
    assert torch.equal(rep.forward(xp).float().cpu()[:, :6], want)
    full = _pad(torch.cat([w[:3], torch.zeros(13, 4), w[3:], torch.zeros(13, 4)]), 32, 64)   # 2 shards of 16 rows
    for r in range(2):
        col = L.ColumnParallelLinear(ctx, 64, 32, tp=tp_mod.TPConfig(2, r, 0))
        col.load_weights(full)
        assert col.weight.shape == (16, 64)
        assert torch.equal(col.forward(xp).float().cpu()[:, :3], want[:, 3 * r:3 * r + 3])


def test_qkv_parallel_linear_split(ctx):
    """linear.rs:325-353: the fused projection equals three separate matmuls (head size 128)."""
    g = torch.Generator().manual_seed(4)
    H, hd, nh, kv = 256, 128, 2, 1
    wq, wk, wv = [torch.randn(n * hd, H, generator=g) / 16 for n in (nh, kv, kv)]
    qkv = L.QKVParallelLinear(ctx, H, hd, nh, kv)
    qkv.load_weights(torch.cat([wq, wk, wv]).to(BF16).cuda())
    x = torch.randn(3, H, generator=g).to(BF16)
    out = qkv.forward(x.cuda()).float().cpu()
    for sl, w in ((slice(0, 256), wq), (slice(256, 384), wk), (slice(384, 512), wv)):
        _close(out[:, sl], O.linear(x.float(), w.to(BF16).float()), 2e-3, 2 ** -7, "layers qkv split")


def test_row_parallel_ranks_sum_to_unsharded(ctx):
    """Row-parallel shards (input columns) summed == unsharded (the all-reduce, N4, is emulated
    by adding the two ranks' outputs on one GPU)."""
    g = torch.Generator().manual_seed(5)
    w = (torch.randn(64, 256, generator=g) / 16).to(BF16)
    x = torch.randn(5, 256, generator=g).to(BF16)
    total = torch.zeros(5, 64)
    for r in range(2):
        row = L.RowParallelLinear(ctx, 256, 64, tp=tp_mod.TPConfig(1, 0, 1))     # tp_size 1 ctx: no collective
        row.tp_rank, row.tp_size = r, 2
        row.load_weights(w.cuda())
        row.tp_size = 1                                                          # skip the NCCL call
        total += row.forward(x[:, 128 * r:128 * (r + 1)].contiguous().cuda()).float().cpu()
    _close(total, O.linear(x.float(), w.float()), 2e-2, 2 ** -6, "layers row parallel")


def test_attention_prefill_then_decode_matches_dense(ctx):
    """Attention.forward with the per-step Context: prefill writes the cache through slot_mapping,
    one decode step reads it back through block_tables; both equal the reference's dense causal
    attention (qwen3.rs:236-277) over the same tokens."""
    nh, kv, hd, bs = 4, 2, 128, 16
    lens = [21, 40]
    g = torch.Generator().manual_seed(6)
    attn = L.Attention(ctx, nh, hd, hd ** -0.5, kv, block_size=bs)
    attn.k_cache = torch.zeros(16, bs, kv, hd, dtype=BF16, device="cuda")
    attn.v_cache = torch.zeros_like(attn.k_cache)
    tables = [[3, 7], [1, 9, 4]]
    N = sum(lens)
    q, k, v = [torch.randn(N + 2, n * hd, generator=g).to(BF16) for n in (nh, kv, kv)]   # +2 rows: the decode tokens
    slots = [tables[b][t // bs] * bs + t % bs for b, n in enumerate(lens) for t in range(n)]
    cu = torch.tensor([0, lens[0], N], dtype=torch.int32)
    L.set_context(is_prefill=True, cu_seqlens=cu.cuda(), max_seqlen=max(lens),
                  slot_mapping=torch.tensor(slots, dtype=torch.int32).cuda())
    out = attn.forward(q[:N].cuda(), k[:N].cuda(), v[:N].cuda())

    def dense(qs, ks, vs):
        T = ks.shape[0]
        kh = O.gqa_expand(ks.reshape(T, kv, hd).permute(1, 0, 2)[None].float(), nh // kv)[0]
        vh = O.gqa_expand(vs.reshape(T, kv, hd).permute(1, 0, 2)[None].float(), nh // kv)[0]
        qh = qs.reshape(-1, nh, hd).permute(1, 0, 2).float()
        s = qh @ kh.transpose(1, 2) * hd ** -0.5
        Tq = qh.shape[1]
        i = torch.arange(Tq)[:, None] + (T - Tq)
        s = s + torch.where(torch.arange(T)[None, :] > i, torch.tensor(-1e9), torch.tensor(0.0))
        return (torch.softmax(s, -1) @ vh).permute(1, 0, 2).reshape(Tq, nh * hd)

    ref = torch.cat([dense(q[s:e], k[s:e], v[s:e]) for s, e in ((0, lens[0]), (lens[0], N))])
    _close(out, ref, 2e-2, 2 ** -6, "layers attention prefill")
    # decode: token N+b is the next token of sequence b
    dslots = [tables[b][n // bs] * bs + n % bs for b, n in enumerate(lens)]
    bt = torch.tensor([[3, 7, -1], [1, 9, 4]], dtype=torch.int32)
    L.set_context(is_prefill=False, slot_mapping=torch.tensor(dslots, dtype=torch.int32).cuda(),
                  context_lens=torch.tensor([n + 1 for n in lens], dtype=torch.int32).cuda(), block_tables=bt.cuda())
    out = attn.forward(q[N:].cuda(), k[N:].cuda(), v[N:].cuda())
    ref = torch.cat([dense(q[N + b:N + b + 1], torch.cat([k[s:e], k[N + b:N + b + 1]]), torch.cat([v[s:e], v[N + b:N + b + 1]]))
                     for b, (s, e) in enumerate(((0, lens[0]), (lens[0], N)))])
    _close(out, ref, 1e-2, 2 ** -7, "layers attention decode")
