"""GPU parity tests: each CUDA kernel, called through the C ABI, against the CPU oracle
(oracle/qwen3_oracle.py, oracle/engine_oracle.py) on the same seeded inputs.
Bar: bit-exact for integer / copy work (store_kvcache, gather, argmax index);
stated tolerances for floating point."""
import math

import numpy as np
import pytest
import torch

from oracle import qwen3_oracle as O
from oracle import engine_oracle as E
from conftest import diag

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


def _bf(x):  # f32 cpu -> bf16 cuda
    return x.to(BF16).cuda()


def _close(got, ref, atol, rtol, name):
    got, ref = got.float().cpu(), ref.float().cpu()
    err = (got - ref).abs()
    tol = atol + rtol * ref.abs()
    bad = err > tol
    if bad.any():
        idx = bad.nonzero()[:8].tolist()
        diag(f"[{name}] FAIL max_err={err.max().item():.4e} bad={bad.float().mean().item():.4f} "
             f"first_bad={idx} got={[got[tuple(i)].item() for i in idx[:4]]} ref={[ref[tuple(i)].item() for i in idx[:4]]}")
        if got.dim() == 2:
            rows_bad = bad.any(dim=1).nonzero().flatten().tolist()
            cols_bad = bad.any(dim=0).nonzero().flatten().tolist()
            diag(f"[{name}] bad rows (first 16) {rows_bad[:16]} of {len(rows_bad)}; bad cols (first 16) {cols_bad[:16]} of {len(cols_bad)}")
    else:
        diag(f"[{name}] ok max_err={err.max().item():.4e}")
    assert not bad.any(), f"{name}: max err {err.max().item():.4e}, {bad.float().mean().item():.2%} elements out of tolerance"


# ----------------------------------------------------------------------------- K0
def test_embed_gather_bit_exact(ctx):
    g = torch.Generator().manual_seed(0)
    table = torch.randn(1000, 256, generator=g).to(BF16)
    ids = torch.randint(0, 1000, (77,), generator=g, dtype=torch.int32)
    out = ctx.embed_gather(ids.cuda(), table.cuda())
    assert torch.equal(out.cpu(), table[ids.long()])


# ----------------------------------------------------------------------------- K1
@pytest.mark.parametrize("N,H", [(1, 1024), (64, 1024), (33, 4096), (5, 5120), (3, 256), (2, 16384)])
@pytest.mark.parametrize("with_res", [False, True])
def test_fused_add_rmsnorm(ctx, N, H, with_res):
    g = torch.Generator().manual_seed(N * 7 + H)
    x = torch.randn(N, H, generator=g).to(BF16)
    w = (1 + 0.1 * torch.randn(H, generator=g)).to(BF16)
    res = torch.randn(N, H, generator=g) if with_res else None
    ref, ref_res = O.rms_norm(x.float(), w.float().reshape(1, -1), 1e-6, res)
    res_d = res.cuda() if with_res else None
    res_out = torch.empty(N, H, dtype=torch.float32, device="cuda")
    out = ctx.fused_add_rmsnorm(x.cuda(), w.cuda(), 1e-6, residual_in=res_d, residual_out=res_out)
    # residual stream is f32 and must be the exact f32 sum
    want_res = ref_res if with_res else x.float()
    assert torch.equal(res_out.cpu(), want_res), "residual sum must be bit-exact"
    # output: one bf16 rounding of the f32 result -> half an ulp of bf16 (2^-9 relative)
    _close(out, ref, 1e-6, 2 ** -8, f"rmsnorm N={N} H={H} res={with_res}")


def test_rmsnorm_reference_goldens(ctx):
    """src/layers/layernorm.rs:67-106 vectors, through the CUDA kernel (row padded to 8)."""
    x = torch.tensor([[1., 2., 3., 4., 0, 0, 0, 0]])
    # mean over 8 instead of 4 -> compare against the oracle on the same padded row
    w = torch.ones(8)
    ref, _ = O.rms_norm(x, w.reshape(1, -1), 1e-5)
    out = ctx.fused_add_rmsnorm(_bf(x), _bf(w), 1e-5)
    _close(out, ref, 1e-6, 2 ** -8, "rmsnorm golden")
    x = torch.full((1, 8), 0.5)
    res_out = torch.empty(1, 8, dtype=torch.float32, device="cuda")
    out = ctx.fused_add_rmsnorm(_bf(x), _bf(w), 1e-6, residual_in=x.cuda(), residual_out=res_out)
    assert torch.equal(res_out.cpu(), torch.ones(1, 8))             # layernorm.rs:99-102
    assert (out.float().cpu() - 1.0).abs().max() < 1e-2            # layernorm.rs:104-106 (bf16 output)


# ----------------------------------------------------------------------------- K6
@pytest.mark.parametrize("N,I", [(1, 3072), (64, 3072), (7, 12288), (3, 8)])
def test_silu_and_mul(ctx, N, I):
    g = torch.Generator().manual_seed(I + N)
    x = (2 * torch.randn(N, 2 * I, generator=g)).to(BF16)
    ref = O.silu_and_mul(x.float())
    out = ctx.silu_and_mul(x.cuda())
    _close(out, ref, 1e-6, 2 ** -8, f"silu_and_mul N={N} I={I}")


def test_silu_reference_golden(ctx):
    """src/layers/activation.rs:26-35: [0,1,-1,2] -> [0, 1.4621172] (d=2; padded to d=8)."""
    x = torch.zeros(1, 16)
    x[0, 0], x[0, 1], x[0, 8], x[0, 9] = 0., 1., -1., 2.
    out = ctx.silu_and_mul(_bf(x)).float().cpu()
    assert abs(out[0, 0].item()) < 1e-6
    assert abs(out[0, 1].item() - 1.4621172) < 1e-2   # bf16 output resolution


# ----------------------------------------------------------------------------- K2 / K2'
def _rope_ref(qkv, qw, kw, eps, positions, nh, kv, hd, theta):
    """oracle: per-head rmsnorm (layernorm.rs:44-60) + neox rope (rotary_embedding.rs:56-107)."""
    N = qkv.shape[0]
    q = qkv[:, :nh * hd].reshape(N, nh, hd).float()
    k = qkv[:, nh * hd:(nh + kv) * hd].reshape(N, kv, hd).float()
    qn, _ = O.rms_norm(q.reshape(-1, hd), qw.float().reshape(1, -1), eps)
    kn, _ = O.rms_norm(k.reshape(-1, hd), kw.float().reshape(1, -1), eps)
    cos, sin = O.build_cos_sin(int(positions.max()) + 1, hd, theta)
    c = cos[positions.long()][:, None, :]
    s = sin[positions.long()][:, None, :]
    return O.apply_rotary(qn.reshape(N, nh, hd), c, s), O.apply_rotary(kn.reshape(N, kv, hd), c, s)


@pytest.mark.parametrize("N,nh,kv", [(1, 16, 8), (64, 16, 8), (37, 32, 8), (5, 4, 2), (3, 5, 1),
                                     (1500, 16, 8)])   # > 1024 CTAs: the prefill ordering of the static loads
def test_qk_norm_rope_store(ctx, N, nh, kv):
    hd, bs, theta = 128, 16, 1e6
    nblk = max(12, (N + bs - 1) // bs + 2)                 # room for one slot per token
    g = torch.Generator().manual_seed(N + nh)
    qkv = torch.randn(N, (nh + 2 * kv) * hd, generator=g).to(BF16)
    qw = (1 + 0.2 * torch.randn(hd, generator=g)).to(BF16)
    kw = (1 + 0.2 * torch.randn(hd, generator=g)).to(BF16)
    positions = torch.randint(0, 3000, (N,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblk * bs, generator=g)[:N].to(torch.int32)
    if N > 2:
        slots[1] = -1                                      # skipped token
    tab = ctx.rope_table(theta, hd, 4096)
    # table itself: host f32 math vs the oracle's f32 math (1e-6: libm vs numpy ulp differences)
    cos, sin = O.build_cos_sin(4096, hd, theta)
    assert (tab.cpu()[:, :64] - cos).abs().max() < 2e-6 and (tab.cpu()[:, 64:] - sin).abs().max() < 2e-6
    kc = torch.zeros(nblk * bs, kv, hd, dtype=BF16, device="cuda")
    vc = torch.zeros_like(kc)
    d = qkv.cuda().clone()
    ctx.qk_norm_rope_store(d, qw.cuda(), kw.cuda(), 1e-6, tab, positions.cuda(), nh, kv, hd,
                           slot_mapping=slots.cuda(), k_cache=kc, v_cache=vc)
    q_ref, k_ref = _rope_ref(qkv, qw, kw, 1e-6, positions, nh, kv, hd, theta)
    got = d.cpu()
    _close(got[:, :nh * hd].reshape(N, nh, hd), q_ref, 1e-5, 2 ** -7, f"q rope N={N}")
    _close(got[:, nh * hd:(nh + kv) * hd].reshape(N, kv, hd), k_ref, 1e-5, 2 ** -7, f"k rope N={N}")
    assert torch.equal(got[:, (nh + kv) * hd:], qkv[:, (nh + kv) * hd:]), "v must be untouched"
    # cache scatter is a bit-exact copy of what the kernel wrote in place (N3)
    kc_ref = np.zeros((nblk * bs, kv, hd), dtype=np.float32)
    vc_ref = np.zeros_like(kc_ref)
    E.store_kvcache(got[:, nh * hd:(nh + kv) * hd].reshape(N, kv, hd).float().numpy(),
                    got[:, (nh + kv) * hd:].reshape(N, kv, hd).float().numpy(), kc_ref, vc_ref, slots.numpy())
    assert np.array_equal(kc.float().cpu().numpy(), kc_ref)
    assert np.array_equal(vc.float().cpu().numpy(), vc_ref)


def test_rope_preserves_norm(ctx):
    """src/layers/rotary_embedding.rs:115-137 / tests/layer_test.rs:441-503 property, on the
    CUDA kernel: with unit norm weights the rotation preserves each head's norm ratio."""
    N, nh, kv, hd = 4, 2, 2, 128
    qkv = torch.randn(N, (nh + 2 * kv) * hd).to(BF16)
    ones = torch.ones(hd, dtype=BF16, device="cuda")
    tab = ctx.rope_table(1e6, hd, 64)
    pos = torch.arange(N, dtype=torch.int32, device="cuda")
    d = qkv.cuda().clone()
    ctx.qk_norm_rope_store(d, ones, ones, 1e-6, tab, pos, nh, kv, hd)
    q = d[:, :nh * hd].reshape(N, nh, hd).float()
    # after rmsnorm every head has mean square 1 -> norm^2 = 128, rotation keeps it
    assert (q.square().sum(-1) - 128).abs().max() < 1.5
    same = torch.ones(N, (nh + 2 * kv) * hd, dtype=BF16, device="cuda")
    ctx.qk_norm_rope_store(same, ones, ones, 1e-6, tab, pos, nh, kv, hd)
    assert (same[0, :hd].float() - same[1, :hd].float()).abs().sum() > 1.0   # positions differ


@pytest.mark.parametrize("N,kv", [(1, 8), (300, 8), (17, 2), (0, 8)])
def test_store_kvcache_bit_exact(ctx, N, kv):
    hd, nslots = 128, 2048
    g = torch.Generator().manual_seed(N)
    buf = torch.randn(max(N, 1), 3 * kv * hd, generator=g).to(BF16)[:N]     # k, v live inside wider rows
    k, v = buf[:, kv * hd:2 * kv * hd], buf[:, 2 * kv * hd:]
    slots = torch.randperm(nslots, generator=g)[:N].to(torch.int32)
    if N > 3:
        slots[2] = -1
    kc = torch.full((nslots, kv, hd), 7.0, dtype=BF16, device="cuda")
    vc = torch.full((nslots, kv, hd), -3.0, dtype=BF16, device="cuda")
    bd = buf.cuda()
    if N:
        ctx.store_kvcache(bd[:, kv * hd:2 * kv * hd], bd[:, 2 * kv * hd:], kc, vc, slots.cuda(), kv, hd)
    kr = np.full((nslots, kv, hd), 7.0, dtype=np.float32)
    vr = np.full((nslots, kv, hd), -3.0, dtype=np.float32)
    E.store_kvcache(k.reshape(N, kv, hd).float().numpy(), v.reshape(N, kv, hd).float().numpy(), kr, vr, slots.numpy())
    assert np.array_equal(kc.float().cpu().numpy(), kr) and np.array_equal(vc.float().cpu().numpy(), vr)


# ----------------------------------------------------------------------------- K7 / K8
def test_gather_rows_bit_exact(ctx):
    x = torch.randn(100, 1024).to(BF16)
    r = torch.randn(100, 1024)
    rows = torch.tensor([99, 0, 5, 5, 42], dtype=torch.int32)
    xo, ro = ctx.gather_rows(x.cuda(), r.cuda(), rows.cuda())
    assert torch.equal(xo.cpu(), x[rows.long()]) and torch.equal(ro.cpu(), r[rows.long()])


@pytest.mark.parametrize("B,V", [(1, 151936), (64, 151936), (3, 1000), (2, 18992), (100, 151936), (200, 32000), (7, 18998),
                                 (74, 16385)])
def test_argmax_last_index(ctx, B, V):
    """rows are split over clusters of 4 / 2 / 1 CTAs by batch (elementwise.cu nvc_argmax_last): every width, ragged V"""
    g = torch.Generator().manual_seed(V)
    logits = torch.randn(B, (V + 3) // 4 * 4, generator=g)[:, :V]
    # force ties: the reference's Iterator::max_by keeps the LAST maximum (llm_engine.rs:135-142)
    mx = logits.max(dim=1).values
    for b in range(B):
        logits[b, (b * 37) % V] = mx[b]
        logits[b, V - 1 - (b * 11) % 50] = mx[b]
    dev = torch.full((B, (V + 3) // 4 * 4), float("inf"), device="cuda")[:, :V]      # +inf in the row padding: never read
    dev.copy_(logits)
    idx, val = ctx.argmax_last(dev, index_offset=5)
    want = [O.argmax_last(logits[b].numpy()) + 5 for b in range(B)]
    assert idx.cpu().tolist() == want
    assert torch.equal(val.cpu(), logits.max(dim=1).values)


# ----------------------------------------------------------------------------- K8s sampling
@pytest.mark.parametrize("B,V,offset", [(5, 2048, 0), (3, 151936, 0), (4, 18992, 37984), (100, 32768, 0), (64, 151936, 0)])
def test_sample_gumbel_matches_oracle(ctx, B, V, offset):
    """Device Gumbel-max draw against oracle/sampling_oracle.py (same Philox counters).  The noise
    bits are exact; logf differs from numpy's float64 log by ulps, so the device's token must be
    the oracle's argmax or lie within 1e-4 of the oracle's best perturbed score; the device's
    winning score must equal the oracle's score of that token to 1e-4."""
    from oracle import sampling_oracle as S
    g = torch.Generator().manual_seed(B * 7 + V)
    logits = torch.randn(B, V, generator=g) * 3.0
    temps = torch.tensor([[0.5, 1.0, 2.0, 1e-9, 0.8][b % 5] for b in range(B)], dtype=torch.float32)
    streams = torch.tensor([S.stream_id(10 + b, 3 * b) for b in range(B)], dtype=torch.int64)
    seed = 0x1234_5678_9ABC_DEF0
    idx, val = ctx.sample_gumbel(logits.cuda(), temps.cuda(), streams.cuda(), seed, index_offset=offset)
    torch.cuda.synchronize()
    exact = 0
    for b in range(B):
        sc = S.gumbel_scores(logits[b].numpy(), float(temps[b]), seed, int(streams[b]), offset)
        t = int(idx[b]) - offset
        assert 0 <= t < V
        tol = 1e-4 * max(1.0, abs(sc.max()))
        assert sc.max() - sc[t] <= tol, f"row {b}: token {t} scores {sc[t]} vs best {sc.max()}"
        assert abs(float(val[b]) - sc[t]) <= tol
        exact += int(S.sample(logits[b].numpy(), float(temps[b]), seed, int(streams[b]), offset) == int(idx[b]))
    diag(f"[sample_gumbel B={B} V={V}] {exact}/{B} draws identical to the oracle")
    assert exact >= B - 1 - B // 32


@pytest.mark.parametrize("B,V", [(64, 151936), (3, 18992), (150, 32768), (8, 65536)])
def test_sample_gumbel_screening_is_exact(ctx, B, V):
    """The fast upper-bound screen (elementwise.cu sample_gumbel_kernel) only skips elements that cannot win: tokens and
    winning scores are bit-identical with the screen switched off, incl. rows of huge / tiny temperature, -inf masks, ties."""
    import importlib
    nvc = importlib.import_module("nano-vllm-candle_b200")
    g = torch.Generator().manual_seed(B + V)
    logits = torch.randn(B, V, generator=g) * 4.0
    logits[0, : V // 2] = float("-inf")                       # a masked half row
    logits[1 % B] = 1.25                                      # a constant row: the noise alone decides
    logits[2 % B, ::7] = 30.0                                 # many equal front-runners
    temps = torch.tensor([[0.5, 1.0, 2.0, 1e-9, 0.8, 50.0][b % 6] for b in range(B)], dtype=torch.float32).cuda()
    streams = (torch.arange(B, dtype=torch.int64) * 977 + 5).cuda()
    dev = logits.cuda()
    outs = []
    for off in (1, 0):
        nvc.lib().nvc_debug_set(6, off)
        try:
            outs.append([ctx.sample_gumbel(dev, temps, streams, seed, index_offset=8) for seed in (1, 2, 0xDEADBEEF)])
            torch.cuda.synchronize()
        finally:
            nvc.lib().nvc_debug_set(6, 0)
    for (i0, v0), (i1, v1) in zip(*outs):
        assert torch.equal(i0, i1) and torch.equal(v0, v1)


def test_sample_gumbel_distribution_and_reproducibility(ctx):
    """4096 independent streams over a 16-way row: frequencies follow the distribution the
    reference samples from (llm_engine.rs:100-125); same (seed, stream) -> same token; the draw
    of a row does not depend on which other rows share the launch."""
    from oracle import sampling_oracle as S
    rng = np.random.default_rng(3)
    row = rng.normal(0, 1.5, size=16).astype(np.float32)
    n = 16384
    logits = torch.from_numpy(np.tile(row, (n, 1))).cuda()
    temps = torch.full((n,), 0.9, dtype=torch.float32).cuda()
    streams = torch.arange(n, dtype=torch.int64).cuda()
    idx, _ = ctx.sample_gumbel(logits, temps, streams, 42)
    idx2, _ = ctx.sample_gumbel(logits, temps, streams, 42)
    sub, _ = ctx.sample_gumbel(logits[100:110], temps[100:110], streams[100:110].contiguous(), 42)
    other, _ = ctx.sample_gumbel(logits, temps, streams, 43)
    torch.cuda.synchronize()
    assert torch.equal(idx, idx2) and torch.equal(sub, idx[100:110]) and not torch.equal(idx, other)
    counts = np.bincount(idx.cpu().numpy(), minlength=16)
    e = n * S.reference_probs(row, 0.9)
    small = e < 5
    c2 = np.append(counts[~small], counts[small].sum()); e2 = np.append(e[~small], e[small].sum())
    chi2 = float((((c2 - e2) ** 2)[e2 > 0] / e2[e2 > 0]).sum())
    diag(f"[sample_gumbel distribution] chi2 {chi2:.1f} over {len(e2)} bins")
    assert chi2 < 45.0
