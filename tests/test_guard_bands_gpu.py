"""Memory-safety evidence without compute-sanitizer (the GPU pool refuses it: profiles/r02_compute_sanitizer_closed.txt).
Every kernel writes its output into the middle of a larger allocation whose bands on both sides hold a sentinel
pattern; after the launch the bands must be untouched, and the inputs sit at the very end of their own allocations
preceded / followed by NaN-filled bands, so a read past an edge poisons the result the parity tests check.  Shapes are
the ragged ones: M, N and sequence lengths that are not multiples of any tile."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
BAND = 4096          # elements on each side
SENT16, SENT32 = 0x7B5A, 0x7B5A7B5A


class Guarded:
    """a [rows, cols] output in the middle of a sentinel-filled buffer"""

    def __init__(self, rows, cols, dtype):
        self.n = rows * cols
        self.int_dtype = torch.int16 if dtype == BF16 else torch.int32
        self.sent = SENT16 if dtype == BF16 else SENT32
        raw = torch.full((2 * BAND + self.n,), self.sent, dtype=self.int_dtype, device="cuda")
        self.raw = raw
        self.view = raw[BAND:BAND + self.n].view(dtype).view(rows, cols)

    def check(self, what):
        torch.cuda.synchronize()
        lo, hi = self.raw[:BAND], self.raw[BAND + self.n:]
        assert bool((lo == self.sent).all()), f"{what}: wrote BEFORE its output buffer"
        assert bool((hi == self.sent).all()), f"{what}: wrote PAST its output buffer"


def edged(t: torch.Tensor) -> torch.Tensor:
    """a copy of t that ends exactly at the end of its allocation's useful part, with NaN bands around it"""
    flat = t.reshape(-1)
    fill = float("nan") if t.is_floating_point() else -(2 ** 30)
    buf = torch.full((2 * BAND + flat.numel(),), fill, dtype=t.dtype, device="cuda")
    buf[BAND:BAND + flat.numel()] = flat.cuda()
    return buf[BAND:BAND + flat.numel()].view(t.shape)


def test_guard_elementwise(ctx):
    g = torch.Generator().manual_seed(1)
    N, H, V = 37, 1000, 513                                   # H = 125 vectors: not a multiple of the warp
    table = edged(torch.randn(V, H, generator=g).to(BF16))
    ids = edged(torch.randint(0, V, (N,), generator=g, dtype=torch.int32))
    out = Guarded(N, H, BF16)
    ctx.embed_gather(ids, table, out=out.view)
    out.check("embed_gather")
    assert torch.equal(out.view.cpu(), table.cpu()[ids.cpu().long()])
    x, w = edged(torch.randn(N, H, generator=g).to(BF16)), edged(torch.randn(H, generator=g).to(BF16))
    res = edged(torch.randn(N, H, generator=g))
    o, ro = Guarded(N, H, BF16), Guarded(N, H, torch.float32)
    ctx.fused_add_rmsnorm(x, w, 1e-6, residual_in=res, residual_out=ro.view, out=o.view)
    o.check("fused_add_rmsnorm out"); ro.check("fused_add_rmsnorm residual_out")
    assert not torch.isnan(o.view.float()).any() and not torch.isnan(ro.view).any()
    act = Guarded(N, 496, BF16)
    ctx.silu_and_mul(edged(torch.randn(N, 992, generator=g).to(BF16)), out=act.view)
    act.check("silu_and_mul")
    assert not torch.isnan(act.view.float()).any()


@pytest.mark.parametrize("M,N,K,out_dtype", [(1, 1024, 1024, BF16), (64, 2000, 512, BF16), (100, 1040, 192, torch.float32),
                                             (129, 272, 64, BF16), (333, 1040, 192, BF16), (300, 4112, 1024, torch.float32),
                                             (64, 1024, 2048, BF16), (64, 65616, 256, torch.float32),
                                             (100, 65552, 128, torch.float32)])   # the last two: swapped-operand lm_head route
def test_guard_gemm(ctx, M, N, K, out_dtype):
    g = torch.Generator().manual_seed(M + N)
    a = edged(torch.randn(M, K, generator=g).to(BF16))
    w = edged((torch.randn(N, K, generator=g) / math.sqrt(K)).to(BF16))
    out = Guarded(M, N, out_dtype)
    ctx.gemm(a, w, out=out.view)
    out.check(f"gemm {M}x{N}x{K}")
    ref = a.float() @ w.float().t()
    assert not torch.isnan(out.view.float()).any()
    assert (out.view.float() - ref).abs().max().item() < 3e-2


@pytest.mark.parametrize("M,I,K", [(64, 3072, 1024), (1, 160, 64), (200, 1008, 320)])
def test_guard_gemm_silu_mul(ctx, M, I, K):
    g = torch.Generator().manual_seed(M + I)
    a = edged(torch.randn(M, K, generator=g).to(BF16))
    w_gu = (torch.randn(2 * I, K, generator=g) / math.sqrt(K)).to(BF16)
    out = Guarded(M, I, BF16)
    ctx.gemm_silu_mul(a, edged(ctx.interleave_gate_up(w_gu)), out=out.view)
    out.check(f"gemm_silu_mul {M}x{I}x{K}")
    assert not torch.isnan(out.view.float()).any()


def test_guard_qk_norm_rope_store_and_attention(ctx):
    """K2 in place + cache scatter, then K4 (varlen and paged) and K5 on the same guarded buffers"""
    g = torch.Generator().manual_seed(3)
    nh, kv, hd, bs = 4, 2, 128, 16
    lens = [17, 130, 1, 65]
    N = sum(lens)
    width = (nh + 2 * kv) * hd
    qkv = Guarded(N, width, BF16)
    qkv.view.copy_(torch.randn(N, width, generator=g).to(BF16))
    nblk = sum((l + bs - 1) // bs for l in lens) + 1
    kc, vc = Guarded(nblk * bs * kv, hd, BF16), Guarded(nblk * bs * kv, hd, BF16)
    kc.view.zero_(); vc.view.zero_()
    tab = ctx.rope_table(1e6, hd, 256)
    pos = edged(torch.cat([torch.arange(l, dtype=torch.int32) for l in lens]))
    bt = torch.full((len(lens), max((l + bs - 1) // bs for l in lens)), -1, dtype=torch.int32)
    slots, nb = [], 0
    for b, l in enumerate(lens):
        n = (l + bs - 1) // bs
        bt[b, :n] = torch.arange(nb, nb + n, dtype=torch.int32)
        slots += [(nb + t // bs) * bs + t % bs for t in range(l)]
        nb += n
    slots = edged(torch.tensor(slots, dtype=torch.int32))
    ones = torch.ones(hd, dtype=BF16, device="cuda")
    kc4, vc4 = kc.view.view(nblk, bs, kv, hd), vc.view.view(nblk, bs, kv, hd)
    ctx.qk_norm_rope_store(qkv.view, ones, ones, 1e-6, tab, pos, nh, kv, hd, slot_mapping=slots, k_cache=kc4, v_cache=vc4)
    qkv.check("qk_norm_rope_store qkv"); kc.check("qk_norm_rope_store k_cache"); vc.check("qk_norm_rope_store v_cache")
    cu = edged(torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32))
    d = qkv.view
    out = Guarded(N, nh * hd, BF16)
    ctx.prefill_attn(d[:, :nh * hd], d[:, nh * hd:(nh + kv) * hd], d[:, (nh + kv) * hd:], cu, max(lens), nh, kv, hd, hd ** -0.5,
                     out=out.view)
    out.check("prefill_attn_varlen")
    assert not torch.isnan(out.view.float()).any()
    ctx_lens = edged(torch.tensor(lens, dtype=torch.int32))
    out2 = Guarded(N, nh * hd, BF16)
    q_rows = d[:, :nh * hd].contiguous()
    ctx.prefill_attn_paged(q_rows, kc4, vc4, edged(bt), ctx_lens, cu, max(lens), nh, kv, hd, bs, hd ** -0.5, out=out2.view)
    out2.check("prefill_attn_paged")
    assert (out2.view.float() - out.view.float()).abs().max().item() < 2e-2
    # decode: the last token of every sequence against its cache
    last = torch.tensor(np.cumsum(lens) - 1, dtype=torch.long, device="cuda")
    out3 = Guarded(len(lens), nh * hd, BF16)
    ctx.paged_decode_attn(edged(q_rows[last]), kc4, vc4, edged(bt), ctx_lens, nh, kv, hd, bs, hd ** -0.5, out=out3.view)
    out3.check("paged_decode_attn")
    assert (out3.view.float() - out.view[last].float()).abs().max().item() < 2e-2


def test_guard_qkv_proj_norm_rope_fused(ctx):
    """K3 + K2 as one kernel (M >= 8192 takes the fused epilogue): qkv rows and both caches inside sentinel bands, M ragged"""
    g = torch.Generator().manual_seed(9)
    nh, kv, hd, bs, K = 4, 2, 128, 16, 256
    M = 8192 + 77
    width = (nh + 2 * kv) * hd
    a = edged(torch.randn(M, K, generator=g).to(BF16))
    w = edged((torch.randn(width, K, generator=g) / math.sqrt(K)).to(BF16))
    qw, kw = edged(torch.randn(hd, generator=g).to(BF16)), edged(torch.randn(hd, generator=g).to(BF16))
    tab = ctx.rope_table(1e6, hd, M)
    pos = edged(torch.arange(M, dtype=torch.int32))
    nblk = (M + bs - 1) // bs
    slots = edged(torch.randperm(nblk * bs, generator=g)[:M].to(torch.int32))
    kc, vc = Guarded(nblk * bs * kv, hd, BF16), Guarded(nblk * bs * kv, hd, BF16)
    kc.view.zero_(); vc.view.zero_()
    kc4, vc4 = kc.view.view(nblk, bs, kv, hd), vc.view.view(nblk, bs, kv, hd)
    qkv = Guarded(M, width, BF16)
    ctx.qkv_proj_norm_rope(a, w, qw, kw, 1e-6, tab, pos, nh, kv, hd, slot_mapping=slots, k_cache=kc4, v_cache=vc4, out=qkv.view)
    qkv.check("qkv_proj_norm_rope qkv"); kc.check("qkv_proj_norm_rope k_cache"); vc.check("qkv_proj_norm_rope v_cache")
    assert not torch.isnan(qkv.view.float()).any()
    # the two-kernel path on the same inputs is bit-identical
    ref = ctx.gemm(a, w)
    kr, vr = torch.zeros_like(kc4), torch.zeros_like(vc4)
    ctx.qk_norm_rope_store(ref, qw, kw, 1e-6, tab, pos, nh, kv, hd, slot_mapping=slots, k_cache=kr, v_cache=vr)
    assert torch.equal(ref, qkv.view) and torch.equal(kr, kc4) and torch.equal(vr, vc4)


def test_guard_head(ctx):
    g = torch.Generator().manual_seed(5)
    B, V = 7, 1040
    logits = edged(torch.randn(B, V, generator=g))
    tok, val = ctx.argmax_last(logits)
    torch.cuda.synchronize()
    assert torch.equal(tok.cpu().long(), logits.cpu().argmax(dim=1))
