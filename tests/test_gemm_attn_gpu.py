"""GPU parity tests for the tensor-core kernels (tcgen05 GEMM, tcgen05 prefill attention) and
the paged decode attention, against the fp32 oracle."""
import math

import numpy as np
import pytest
import torch

from oracle import qwen3_oracle as O
from conftest import diag
from test_kernels_gpu import _close

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


# ----------------------------------------------------------------------------- K3 GEMM
GEMM_SHAPES = [
    # (M, N, K, force_bn, force_splits)
    (128, 128, 64, 128, 1),        # one tile, one k-block
    (128, 128, 512, 128, 1),       # pipeline wraps (8 k-blocks > 6 stages)
    (128, 256, 256, 256, 1),
    (128, 64, 256, 64, 1),
    (128, 32, 256, 32, 1),
    (256, 512, 1024, 128, 1),      # several tiles
    (64, 1024, 2048, 0, 0),        # decode shapes, auto config (0.6B o_proj)
    (64, 4096, 1024, 0, 0),        # 0.6B qkv
    (64, 6144, 1024, 0, 0),        # 0.6B gate_up
    (64, 1024, 3072, 0, 0),        # 0.6B down
    (1, 1024, 1024, 0, 0),         # M = 1
    (200, 4096, 1024, 0, 0),       # ragged M
    (333, 1040, 192, 128, 1),      # ragged M and N (N % 16 == 0 only)
    (64, 1024, 2048, 64, 4),       # forced split-K
    (130, 256, 1024, 128, 8),      # split-K with 2 m-tiles
    (64, 2000, 512, 32, 2),        # split-K, ragged N, BN=32
    (64, 1024, 3072, 2128, 4),     # two k-blocks per stage (KPS=2) + split-K, ragged last stage (12 k-blocks / 4 / 2)
    (64, 1024, 1088, 4064, 1),     # KPS=4 with a partial last stage (17 k-blocks)
    (64, 4096, 1024, 4032, 1),     # KPS=4, BN=32
    (100, 512, 320, 1128, 1),      # KPS=1, 5 k-blocks
    (64, 1024, 2048, 84016, 1),    # cluster of 8 CTAs, activation tile multicast, BN=16
    (64, 4096, 1024, 44032, 1),    # cluster of 4, BN=32
    (64, 6144, 1024, 44064, 1),    # cluster of 4, BN=64
    (1, 1024, 1024, 84016, 1),     # M=1 through the cluster path (rows zero-filled up to 8 per CTA)
    (130, 1024, 512, 44032, 1),    # two m-tiles, second one ragged, clusters along N
    (256, 6144, 4096, 42096, 1),   # BN=96 (UMMA N=96), cluster 4: 8B qkv at batch 256
    (256, 4096, 4096, 42064, 1),   # 8B o_proj at batch 256
    (256, 2048, 512, 21256, 1),    # BN=256 cluster of 2
    (256, 6144, 4096, 0, 0),       # 8B qkv at batch 256
    (2048, 1024, 1024, 0, 0),      # prefill-sized
    # split-K inside a cluster (1,1,S), partial tiles reduced over distributed shared memory: force_splits = 100 + S
    (64, 1024, 2048, 4032, 104),   # 0.6B o_proj: BN=32, 4 splits of 8 k-blocks
    (64, 1024, 3072, 4032, 104),   # 0.6B down_proj
    (64, 1024, 2048, 2064, 108),   # BN=64, cluster of 8
    (1, 1024, 1024, 4032, 104),    # M = 1
    (100, 2000, 1088, 4032, 104),  # ragged N, 17 k-blocks (5+5+5+2)
    (128, 512, 4096, 2128, 102),   # BN=128, 2 splits
    (64, 1024, 2048, 1256, 103),   # BN=256 (128 KB partial tile in the ring), 3 splits
    # CTA-pair kernel (gemm_pair.cu): force_bn = 512, force_splits = in-cluster split-K factor (0 = auto)
    (256, 256, 64, 512, 1),        # one pair tile, one k-block
    (129, 256, 64, 512, 1),        # second CTA of the pair holds a single valid row
    (128, 256, 128, 512, 1),       # second CTA's activation box entirely out of bounds (zero-filled)
    (256, 512, 1024, 512, 1),      # ring wraps (16 k-blocks > 6 stages)
    (300, 1040, 192, 512, 1),      # ragged M and N (N % 16 == 0 only)
    (4096, 2048, 512, 512, 1),     # 128 tiles: persistent, two tiles per cluster (both accumulators)
    (8192, 4096, 256, 512, 1),     # 512 tiles: ~7 per cluster, accumulator + ring phases wrap
    (256, 4096, 4096, 512, 4),     # 8B o_proj at batch 256: split-K 4 in clusters of 8, DSMEM exchange
    (256, 6144, 4096, 512, 3),     # 8B qkv: split 3 (uneven k ranges 22/22/20), clusters of 6
    (256, 4096, 12288, 512, 4),    # 8B down_proj
    (256, 24576, 4096, 512, 2),    # 8B gate_up: 96 tiles x 2 splits, several waves of clusters
    (200, 784, 448, 512, 2),       # split 2, ragged M and N, 7 k-blocks (4 + 3)
    (256, 512, 64, 512, 4),        # more splits than k-blocks: clamps to 1
    (256, 4096, 4096, 512, 0),     # auto split choice
    (512, 151936 // 8, 1024, 512, 0),   # lm_head shard, ragged last n-tile
    (256, 24576, 4096, 0, 0),      # the same through the schedule choice
    # narrow pair tiles (one MMA of N = NT < 256, two accumulator stages): force_splits = 10000 * S + 1000 + NT
    (256, 1536, 512, 512, 1192),   # 8 tiles of 192
    (300, 2240, 192, 512, 1224),   # NT = 224, ragged M, exact N
    (256, 1040, 256, 512, 1160),   # NT = 160, ragged last tile (80 valid columns)
    (512, 1024, 1024, 512, 1128),  # NT = 128, two m-tiles, ring wraps
    (4096, 3072, 256, 512, 1192),  # 256 tiles of 192: persistent, accumulator stages alternate
    (256, 4096, 4096, 512, 31192), # 8B o_proj: 22 tiles of 192 (last one 64 wide) x split 3
    (256, 6144, 4096, 512, 21192), # 8B qkv: 32 tiles x split 2
    (256, 4096, 12288, 512, 41128),# 8B down: NT 128, split 4
    (256, 24576, 4096, 512, 1192), # 8B gate_up shape: 128 tiles, two full waves
]


@pytest.mark.parametrize("M,N,K,bn,splits", GEMM_SHAPES)
@pytest.mark.parametrize("out_dtype", [BF16, torch.float32])
def test_gemm_bf16(ctx, M, N, K, bn, splits, out_dtype):
    g = torch.Generator().manual_seed(M * 31 + N * 7 + K)
    a = torch.randn(M, K, generator=g).to(BF16)
    w = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(BF16)
    ref = O.linear(a.float(), w.float())
    out = ctx.gemm(a.cuda(), w.cuda(), out_dtype=out_dtype, force_bn=bn, force_splits=splits)
    torch.cuda.synchronize()
    # fp32 accumulation of exact bf16 products: only summation order differs (1e-3 abs on O(1)
    # values); bf16 output adds one rounding (2^-8 relative)
    rtol = 2 ** -8 if out_dtype == BF16 else 1e-5
    _close(out, ref, 2e-3, rtol, f"gemm M={M} N={N} K={K} bn={bn} s={splits} {out_dtype}")


@pytest.mark.parametrize("M,N,K", [
    (64, 151936, 1024),      # 0.6B lm_head at the bench's decode batch: 594 tiles of 256 weight rows, the last one half full
    (40, 65616, 512),        # tokens not a multiple of the 32-row box (second CTA's token rows 40..63 zero-filled), ragged vocab
    (100, 70000, 1024),      # 65..128 tokens: 128-column tiles
    (128, 65536, 256),       # full 128 tokens, 4 k-blocks (shorter than the ring)
    (1, 66000, 1024),        # a single token
    (64, 1040, 192),         # small N: 5 tiles, fewer than the pairs (forced only)
    (48, 37984, 1024),       # the TP 4 vocabulary shard
    (128, 65536, 2048),      # 65..128 tokens take it from K = 2 048
])
def test_gemm_lm_head_swapped_operands(ctx, M, N, K):
    """lm_head at decode batch sizes with the operands swapped in the pair kernel (weights as the 256-row MMA operand, fp32
    logits stored transposed; gemm_pair.cu nvc_gemm_pair_swapped): against the fp32 oracle, against the unswapped pair
    schedule, with a row stride on both sides, and as the automatic route of vocabulary-sized fp32 outputs."""
    g = torch.Generator().manual_seed(M + N + K)
    a_full = torch.randn(M, K + 64, generator=g).to(BF16).cuda()
    a = a_full[:, :K]                                          # lda > K
    w = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(BF16).cuda()
    ref = O.linear(a.float().cpu(), w.float().cpu())
    buf = torch.full((M, N + 16), float("nan"), dtype=torch.float32, device="cuda")
    out = ctx.gemm(a, w, out=buf[:, :N], force_bn=513)         # ldc > N: the padding columns must stay NaN
    torch.cuda.synchronize()
    _close(out, ref, 2e-3, 1e-5, f"swapped lm_head M={M} N={N} K={K}")
    assert bool(torch.isnan(buf[:, N:]).all())
    plain = ctx.gemm(a, w, out_dtype=torch.float32, force_bn=512)
    diag(f"[lm_head swapped M={M} N={N} K={K}] bit-identical with the unswapped pair schedule: {torch.equal(plain, out)}, "
         f"max diff {(plain - out).abs().max().item():.2e}")
    assert (plain - out).abs().max().item() < 1e-3
    auto = ctx.gemm(a, w, out_dtype=torch.float32)
    if (M <= 64 and N >= 32768) or N >= 131072 or (N >= 65536 and K >= 2048):
        assert torch.equal(auto, out)                          # the automatic route IS the swapped kernel


@pytest.mark.parametrize("M,nh,kv,K", [
    (8200, 16, 8, 1024),      # 0.6B qkv at a prefill size: 33 m-tiles (last one ragged) x 16 tiles of two heads
    (32768, 16, 8, 1024),     # the bench's prefill step
    (8192, 5, 1, 512),        # 7 heads = 896 columns: the last 256-column tile holds one head
    (9000, 32, 8, 4096),      # 8B shape, M not a multiple of 128
    (1500, 16, 8, 1024),      # below the fusion threshold: the two-kernel path
    (300, 4, 2, 256),         # few tiles: whatever schedule the library picks, the results must not change
    (64, 16, 8, 1024),        # decode-sized: the two-kernel path
])
def test_qkv_proj_norm_rope_fused(ctx, nvc, M, nh, kv, K):
    """K3 + K2 as one kernel (the QKV GEMM's persistent epilogue applies the q/k-norm, RoPE and the KV scatter) against
    the two-kernel sequence on the same operands: qkv rows and both caches bit-identical; and against the oracle
    (qwen3.rs:205-234, rotary_embedding.rs:82-107) through the existing K2 reference."""
    hd, bs, theta, eps = 128, 16, 1e6, 1e-6
    g = torch.Generator().manual_seed(M + nh + K)
    N = (nh + 2 * kv) * hd
    a = torch.randn(M, K, generator=g).to(BF16).cuda()
    w = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(BF16).cuda()
    qw = (1 + 0.2 * torch.randn(hd, generator=g)).to(BF16).cuda()
    kw = (1 + 0.2 * torch.randn(hd, generator=g)).to(BF16).cuda()
    pos = torch.randint(0, 4000, (M,), generator=g, dtype=torch.int32).cuda()
    nslots = ((M + bs - 1) // bs + 2) * bs
    slots = torch.randperm(nslots, generator=g)[:M].to(torch.int32)
    if M > 2:
        slots[1] = -1                                         # skipped token
    slots = slots.cuda()
    tab = ctx.rope_table(theta, hd, 4096)
    res = {}
    for fused in (0, 1):
        nvc.lib().nvc_debug_set(7, 0 if fused else 1)
        try:
            kc = torch.zeros(nslots, kv, hd, dtype=BF16, device="cuda")
            vc = torch.zeros_like(kc)
            qkv = ctx.qkv_proj_norm_rope(a, w, qw, kw, eps, tab, pos, nh, kv, hd, slot_mapping=slots, k_cache=kc, v_cache=vc)
            torch.cuda.synchronize()
            res[fused] = (qkv, kc, vc)
        finally:
            nvc.lib().nvc_debug_set(7, 0)
    for name, x, y in zip(("qkv", "k_cache", "v_cache"), res[0], res[1]):
        assert torch.equal(x.view(torch.int16), y.view(torch.int16)), f"fused QKV epilogue: {name} differs from GEMM + K2"
    # oracle: linear, then per-head rmsnorm + rope on the bf16 projection
    from test_kernels_gpu import _rope_ref
    proj = ctx.gemm(a, w).cpu()
    q_ref, k_ref = _rope_ref(proj, qw.cpu(), kw.cpu(), eps, pos.cpu(), nh, kv, hd, theta)
    got = res[1][0].cpu()
    _close(got[:, :nh * hd].reshape(M, nh, hd), q_ref, 1e-5, 2 ** -7, f"fused qkv epilogue q M={M}")
    _close(got[:, nh * hd:(nh + kv) * hd].reshape(M, kv, hd), k_ref, 1e-5, 2 ** -7, f"fused qkv epilogue k M={M}")
    assert torch.equal(got[:, (nh + kv) * hd:], proj[:, (nh + kv) * hd:]), "v columns must be the projection itself"


def test_gemm_reference_goldens(ctx):
    """src/layers/linear.rs:231-267 (replicated, exact small integers; bias added on the host
    here since the Qwen3 path has no biases) and :274-319 (two hand-sharded column ranks),
    zero-padded to the kernel's K % 64 / N % 16 granularity."""
    x = torch.tensor([[1., 2., 3., 4.], [-1., 0., 1., 2.]])
    w = torch.tensor([[1., 0., -1., 2.], [0., 1., 2., -1.], [2., -1., 0., 1.],
                      [-2., 1., 1., 0.], [1., 1., 1., 1.], [3., 0., -2., 1.]])
    bias = torch.tensor([1., -2., 0., 3., -1., 2.])
    want = torch.tensor([[7., 2., 4., 6., 9., 3.], [3., -2., 0., 6., 1., -1.]])
    xp = torch.zeros(2, 64); xp[:, :4] = x
    wp = torch.zeros(16, 64); wp[:6, :4] = w
    out = ctx.gemm(xp.to(BF16).cuda(), wp.to(BF16).cuda(), out_dtype=torch.float32).cpu()
    assert torch.equal(out[:, :6] + bias, want)
    # column-parallel: rank r owns rows [3r, 3r+3) of W; concat of the shards == unsharded
    for r in range(2):
        ws = torch.zeros(16, 64); ws[:3, :4] = w[3 * r:3 * r + 3]
        o = ctx.gemm(xp.to(BF16).cuda(), ws.to(BF16).cuda(), out_dtype=torch.float32).cpu()
        assert torch.equal(o[:, :3] + bias[3 * r:3 * r + 3], want[:, 3 * r:3 * r + 3])


SILU_SHAPES = [
    # (M, I, K, force_bn, force_splits)
    (64, 3072, 1024, 0, 0),        # 0.6B gate_up at batch 64 (1-CTA tiles)
    (1, 128, 64, 0, 0),
    (100, 1008, 192, 2064, 1),     # ragged: last 64-wide tile holds 32 valid columns
    (128, 1024, 512, 1256, 1),     # BN = 256
    (64, 768, 512, 42096, 1),      # BN = 96, activation multicast cluster of 4
    (256, 12288, 4096, 0, 0),      # 8B gate_up at batch 256 (pair kernel, schedule chosen)
    (256, 1600, 1024, 512, 2),     # pair kernel, split-K 2 (32B/TP8-like shard), ragged last tile
    (300, 512, 448, 512, 3),       # pair kernel, split-K 3, ragged M
    (2048, 3072, 1024, 0, 0),      # prefill-sized, persistent pair schedule
    (256, 12288, 4096, 512, 1192), # 8B gate_up at batch 256 on 192-wide pair tiles (6 gate/up groups per tile)
    (256, 880, 256, 512, 21160),   # NT = 160 (5 units), split 2, ragged last tile
    (300, 448, 320, 512, 1224),    # NT = 224, ragged M
]


@pytest.mark.parametrize("M,I,K,bn,splits", SILU_SHAPES)
def test_gemm_silu_mul_fused(ctx, M, I, K, bn, splits):
    """K3 + K6 fused: the oracle computes silu(x Wg^T) * (x Wu^T) in fp32 (qwen3.rs:323-327, activation.rs:13-18)."""
    g = torch.Generator().manual_seed(M * 17 + I * 3 + K)
    a = torch.randn(M, K, generator=g).to(BF16)
    w = (torch.randn(2 * I, K, generator=g) / math.sqrt(K)).to(BF16)
    gu = O.linear(a.float(), w.float())
    ref = O.silu_and_mul(gu)
    out = ctx.gemm_silu_mul(a.cuda(), ctx.interleave_gate_up(w.cuda()), force_bn=bn, force_splits=splits)
    torch.cuda.synchronize()
    _close(out, ref, 2e-3, 2 ** -8, f"gemm+silu_mul M={M} I={I} K={K} bn={bn} s={splits}")


def test_gemm_pair_split_determinism(ctx):
    """in-cluster split-K sums the partials in split order whichever CTA reduces: bit-identical reruns"""
    g = torch.Generator().manual_seed(11)
    a = torch.randn(256, 4096, generator=g).to(BF16).cuda()
    w = (torch.randn(4096, 4096, generator=g) / 64).to(BF16).cuda()
    outs = [ctx.gemm(a, w, out_dtype=torch.float32, force_bn=512, force_splits=4) for _ in range(3)]
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    big = torch.randn(256, 3 * 1024, generator=g).to(BF16).cuda()
    av = big[:, 1024:2048]                                   # row stride 3072
    w2 = (torch.randn(512, 1024, generator=g) / 32).to(BF16).cuda()
    ref = O.linear(av.float().cpu(), w2.float().cpu())
    _close(ctx.gemm(av, w2, force_bn=512, force_splits=2), ref, 2e-3, 2 ** -8, "pair gemm strided split-K")


def test_gemm_strided_input_and_determinism(ctx):
    g = torch.Generator().manual_seed(5)
    big = torch.randn(96, 3 * 512, generator=g).to(BF16).cuda()
    a = big[:, 512:1024]                                    # row stride 1536
    w = (torch.randn(1024, 512, generator=g) / 22).to(BF16).cuda()
    ref = O.linear(a.float().cpu(), w.float().cpu())
    o1 = ctx.gemm(a, w, force_bn=64, force_splits=4)
    o2 = ctx.gemm(a, w, force_bn=64, force_splits=4)
    _close(o1, ref, 2e-3, 2 ** -8, "gemm strided split-K")
    assert torch.equal(o1, o2), "split-K must be deterministic (fixed reduction order)"


# ----------------------------------------------------------------------------- K4 prefill
def _dense_causal_ref(q, k, v, nh, kv, scale):
    """oracle attention core, src/models/qwen3.rs:236-277, one sequence: q [T,nh,hd], k/v [T,kv,hd]."""
    T = q.shape[0]
    qh = q.permute(1, 0, 2)[None].float()
    kh = O.gqa_expand(k.permute(1, 0, 2)[None].float(), nh // kv)
    vh = O.gqa_expand(v.permute(1, 0, 2)[None].float(), nh // kv)
    s = (qh @ kh.transpose(2, 3)) * scale
    idx = torch.arange(T)
    s = s + torch.where(idx[None, :] > idx[:, None], torch.tensor(-1e9), torch.tensor(0.0))
    p = torch.softmax(s, dim=-1)
    return (p @ vh)[0].permute(1, 0, 2).reshape(T, -1)


@pytest.mark.parametrize("lens,nh,kv", [
    ([128], 2, 1),
    ([1], 2, 2),
    ([17, 128, 129, 300, 1, 512], 4, 2),
    ([640, 33], 16, 8),
    ([1000], 8, 1),
])
def test_prefill_attn_varlen(ctx, nvc, lens, nh, kv):
    hd = 128
    N = sum(lens)
    g = torch.Generator().manual_seed(N + nh)
    qkv = torch.randn(N, (nh + 2 * kv) * hd, generator=g).to(BF16)
    cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
    scale = hd ** -0.5
    ref = torch.cat([_dense_causal_ref(
        qkv[s:e, :nh * hd].reshape(-1, nh, hd), qkv[s:e, nh * hd:(nh + kv) * hd].reshape(-1, kv, hd),
        qkv[s:e, (nh + kv) * hd:].reshape(-1, kv, hd), nh, kv, scale) for s, e in zip(cu[:-1], cu[1:])])
    d = qkv.cuda()
    out = ctx.prefill_attn(d[:, :nh * hd], d[:, nh * hd:(nh + kv) * hd], d[:, (nh + kv) * hd:],
                           cu.cuda(), max(lens), nh, kv, hd, scale)
    # P is rounded to bf16 before P*V and the output is bf16: 2e-2 abs on O(1) values
    _close(out, ref, 2e-2, 2 ** -6, f"prefill lens={lens} nh={nh} kv={kv}")


@pytest.mark.parametrize("nh,kv,block_size,cases", [
    # cases: (context_len, q_len) per sequence; the q rows are the LAST q_len positions of the context
    (4, 2, 16, [(48, 16), (33, 1), (200, 200), (131, 3), (16, 16)]),
    (4, 2, 32, [(300, 44), (128, 128), (129, 1)]),
    (16, 8, 256, [(768, 256), (1000, 232), (256, 1), (257, 257), (640, 130)]),   # block-aligned and unaligned prefixes
    (8, 1, 128, [(1500, 476), (5, 5)]),
    (2, 2, 64, [(513, 385)]),
])
def test_prefill_attn_paged(ctx, nh, kv, block_size, cases):
    """K4 over the paged cache (prefix caching): rows of dense causal attention over the whole
    context (qwen3.rs:236-277) for the last q_len positions; unwritten cache slots are NaN."""
    hd = 128
    B = len(cases)
    g = torch.Generator().manual_seed(nh * 1000 + block_size + B)
    nblk_per = [(c + block_size - 1) // block_size for c, _ in cases]
    total_blk = sum(nblk_per) + 2
    perm = torch.randperm(total_blk, generator=g)
    bt = torch.full((B, max(nblk_per) + 1), -1, dtype=torch.int32)
    kc = torch.full((total_blk * block_size, kv, hd), float("nan")).to(BF16)
    vc = torch.full((total_blk * block_size, kv, hd), float("nan")).to(BF16)
    scale = hd ** -0.5
    qs, refs, p = [], [], 0
    for b, (c, ql) in enumerate(cases):
        blocks = perm[p:p + nblk_per[b]]; p += nblk_per[b]
        bt[b, :len(blocks)] = blocks.to(torch.int32)
        q = torch.randn(c, nh, hd, generator=g).to(BF16)
        k = torch.randn(c, kv, hd, generator=g).to(BF16)
        v = torch.randn(c, kv, hd, generator=g).to(BF16)
        t = torch.arange(c)
        slots = blocks[t // block_size] * block_size + t % block_size
        kc[slots], vc[slots] = k, v
        refs.append(_dense_causal_ref(q, k, v, nh, kv, scale)[c - ql:])
        qs.append(q[c - ql:].reshape(ql, nh * hd))
    qd = torch.cat(qs).cuda()
    ref = torch.cat(refs)
    cu = torch.tensor([0] + list(np.cumsum([ql for _, ql in cases])), dtype=torch.int32)
    ctx_lens = torch.tensor([c for c, _ in cases], dtype=torch.int32)
    out = ctx.prefill_attn_paged(qd, kc.reshape(total_blk, block_size, kv, hd).cuda(),
                                 vc.reshape(total_blk, block_size, kv, hd).cuda(), bt.cuda(), ctx_lens.cuda(), cu.cuda(),
                                 max(ql for _, ql in cases), nh, kv, hd, block_size, scale)
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "poisoned (unwritten) cache slots leaked into the output"
    err = (out.float().cpu() - ref).abs().max().item()
    diag(f"[prefill paged nh={nh} kv={kv} bs={block_size} cases={cases}] max_err={err:.4e}")
    _close(out, ref, 2e-2, 2 ** -6, f"paged prefill nh={nh} kv={kv} bs={block_size} cases={cases}")


def test_prefill_attn_persistent_many_items(ctx, nvc):
    """more work items than resident CTAs (2 per SM): every CTA of the persistent grid walks several (q tile, head,
    sequence) items - ragged lengths, q tiles past the end of short sequences (skipped items), GQA groups."""
    test_prefill_attn_varlen(ctx, nvc, [513, 1, 700, 128, 129, 64, 1024, 300, 257, 33, 900, 512], 16, 8)
    test_prefill_attn_varlen(ctx, nvc, [200] * 40, 8, 2)


@pytest.mark.parametrize("lens,nh,kv", [([2048, 1500], 8, 2), ([4096], 4, 1), ([64, 65, 63, 191, 192, 193], 2, 1)])
def test_prefill_attn_long_sequences(ctx, nvc, lens, nh, kv):
    """Many kv tiles per q tile: the tensor-memory output accumulator is rescaled across up to 64 tiles;
    tile-boundary lengths around the 64-token kv tile."""
    test_prefill_attn_varlen(ctx, nvc, lens, nh, kv)


def test_prefill_attn_paged_equals_varlen_without_prefix(ctx):
    """With no cached prefix the paged kernel must reproduce the in-place varlen kernel bit for bit
    (same tiles, same order of operations)."""
    hd, nh, kv, bs = 128, 8, 4, 256
    lens = [300, 128, 77, 513]
    N = sum(lens)
    g = torch.Generator().manual_seed(7)
    qkv = torch.randn(N, (nh + 2 * kv) * hd, generator=g).to(BF16).cuda()
    cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32).cuda()
    scale = hd ** -0.5
    a = ctx.prefill_attn(qkv[:, :nh * hd], qkv[:, nh * hd:(nh + kv) * hd], qkv[:, (nh + kv) * hd:], cu, max(lens), nh, kv, hd, scale)
    nblk = [(l + bs - 1) // bs for l in lens]
    total = sum(nblk)
    kc = torch.zeros(total * bs, kv * hd, dtype=BF16, device="cuda")
    vc = torch.zeros(total * bs, kv * hd, dtype=BF16, device="cuda")
    bt = torch.full((len(lens), max(nblk)), -1, dtype=torch.int32)
    blk, s = 0, 0
    for b, l in enumerate(lens):
        bt[b, :nblk[b]] = torch.arange(blk, blk + nblk[b], dtype=torch.int32)
        kc[blk * bs:blk * bs + l] = qkv[s:s + l, nh * hd:(nh + kv) * hd]
        vc[blk * bs:blk * bs + l] = qkv[s:s + l, (nh + kv) * hd:]
        blk += nblk[b]; s += l
    b_out = ctx.prefill_attn_paged(qkv[:, :nh * hd], kc.reshape(total, bs, kv, hd), vc.reshape(total, bs, kv, hd), bt.cuda(),
                                   torch.tensor(lens, dtype=torch.int32).cuda(), cu, max(lens), nh, kv, hd, bs, scale)
    torch.cuda.synchronize()
    assert torch.equal(a, b_out)


# ----------------------------------------------------------------------------- K5 decode
@pytest.mark.parametrize("nh,kv,block_size,ctx_lens", [
    (16, 8, 256, [1, 16, 17, 255, 256, 257, 1000, 64]),
    (32, 8, 256, [700, 3, 1025]),
    (4, 4, 16, [5, 40]),
    (40, 8, 256, [300, 1]),          # group 5 (Qwen3-14B)
    (64, 8, 256, [513]),             # group 8 (Qwen3-32B): two chunks of 4
    (16, 8, 256, [2048]),            # single sequence -> split-KV + combine
])
def test_paged_decode_attn(ctx, nh, kv, block_size, ctx_lens):
    hd = 128
    B = len(ctx_lens)
    g = torch.Generator().manual_seed(nh * 100 + B)
    nblk_per = [(c + block_size - 1) // block_size for c in ctx_lens]
    total_blk = sum(nblk_per) + 3
    perm = torch.randperm(total_blk, generator=g).tolist()
    bt = torch.full((B, max(nblk_per)), -1, dtype=torch.int32)
    kc = torch.full((total_blk * block_size, kv, hd), float("nan")).to(BF16)    # unwritten slots are poison
    vc = torch.full((total_blk * block_size, kv, hd), float("nan")).to(BF16)
    q = torch.randn(B, nh, hd, generator=g).to(BF16)
    ref = torch.empty(B, nh * hd)
    scale = hd ** -0.5
    p = 0
    for b, c in enumerate(ctx_lens):
        blocks = perm[p:p + nblk_per[b]]; p += nblk_per[b]
        bt[b, :len(blocks)] = torch.tensor(blocks, dtype=torch.int32)
        k = torch.randn(c, kv, hd, generator=g).to(BF16)
        v = torch.randn(c, kv, hd, generator=g).to(BF16)
        for t in range(c):
            slot = blocks[t // block_size] * block_size + t % block_size
            kc[slot], vc[slot] = k[t], v[t]
        # oracle: last row of dense causal attention == attention of q over all c cached tokens
        kh = O.gqa_expand(k.permute(1, 0, 2)[None].float(), nh // kv)[0]
        vh = O.gqa_expand(v.permute(1, 0, 2)[None].float(), nh // kv)[0]
        s = torch.einsum("hd,htd->ht", q[b].float(), kh) * scale
        ref[b] = torch.einsum("ht,htd->hd", torch.softmax(s, dim=-1), vh).reshape(-1)
    out = ctx.paged_decode_attn(q.reshape(B, nh * hd).cuda(), kc.cuda(), vc.cuda(), bt.cuda(),
                                torch.tensor(ctx_lens, dtype=torch.int32).cuda(), nh, kv, hd, block_size, scale)
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "poisoned (unwritten) cache slots leaked into the output"
    _close(out, ref, 1e-2, 2 ** -7, f"paged decode nh={nh} kv={kv} bs={block_size} ctx={ctx_lens}")


def _paged_case(nh, kv, block_size, ctx_lens, seed):
    """random paged KV cache (poisoned where unwritten) + the oracle's attention of q over each sequence's cache"""
    hd = 128
    B = len(ctx_lens)
    g = torch.Generator().manual_seed(seed)
    nblk_per = [(c + block_size - 1) // block_size for c in ctx_lens]
    total_blk = sum(nblk_per) + 3
    perm = torch.randperm(total_blk, generator=g)
    bt = torch.full((B, max(max(nblk_per), 1)), -1, dtype=torch.int32)
    kc = torch.full((total_blk * block_size, kv, hd), float("nan")).to(BF16)
    vc = torch.full((total_blk * block_size, kv, hd), float("nan")).to(BF16)
    q = torch.randn(B, nh, hd, generator=g).to(BF16)
    ref = torch.zeros(B, nh * hd)
    scale = hd ** -0.5
    p = 0
    for b, c in enumerate(ctx_lens):
        blocks = perm[p:p + nblk_per[b]]; p += nblk_per[b]
        bt[b, :len(blocks)] = blocks.to(torch.int32)
        if c == 0:
            continue
        k = torch.randn(c, kv, hd, generator=g).to(BF16)
        v = torch.randn(c, kv, hd, generator=g).to(BF16)
        t = torch.arange(c)
        slots = blocks[t // block_size] * block_size + t % block_size
        kc[slots], vc[slots] = k, v
        kh = O.gqa_expand(k.permute(1, 0, 2)[None].float(), nh // kv)[0]
        vh = O.gqa_expand(v.permute(1, 0, 2)[None].float(), nh // kv)[0]
        sc = torch.einsum("hd,htd->ht", q[b].float(), kh) * scale
        ref[b] = torch.einsum("ht,htd->hd", torch.softmax(sc, dim=-1), vh).reshape(-1)
    return q, kc, vc, bt, ref, scale


def _lens(lens_spec):
    if lens_spec[0] == "uniform":
        _, n, lo, hi, seed = lens_spec
        return torch.randint(lo, hi + 1, (n,), generator=torch.Generator().manual_seed(seed)).tolist()
    if lens_spec[0] == "const":
        return [lens_spec[2]] * lens_spec[1]
    return lens_spec[1]


@pytest.mark.parametrize("nh,kv,block_size,lens_spec,splits", [
    (16, 8, 256, ("uniform", 48, 1, 700, 1), 0),     # ragged batch
    (16, 8, 256, ("const", 64, 581), 0),             # the bench shape: 512 CTAs, one resident wave of 2-warp CTAs
    (32, 8, 64, ("uniform", 20, 60, 3000, 2), 3),    # long ragged contexts, forced KV splits (some own no tile)
    (8, 1, 256, ("uniform", 100, 1, 300, 3), 0),     # one kv head (a TP=8 shard), group 8
    (16, 8, 256, ("list", [300, 0, 5, 0, 129]), 0),  # empty sequences: zero rows
    (40, 8, 16, ("uniform", 9, 1, 200, 4), 2),       # group 5, small blocks
    (40, 2, 256, ("list", [33, 700]), 0),            # group 20: two head chunks (16 + 4) per kv head
    (16, 8, 256, ("list", [2048]), 8),               # one sequence: the small-batch case the splits exist for
    (2, 1, 256, ("const", 64, 581), 4),              # a TP=8 shard of the bench shape
])
def test_paged_decode_warps_and_splits(ctx, nvc, nh, kv, block_size, lens_spec, splits):
    """both CTA shapes (4 warps x 2 per SM, 2 warps x 4 per SM), with and without KV splits, against the oracle;
    each configuration is launched twice and must be bit-identical between launches (fixed merge order)"""
    import os
    lens = _lens(lens_spec)
    q, kc, vc, bt, ref, scale = _paged_case(nh, kv, block_size, lens, 7 + len(lens))
    B = len(lens)
    args = (q.reshape(B, nh * 128).cuda(), kc.cuda(), vc.cuda(), bt.cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
            nh, kv, 128, block_size, scale)
    if splits:
        os.environ["NVC_PD_SPLITS"] = str(splits)
    try:
        for warps in (2, 4):
            nvc.lib().nvc_debug_set(5, warps)
            o1 = ctx.paged_decode_attn(*args).clone()
            o2 = ctx.paged_decode_attn(*args).clone()
            torch.cuda.synchronize()
            assert not torch.isnan(o1.float()).any(), "poisoned (unwritten) cache slots leaked into the output"
            _close(o1, ref, 1e-2, 2 ** -7, f"paged decode warps={warps} splits={splits} nh={nh} kv={kv} bs={block_size} B={B}")
            assert torch.equal(o1, o2), "paged decode must be bit-identical between launches"
    finally:
        nvc.lib().nvc_debug_set(5, 0)
        os.environ.pop("NVC_PD_SPLITS", None)


