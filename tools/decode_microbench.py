"""GPU microbenchmark of the paged decode attention kernel: achieved HBM GB/s for the BASELINE
shapes, timed with CUDA graphs over several distinct KV pools (working set >> L2)."""
import importlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ops = importlib.import_module("nano-vllm-candle_b200.ops")
ctx = ops.Context(0)
BF16 = torch.bfloat16
PEAK = 6545.6
if os.path.exists("MEASURED_PEAKS.json"):
    PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]


def bench(B, nh, kv, ctx_len, ragged=False, block_size=256, pools=None):
    hd = 128
    g = torch.Generator().manual_seed(0)
    lens = torch.randint(max(1, ctx_len // 8), ctx_len + 1, (B,), generator=g) if ragged else torch.full((B,), ctx_len)
    nblk = [(int(l) + block_size - 1) // block_size for l in lens]
    total = sum(nblk)
    bytes_per_pool = total * block_size * kv * hd * 2 * 2
    pools = pools or max(2, min(28, int(6e9 // bytes_per_pool)))
    bt = torch.full((B, max(nblk)), -1, dtype=torch.int32)
    perm = torch.randperm(total, generator=g)
    p = 0
    for b in range(B):
        bt[b, :nblk[b]] = perm[p:p + nblk[b]].to(torch.int32)
        p += nblk[b]
    kcs = [torch.randn(total * block_size, kv, hd, device="cuda", dtype=torch.float16).to(BF16) for _ in range(pools)]
    vcs = [torch.randn(total * block_size, kv, hd, device="cuda", dtype=torch.float16).to(BF16) for _ in range(pools)]
    q = torch.randn(B, nh * hd, device="cuda").to(BF16)
    out = torch.empty(B, nh * hd, device="cuda", dtype=BF16)
    btd, ld = bt.cuda(), lens.to(torch.int32).cuda()
    scale = hd ** -0.5
    call = lambda i: ctx.paged_decode_attn(q, kcs[i], vcs[i], btd, ld, nh, kv, hd, block_size, scale, out=out)
    for i in range(2):
        call(i % pools)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for i in range(pools):
            call(i)
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (3 * pools)
    alg = int(lens.sum()) * 2 * kv * hd * 2 + 2 * B * nh * hd * 2
    return us, alg / us / 1e3


def sweep_splits():
    """--splits: warps per CTA (4 warps x 2 CTAs per SM | 2 warps x 4 CTAs per SM) x KV-split count on the bench shape
    and the sharded / small-batch shapes, where the per-(sequence, kv head) grid alone cannot fill 148 SMs."""
    print(f"paged decode attention warps x split sweep, peak {PEAK} GB/s")
    for name, B, nh, kv, c in [("0.6B tp1 bs64 ctx581", 64, 16, 8, 581), ("0.6B tp2 bs64 ctx581", 64, 8, 4, 581),
                               ("0.6B tp4 bs64 ctx581", 64, 4, 2, 581), ("0.6B tp8 bs64 ctx581", 64, 2, 1, 581),
                               ("0.6B bs32 ctx581", 32, 16, 8, 581), ("0.6B bs48 ctx581", 48, 16, 8, 581),
                               ("0.6B bs128 ctx581", 128, 16, 8, 581), ("8B bs64 ctx1024", 64, 32, 8, 1024),
                               ("8B tp8 bs256 ctx1024", 256, 4, 1, 1024), ("8B tp4 bs256 ctx1024", 256, 8, 2, 1024),
                               ("8B bs256 ctx1024", 256, 32, 8, 1024),
                               ("0.6B bs1 ctx2048", 1, 16, 8, 2048), ("8B bs1 ctx1024", 1, 32, 8, 1024), ("8B bs16 ctx1024", 16, 32, 8, 1024)]:
        row = []
        for warps in (0, 4, 2):
            if warps:
                os.environ["NVC_PD_WARPS"] = str(warps)
            else:
                os.environ.pop("NVC_PD_WARPS", None)
            for s in ((0,) if not warps else (1, 2, 3, 4, 6, 8, 16)):
                if s:
                    os.environ["NVC_PD_SPLITS"] = str(s)
                else:
                    os.environ.pop("NVC_PD_SPLITS", None)
                us, gbs = bench(B, nh, kv, c, pools=8)
                row.append(f"w{warps or 'auto'} S={s or 'auto'}: {us:.2f}")
        os.environ.pop("NVC_PD_SPLITS", None)
        os.environ.pop("NVC_PD_WARPS", None)
        print(f"{name:24s} " + " | ".join(row), flush=True)


if __name__ == "__main__":
    if "--splits" in sys.argv:
        sweep_splits()
        sys.exit(0)
    print(f"paged decode attention, peak {PEAK} GB/s; env NVC_PD_SPLITS={os.environ.get('NVC_PD_SPLITS')} NVC_PD_WARPS={os.environ.get('NVC_PD_WARPS')}")
    for name, B, nh, kv, c, rag in [("0.6B bs64 ctx581", 64, 16, 8, 581, False), ("0.6B bs64 ctx1024", 64, 16, 8, 1024, False),
                                    ("0.6B bs1 ctx2048", 1, 16, 8, 2048, False), ("0.6B bs256 ctx640", 256, 16, 8, 640, False),
                                    ("8B bs256 ctx512", 256, 32, 8, 512, False), ("8B bs256 ctx1024", 256, 32, 8, 1024, False),
                                    ("8B bs256 ctx2048", 256, 32, 8, 2048, False), ("8B bs256 ragged[256,2048]", 256, 32, 8, 2048, True),
                                    ("8B bs64 ctx1024", 64, 32, 8, 1024, False), ("8B tp8 bs256 ctx1024", 256, 4, 1, 1024, False),
                                    ("14B bs128 ctx2048", 128, 40, 8, 2048, False), ("32B tp8 bs256 ctx4096", 256, 8, 1, 4096, False)]:
        us, gbs = bench(B, nh, kv, c, rag)
        print(f"{name:28s} {us:9.2f} us  {gbs:8.1f} GB/s  {gbs / PEAK:.3f} of measured peak", flush=True)
