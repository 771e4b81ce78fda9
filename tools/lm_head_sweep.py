"""lm_head GEMM at decode batch sizes ([M, H] x [V, H]^T -> fp32 logits): forced (BN, KPS) tiles of the 1-CTA
kernel against the automatic choice; cold weights (two copies > L2), CUDA-graph timed."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ops = importlib.import_module("nano-vllm-candle_b200.ops")
ctx = ops.Context(0)
BF16 = torch.bfloat16
V = int(sys.argv[1]) if len(sys.argv) > 1 else 151936


def timed(M, K, force_bn, iters=8, force_splits=0):
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [torch.randn(V, K, device="cuda").to(BF16) for _ in range(2)]
    out = torch.empty(M, V, device="cuda", dtype=torch.float32)
    try:
        for w in ws:
            ctx.gemm(a, w, out=out, force_bn=force_bn, force_splits=force_splits)
    except Exception as e:          # a (BN, KPS) pair that is not instantiated
        return None
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            ctx.gemm(a, ws[i % 2], out=out, force_bn=force_bn, force_splits=force_splits)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


for M, K in ((64, 1024), (1, 1024), (16, 1024), (128, 1024), (64, 4096), (128, 5120)):
    gb = (V * K * 2 + M * V * 4) / 1e3
    row = []
    for fb in (0, 1256, 21256):
        us = timed(M, K, fb)
        row.append(f"{fb}:{'-' if us is None else f'{us:.1f}us/{gb / us:.0f}GB/s'}")
    # the CTA-pair kernel (persistent, two accumulator stages) on the same decode shape: force_bn 512, tile width NT
    for nt in (256, 192):
        us = timed(M, K, 512, force_splits=1000 + nt)
        row.append(f"pair{nt}:{'-' if us is None else f'{us:.1f}us/{gb / us:.0f}GB/s'}")
    # operands swapped in the pair kernel (weights as the 256-row MMA operand, transposed fp32 store): force_bn 513
    us = timed(M, K, 513)
    row.append(f"swapped:{'-' if us is None else f'{us:.1f}us/{gb / us:.0f}GB/s'}")
    print(f"V={V} M={M} K={K}  " + "  ".join(row), flush=True)
