"""One swapped-operand lm_head launch per shape (for an ncu capture): fp32 logits = [64, K] x [151936, K]^T, K = 1024 and 4096."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ops = importlib.import_module("nano-vllm-candle_b200.ops")
ctx = ops.Context(0)
V = 151936
for K in (1024, 4096):
    a = torch.randn(64, K, device="cuda").to(torch.bfloat16)
    w = torch.randn(V, K, device="cuda").to(torch.bfloat16)
    out = torch.empty(64, V, device="cuda", dtype=torch.float32)
    for _ in range(2):
        ctx.gemm(a, w, out=out)
    torch.cuda.synchronize()
    print("K", K, "ok", bool(torch.isfinite(out).all()))
