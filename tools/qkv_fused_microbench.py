"""QKV projection + K2 (q/k-norm + RoPE + KV scatter): one kernel (K2 as the pair GEMM's persistent epilogue) against the
two-kernel sequence, at prefill sizes; CUDA events around back-to-back launches over distinct weight / cache copies."""
import importlib
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

nvc = importlib.import_module("nano-vllm-candle_b200")
ops = importlib.import_module("nano-vllm-candle_b200.ops")
ctx = ops.Context(0)
BF16 = torch.bfloat16


def bench(M, nh, kv, K, fused, iters=10):
    hd = 128
    N = (nh + 2 * kv) * hd
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [(torch.randn(N, K, device="cuda") / math.sqrt(K)).to(BF16) for _ in range(3)]
    qw = torch.ones(hd, device="cuda", dtype=BF16)
    tab = ctx.rope_table(1e6, hd, 4096)
    pos = (torch.arange(M, device="cuda") % 512).to(torch.int32)
    slots = torch.arange(M, device="cuda", dtype=torch.int32)
    kc = torch.zeros(M, kv, hd, dtype=BF16, device="cuda")
    vc = torch.zeros_like(kc)
    out = torch.empty(M, N, dtype=BF16, device="cuda")
    nvc.lib().nvc_debug_set(7, 0 if fused else 1)
    run = lambda i: ctx.qkv_proj_norm_rope(a, ws[i % 3], qw, qw, 1e-6, tab, pos, nh, kv, hd, slot_mapping=slots, k_cache=kc,
                                           v_cache=vc, out=out)
    for i in range(3):
        run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(i)
    e1.record()
    torch.cuda.synchronize()
    nvc.lib().nvc_debug_set(7, 0)
    return e0.elapsed_time(e1) * 1e3 / iters


for name, M, nh, kv, K in [("0.6B, 32768 tokens", 32768, 16, 8, 1024), ("8B, 16384 tokens", 16384, 32, 8, 4096),
                           ("14B / TP 2, 16384 tokens", 16384, 20, 4, 5120), ("0.6B, 4096 tokens", 4096, 16, 8, 1024)]:
    u = bench(M, nh, kv, K, False)
    f = bench(M, nh, kv, K, True)
    print(f"{name:28s} GEMM + K2 {u:8.1f} us   fused {f:8.1f} us   ({u / f:.3f} x)", flush=True)
