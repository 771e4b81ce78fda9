"""GPU microbenchmark of the prefill attention kernel (K4): TFLOP/s (causal flops 4 * nh * 128 * T^2 / 2 per
sequence) for the BASELINE shapes, both kernel generations, CUDA events around back-to-back launches over
distinct qkv buffers (working set >> L2 where the shape allows)."""
import importlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ops = importlib.import_module("nano-vllm-candle_b200.ops")
lib = importlib.import_module("nano-vllm-candle_b200._lib")
ctx = ops.Context(0)
BF16 = torch.bfloat16
PEAK = 1394.9
if os.path.exists("MEASURED_PEAKS.json"):
    PEAK = json.load(open("MEASURED_PEAKS.json")).get("bf16_tflops_sustained", PEAK)


def bench(name, B, T, nh, kv, iters=10, bufs=4):
    hd = 128
    width = (nh + 2 * kv) * hd
    N = B * T
    qkvs = [(torch.randn(N, width, device="cuda", dtype=torch.float32) * 0.5).to(BF16) for _ in range(bufs)]
    cu = torch.arange(0, (B + 1) * T, T, dtype=torch.int32, device="cuda")
    out = torch.empty(N, nh * hd, dtype=BF16, device="cuda")
    flops = 4.0 * nh * hd * B * T * T / 2
    row = {"shape": name, "B": B, "T": T, "nh": nh, "kv": kv}
    # (label, first-generation kernel)
    variants = [("v2", 0)]
    if os.environ.get("PREFILL_VARIANTS"):
        variants = [v for v in variants if v[0] in os.environ["PREFILL_VARIANTS"].split(",")]
    for label, v1 in variants:
        def run(d):
            ctx.prefill_attn(d[:, :nh * hd], d[:, nh * hd:(nh + kv) * hd], d[:, (nh + kv) * hd:], cu, T, nh, kv, hd, hd ** -0.5, out=out)
        for d in qkvs:
            run(d)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            run(qkvs[i % bufs])
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        row[f"{label}_us"] = round(us, 1)
        row[f"{label}_tflops"] = round(flops / us / 1e6, 1)
        row[f"{label}_frac"] = round(flops / us / 1e6 / PEAK, 3)
    print(json.dumps(row), flush=True)
    return row


SHAPES = [
    ("0.6B cfg2: 64 x 512", 64, 512, 16, 8, 10),
    ("8B: 32 x 512", 32, 512, 32, 8, 10),
    ("8B: 16 x 1024", 16, 1024, 32, 8, 10),
    ("14B/TP2 cfg4: 8 x 2048", 8, 2048, 20, 4, 10),
    ("32B/TP8 cfg5: 8 x 4096", 8, 4096, 8, 1, 10),
    ("8B: 4 x 8192", 4, 8192, 32, 8, 4),
]
only = os.environ.get("PREFILL_SHAPE")          # index into SHAPES (used for the ncu capture of one shape)
todo = [SHAPES[int(only)]] if only is not None else SHAPES
rows = [bench(n, B, T, nh, kv, iters=it) for n, B, T, nh, kv, it in todo]
if len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)
