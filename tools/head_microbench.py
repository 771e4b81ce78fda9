"""Times the two token-choice kernels (K8 argmax, K8s Gumbel-max sampling) at decode batch sizes.  Usage: python tools/head_microbench.py"""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ops = importlib.import_module("nano-vllm-candle_b200.ops")


def timed(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n


def main():
    ctx = ops.Context(0)
    V = 151936
    for B in (1, 16, 64, 128, 256):
        bufs = [torch.randn(B, V, device="cuda") for _ in range(max(2, 400_000_000 // (B * V * 4)))]   # rotate past L2
        temps = torch.full((B,), 0.7, device="cuda")
        streams = torch.arange(B, dtype=torch.int64, device="cuda")
        i = [0]

        def am():
            i[0] += 1
            ctx.argmax_last(bufs[i[0] % len(bufs)])

        def sm():
            i[0] += 1
            ctx.sample_gumbel(bufs[i[0] % len(bufs)], temps, streams, 7)
        ta, ts = timed(am), timed(sm)
        print(f"B={B:4d} V={V}: argmax {ta:6.2f} us ({B * V * 4 / ta / 1e3:6.0f} GB/s)   sample_gumbel {ts:6.2f} us", flush=True)


if __name__ == "__main__":
    main()
