"""GPU microbenchmark of the tcgen05 GEMM: time vs shape / tile / split-K with cold weights
(rotating over enough weight copies to exceed L2), plus phase timestamps of one CTA."""
import ctypes as C
import importlib
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ops = importlib.import_module("nano-vllm-candle_b200.ops")
ctx = ops.Context(0)
BF16 = torch.bfloat16


def bench(M, N, K, bn=0, splits=0, out_dtype=BF16, iters=20, check=False):
    """GPU time per GEMM: `iters` launches over rotating (cold) weight copies captured in ONE CUDA
    graph, so host launch cost (python + 2 tensor-map encodes, ~11 us) is not what is measured."""
    ncopies = max(2, min(64, int(300e6 // (N * K * 2)) + 1))
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [torch.randn(N, K, device="cuda").to(BF16) for _ in range(ncopies)]
    out = torch.empty(M, N, device="cuda", dtype=out_dtype)
    if check:
        want = ctx.gemm(a, ws[0], force_bn=bn, force_splits=splits).clone()
    for w in ws[:2]:
        ctx.gemm(a, w, out=out, force_bn=bn, force_splits=splits)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            ctx.gemm(a, ws[i % ncopies], out=out, force_bn=bn, force_splits=splits)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    gbs = (N * K * 2 + M * K * 2 + M * N * out.element_size()) / us / 1e3
    tf = 2.0 * M * N * K / us / 1e6
    return us, gbs, tf


def bench_cublas(M, N, K, iters=20):
    """torch.matmul (cuBLAS) on the same shapes with the same cold-weight rotation, for reference."""
    ncopies = max(2, min(64, int(300e6 // (N * K * 2)) + 1))
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [torch.randn(N, K, device="cuda").to(BF16) for _ in range(ncopies)]
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    for w in ws[:2]:
        torch.matmul(a, w.t(), out=out)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            torch.matmul(a, ws[i % ncopies].t(), out=out)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    return us, 2.0 * M * N * K / us / 1e6


def trace(M, N, K, bn, splits, cold=False):
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [torch.randn(N, K, device="cuda").to(BF16) for _ in range(max(2, int(300e6 // (N * K * 2)) + 1) if cold else 1)]
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    t = torch.zeros(128, dtype=torch.int64, device="cuda")
    for w in (ws if cold else ws * 2):     # cold: the last launch reads weights that left L2 long ago
        st = ctx.l.nvc_gemm_bf16_trace(ctx.h, C.c_void_p(a.data_ptr()), K, C.c_void_p(w.data_ptr()), C.c_void_p(out.data_ptr()),
                                       N, 0, M, N, K, bn, splits, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                                       C.c_void_p(t.data_ptr()))
        assert st == 0
    torch.cuda.synchronize()
    t = t.cpu().tolist()
    t0 = t[0]
    kps = max((bn // 1000) % 10, 1)
    kb = -(-(K // 64 // max(splits % 100, 1)) // kps)
    print(f"trace M={M} N={N} K={K} bn={bn} splits={splits} cold={cold} (cycles from kernel start)")
    print("  setup done        ", t[1] - t0)
    print("  tma issued  kb:   ", [t[2 + i] - t0 for i in range(min(kb, 32))])
    print("  full waited kb:   ", [t[40 + i] - t0 for i in range(min(kb, 32))])
    print("  mma all committed ", t[80] - t0)
    print("  epilogue start    ", t[81] - t0)
    print("  epilogue end      ", t[82] - t0)
    print("  kernel end        ", t[83] - t0)


def trace_pair(M, N, K, splits):
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [torch.randn(N, K, device="cuda").to(BF16) for _ in range(max(2, int(300e6 // (N * K * 2)) + 1))]
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    t = torch.zeros(128, dtype=torch.int64, device="cuda")
    for w in ws:       # the last launch reads cold weights
        st = ctx.l.nvc_gemm_bf16_trace(ctx.h, C.c_void_p(a.data_ptr()), K, C.c_void_p(w.data_ptr()), C.c_void_p(out.data_ptr()),
                                       N, 0, M, N, K, 512, splits, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                                       C.c_void_p(t.data_ptr()))
        assert st == 0
    torch.cuda.synchronize()
    t = t.cpu().tolist()
    t0 = t[0]
    S = splits if 1 <= splits <= 4 else (max(1, splits // 10000) if splits >= 1000 else 1)
    kb = min(32, K // 64 // S)
    print(f"pair trace M={M} N={N} K={K} splits={splits} (cycles from kernel start, CTA 0)")
    print("  cluster set up    ", t[1] - t0, " pdl wait done", t[2] - t0)
    print("  tma issued  kb:   ", [t[3 + i] - t0 for i in range(kb)])
    print("  full waited kb:   ", [t[40 + i] - t0 for i in range(kb)])
    print("  mma all committed ", t[80] - t0)
    print("  epilogue start    ", t[81] - t0)
    if S > 1:
        print("  partials parked   ", t[82] - t0, " cluster barrier passed", t[85] - t0)
    print("  epilogue end      ", t[83] - t0)
    if S == 1:
        print("  last tile epilogue", t[86] - t0, "->", t[87] - t0)
    print("  kernel end        ", t[84] - t0)


if __name__ == "__main__":
    print("M N K bn splits -> us, GB/s, TFLOP/s")
    if "--single" in sys.argv:      # one shape, a few plain launches (for ncu)
        i = sys.argv.index("--single")
        M, N, K, bn = [int(x) for x in sys.argv[i + 1:i + 5]]
        sp = int(sys.argv[i + 5]) if len(sys.argv) > i + 5 else 1
        a = torch.randn(M, K, device="cuda").to(BF16)
        ws = [torch.randn(N, K, device="cuda").to(BF16) for _ in range(6)]
        out = torch.empty(M, N, device="cuda", dtype=BF16)
        for w in ws:
            ctx.gemm(a, w, out=out, force_bn=bn, force_splits=sp)
        torch.cuda.synchronize()
        print("single ok")
        sys.exit(0)
    if "--trace-pair" in sys.argv:
        for sp in (1256, 1192, 31192, 21192, 21128):
            trace_pair(256, 24576 if sp < 10000 else 4096, 4096, sp)
        sys.exit(0)
    if "--narrow" in sys.argv:      # narrow pair tiles x in-cluster split-K on the decode (M = 256) shapes
        shapes = [(256, 6144, 4096), (256, 4096, 4096), (256, 24576, 4096), (256, 4096, 12288), (256, 151936, 4096),
                  (256, 5120, 5120), (256, 7168, 5120), (256, 34816, 5120), (256, 5120, 17408),
                  (128 + 64, 24576, 4096), (512, 24576, 4096)]
        for (M, N, K) in shapes:
            us, tf = bench_cublas(M, N, K)
            print(f"{M:6d} {N:6d} {K:6d} cuBLAS        -> {us:9.2f} us {tf:8.1f} TF/s", flush=True)
            us, gbs, tf = bench(M, N, K, 512, 0)
            print(f"{M:6d} {N:6d} {K:6d} auto          -> {us:9.2f} us {tf:8.1f} TF/s", flush=True)
            res = []
            for nt in (128, 160, 192, 224, 256):
                for S in (1, 2, 3, 4):
                    tiles = -(-N // nt) * -(-M // 256)
                    if (S > 1 and tiles > (0, 74, 33, 22, 15)[S] * 2) or (nt > 256 and S > 1):
                        continue
                    try:
                        us, gbs, tf = bench(M, N, K, 512, 10000 * S + 1000 + nt)
                        res.append((us, nt, S))
                        print(f"{M:6d} {N:6d} {K:6d} nt={nt:3d} S={S} tiles={tiles:4d} -> {us:9.2f} us {tf:8.1f} TF/s {gbs:8.1f} GB/s", flush=True)
                    except Exception as ex:
                        print(f"{M} {N} {K} nt={nt} S={S} failed: {ex}")
            res.sort()
            print(f"{M:6d} {N:6d} {K:6d} best: " + ", ".join(f"nt={n} S={s} {u:.2f}us" for u, n, s in res[:4]), flush=True)
        sys.exit(0)
    if "--pair" in sys.argv:        # CTA-pair kernel vs the best 1-CTA tiles vs cuBLAS, M > 128
        shapes = [(256, 6144, 4096), (256, 4096, 4096), (256, 24576, 4096), (256, 4096, 12288), (256, 151936, 4096),
                  (256, 4096, 1024), (256, 1024, 2048), (256, 6144, 1024), (256, 1024, 3072),
                  (192, 4096, 4096), (512, 4096, 4096),
                  (4096, 4096, 4096), (8192, 8192, 8192), (32768, 4096, 1024), (32768, 1024, 2048), (32768, 6144, 1024),
                  (32768, 1024, 3072), (16384, 6144, 4096), (16384, 4096, 12288)]
        for (M, N, K) in shapes:
            us, tf = bench_cublas(M, N, K)
            print(f"{M:6d} {N:6d} {K:6d} cuBLAS                 -> {us:9.2f} us {tf:8.1f} TF/s", flush=True)
            for bn, sp in [(2064, 1), (2128, 1), (1256, 1), (512, 0), (512, 1), (512, 2), (512, 3), (512, 4)]:
                if (M > 512 and (bn == 2064 or sp > 1)) or (sp > 1000 and -(-N // (sp - 1000)) * -(-M // 256) > 74):
                    continue
                try:
                    us, gbs, tf = bench(M, N, K, bn, sp)
                    print(f"{M:6d} {N:6d} {K:6d} bn={bn % 1000:3d} kps={(bn // 1000) % 10} s={sp}        -> {us:9.2f} us {tf:8.1f} TF/s {gbs:8.1f} GB/s", flush=True)
                except Exception as ex:
                    print(f"{M} {N} {K} bn={bn} s={sp} failed: {ex}")
        for sp in (1, 2, 3):
            trace_pair(256, 4096, 4096, sp)
        trace_pair(256, 24576, 4096, 1)
        sys.exit(0)
    if "--trace-decode" in sys.argv:
        for cold in (False, True):
            trace(64, 4096, 4096, 4032, 1, cold)
            trace(64, 4096, 4096, 2128, 104, cold)
            trace(64, 4096, 12288, 2128, 104, cold)
            trace(64, 6144, 4096, 2064, 1, cold)
        sys.exit(0)
    if "--ks" in sys.argv:          # cluster split-K (DSMEM) for the narrow decode GEMMs vs the plain tiles vs cuBLAS
        for (M, N, K) in [(64, 1024, 2048), (64, 1024, 3072), (64, 4096, 1024), (64, 6144, 1024), (1, 1024, 2048),
                          (64, 4096, 4096), (64, 4096, 12288), (64, 6144, 4096), (64, 24576, 4096), (64, 5120, 5120),
                          (128, 1024, 3072), (128, 4096, 4096), (1, 4096, 4096), (1, 4096, 12288), (16, 6144, 4096),
                          (128, 7168, 5120), (128, 5120, 17408), (64, 151936, 4096)]:
            us, tf = bench_cublas(M, N, K)
            print(f"{M:6d} {N:6d} {K:6d} cuBLAS                 -> {us:9.2f} us {tf:8.1f} TF/s", flush=True)
            for bn, sp in [(0, 0), (4032, 1), (2064, 1), (2128, 1), (1256, 1), (4032, 102), (4032, 104), (2064, 102), (2064, 104),
                           (2128, 102), (2128, 104), (1256, 102), (1256, 104)]:
                if N > 100000 and (sp > 1 or bn == 4032):
                    continue
                try:
                    us, gbs, tf = bench(M, N, K, bn, sp)
                    print(f"{M:6d} {N:6d} {K:6d} bn={bn % 1000:3d} kps={(bn // 1000) % 10} s={sp:3d}      -> {us:9.2f} us {tf:8.1f} TF/s {gbs:8.1f} GB/s", flush=True)
                except Exception as ex:
                    print(f"{M} {N} {K} bn={bn} s={sp} failed: {ex}")
        sys.exit(0)
    if "--trace-only" in sys.argv:
        trace(64, 4096, 1024, 4032, 1)
        trace(64, 4096, 1024, 2128, 4)
        trace(4096, 4096, 4096, 1256, 1)
        sys.exit(0)
    small = [(0, 0), (4032, 1), (2064, 1), (4064, 1), (2128, 1), (1256, 1),
             (84016, 1), (44016, 1), (44032, 1), (84032, 1), (44064, 1), (42064, 1), (42128, 1), (21256, 1)]
    mid = [(0, 0), (2064, 1), (1128, 1), (2128, 1), (1256, 1), (42064, 1), (44064, 1), (42096, 1), (42128, 1), (41128, 1), (21256, 1)]
    big = [(0, 0), (1128, 1), (1256, 1)]
    for (M, N, K) in [(64, 4096, 1024), (64, 1024, 2048), (64, 6144, 1024), (64, 1024, 3072), (64, 151936, 1024),
                      (1, 4096, 1024), (16, 1024, 3072),
                      (256, 6144, 4096), (256, 4096, 4096), (256, 24576, 4096), (256, 4096, 12288),
                      (4096, 4096, 4096), (8192, 8192, 8192), (32768, 4096, 1024), (32768, 1024, 3072)]:
        combos = small if M <= 128 else (mid if M <= 256 else big)
        for bn, sp in combos:
            if N > 100000 and sp > 1:
                continue
            cl, b_ = max(bn // 10000, 1), bn % 1000
            if b_ and (-(-N // b_)) % cl:
                continue
            try:
                us, gbs, tf = bench(M, N, K, bn, sp)
                print(f"{M:6d} {N:6d} {K:6d} bn={bn % 1000:3d} kps={(bn // 1000) % 10} cl={max(bn // 10000, 1)} s={sp} -> {us:9.2f} us {gbs:8.1f} GB/s {tf:8.1f} TF/s", flush=True)
            except Exception as ex:
                print(f"{M} {N} {K} bn={bn} s={sp} failed: {ex}")
    trace(64, 4096, 1024, 4032, 1)
    trace(64, 4096, 1024, 44032, 1)
    trace(64, 1024, 2048, 84016, 1)
    trace(256, 6144, 4096, 42096, 1)
