"""CTA-0 phase traces (clock64 stamps) and cuBLAS comparison of the four Qwen3-8B layer GEMMs at batch 256 through the
pair kernel's own schedule choice (force_splits = 0) and forced split factors."""
import ctypes as C
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.argv = sys.argv[:1]
import gemm_microbench as gm
import torch

BF16 = torch.bfloat16
ctx = gm.ctx


def raw_trace(M, N, K, splits):
    a = torch.randn(M, K, device="cuda").to(BF16)
    ws = [torch.randn(N, K, device="cuda").to(BF16) for _ in range(max(2, int(300e6 // (N * K * 2)) + 1))]
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    t = torch.zeros(128, dtype=torch.int64, device="cuda")
    for w in ws:
        st = ctx.l.nvc_gemm_bf16_trace(ctx.h, C.c_void_p(a.data_ptr()), K, C.c_void_p(w.data_ptr()), C.c_void_p(out.data_ptr()),
                                       N, 0, M, N, K, 512, splits, C.c_void_p(torch.cuda.current_stream().cuda_stream),
                                       C.c_void_p(t.data_ptr()))
        assert st == 0
    torch.cuda.synchronize()
    t = t.cpu().tolist()
    t0 = t[0]
    rel = lambda i: (t[i] - t0) if t[i] else None
    tma = [rel(3 + i) for i in range(32) if t[3 + i]]
    full = [rel(40 + i) for i in range(32) if t[40 + i]]
    print(f"M={M} N={N} K={K} splits={splits}: setup {rel(1)} pdl {rel(2)} tma0 {tma[:3]}..{tma[-1:]} n={len(tma)} "
          f"full0 {full[:3]}..{full[-1:]} mma_done {rel(80)} epi_start {rel(81)} parked {rel(82)} barrier {rel(85)} "
          f"epi_end {rel(83)} end {rel(84)}", flush=True)


for (M, N, K) in [(256, 6144, 4096), (256, 4096, 4096), (256, 24576, 4096), (256, 4096, 12288)]:
    us, tf = gm.bench_cublas(M, N, K)
    print(f"{M} {N} {K} cuBLAS {us:.2f} us {tf:.0f} TF", flush=True)
    for sp in (0, 1, 2, 3, 4):
        try:
            us, gbs, tf = gm.bench(M, N, K, 512, sp)
            print(f"{M} {N} {K} pair splits={sp}: {us:.2f} us {tf:.0f} TF", flush=True)
        except Exception as e:
            print("failed", sp, e)
    for sp in (0, 2, 3):
        raw_trace(M, N, K, sp)
