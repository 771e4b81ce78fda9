"""Parity + latency of the NVLS (NVSwitch multicast) all-reduce kernels, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29513 tools/nvls_check.py --out gpurun_out/nvls_N.jsonl

For every (rows, hidden, mode): each rank holds its own partial sums; the kernel's output (sum over ranks
+ residual, RMSNorm) is compared with a torch fp32 evaluation of the same expression on the all-gathered
partials (test driver: torch is the checker here, not the product), and the outputs of all ranks must be
bit-identical.  Then the same call is timed back to back from a CUDA graph (28 calls per replay, max over
ranks), next to the peer-load kernel of round 1 and NCCL all-reduce + the norm kernel."""
import argparse
import ctypes as C
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

MODES = {1: "ll_oneshot", 2: "twoshot_flags", 3: "twoshot_ll"}


def expected(parts, res, w, eps, bf16_sum):
    s = torch.zeros_like(parts[0], dtype=torch.float32)
    for p in parts:                                   # rank order
        s = s + p.float()
    if bf16_sum:                                      # two-shot: the switch returns the sum rounded to bf16
        s = s.bfloat16().float()
    if w is None:
        return s, None
    s = s + res
    o = s * torch.rsqrt((s * s).mean(-1, keepdim=True) + eps) * w.float()
    return o, s


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-timing", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ops = importlib.import_module("nano-vllm-candle_b200.ops")
    ctx = ops.Context(local, rank, world)
    ctx.init_tp(dist, slot_bytes=8 << 20)
    slot_bytes, ll_bytes = 64 << 20, 4 << 20
    ok_nvls = ctx.nvls_init(slot_bytes, ll_bytes)
    lines = [{"what": "nvls_init", "world": world, "ok": ok_nvls, "status": ctx.nvls_status()}]
    if rank == 0:
        print(json.dumps(lines[0]), flush=True)
    fails = 0
    eps = 1e-6
    if ok_nvls:
        # ---- parity.  The order matters: the row count changes between calls (LL buffers are refilled by extent)
        shapes = [(64, 1024), (1, 1024), (16, 1024), (256, 1024), (64, 1024), (37, 5120), (256, 4096), (8, 8192), (512, 4096),
                  (3, 2560), (64, 1024), (2048, 1024), (4096, 4096)]
        gen = torch.Generator(device=dev)
        call = 0
        for (N, H) in shapes:
            msg = N * H * 2
            modes = [m for m in (1, 2, 3) if (m != 1 or msg * world <= ll_bytes) and (m != 3 or msg <= ll_bytes) and msg <= slot_bytes]
            modes.append(0)
            for mode in modes:
                for with_norm in (True, False):
                    call += 1
                    gen.manual_seed(1000 * call + rank)
                    x = (torch.randn((N, H), device=dev, generator=gen) * 0.5).bfloat16()
                    if call % 3 == 0:
                        x.view(torch.int16)[0, :8] = -1          # 0xFFFF halves: the scrubbed pattern
                    gen.manual_seed(77 * call)
                    res = torch.randn((N, H), device=dev, generator=gen)
                    w = (1 + 0.1 * torch.randn((H,), device=dev, generator=gen)).bfloat16() if with_norm else None
                    slot = ctx.nvls_slot(N, H)
                    slot.copy_(x)
                    parts = [torch.empty_like(x) for _ in range(world)]
                    dist.all_gather(parts, x)
                    res_out = torch.empty_like(res) if with_norm else None
                    torch.cuda.synchronize()
                    dist.barrier()
                    out = ctx.nvls_allreduce_rmsnorm(slot, w, eps, res if with_norm else None, res_out, mode=mode)
                    torch.cuda.synchronize()
                    if call % 3 == 0:
                        for p in parts:
                            p.view(torch.int16)[0, :8] = 0x7FFF  # what the sender makes of it (still a NaN)
                    picked = mode if mode else (1 if msg * world <= ll_bytes and msg * (world - 1) <= (3 << 20) else (3 if msg <= ll_bytes else 2))
                    eo, es = expected(parts, res, w, eps, bf16_sum=picked != 1)
                    fin = torch.isfinite(eo)
                    err = float(((out.float() - eo)[fin]).abs().max()) if bool(fin.any()) else 0.0
                    scale = float(eo[fin].abs().max()) if bool(fin.any()) else 1.0
                    tol = 2e-2 * max(scale, 1.0) if picked != 1 else 1e-2 * max(scale, 1.0)
                    res_err = float(((res_out - es)[fin]).abs().max()) if with_norm and bool(fin.any()) else 0.0
                    nan_ok = bool((torch.isfinite(out.float()) == fin).all())
                    outs = [torch.empty_like(out) for _ in range(world)]
                    dist.all_gather(outs, out)
                    same = all(torch.equal(o.view(torch.int16), outs[0].view(torch.int16)) for o in outs)
                    good = err <= tol and res_err <= 2e-2 * max(scale, 1.0) and same and nan_ok
                    fails += int(not good)
                    rec = {"what": "parity", "N": N, "H": H, "mode": mode, "picked": MODES[picked], "norm": with_norm,
                           "max_err": err, "res_err": res_err, "scale": scale, "ranks_bit_identical": same, "ok": good}
                    lines.append(rec)
                    if rank == 0 and not good:
                        print(json.dumps(rec), flush=True)
    # ---- timing
    if not a.no_timing:
        tshapes = [(1, 1024), (64, 1024), (256, 1024), (64, 4096), (256, 4096), (256, 5120), (256, 8192), (1024, 4096), (4096, 4096),
                   (32768, 1024)]
        L = ctx.l
        for (N, H) in tshapes:
            msg = N * H * 2
            res = torch.zeros((N, H), device=dev)
            w = torch.ones((H,), device=dev, dtype=torch.bfloat16)
            out = torch.empty((N, H), device=dev, dtype=torch.bfloat16)
            variants = []
            if ok_nvls:
                slot = ctx.nvls_slot(N, H)
                slot.zero_()
                for m in (1, 2, 3):
                    if (m == 1 and msg * world > ll_bytes) or (m == 3 and msg > ll_bytes):
                        continue
                    variants.append((MODES[m], lambda m=m, slot=slot: ctx.nvls_allreduce_rmsnorm(slot, w, eps, res, res, out=out, mode=m)))
            if ctx.l.nvc_ctx_p2p_ready(ctx.h) and msg <= (8 << 20):
                def p2p(k=[0]):
                    k[0] ^= 1
                    ctx._ck(L.nvc_p2p_allreduce_rmsnorm(ctx.h, k[0], ops._ptr(res), ops._ptr(res), ops._ptr(w), eps, ops._ptr(out), N, H,
                                                        ops._stream()), "p2p")
                variants.append(("p2p_oneshot_r01", p2p))
            xb = torch.zeros((N, H), device=dev, dtype=torch.bfloat16)

            def nccl():
                ctx.allreduce(xb)
                ctx.fused_add_rmsnorm(xb, w, eps, res, res, out=out)
            variants.append(("nccl_plus_norm", nccl))
            for name, fn in variants:
                reps = 28 if msg <= (8 << 20) else 4
                s = torch.cuda.Stream()
                with torch.cuda.stream(s):
                    fn()
                    torch.cuda.synchronize()
                    dist.barrier()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=s):
                        for _ in range(reps):
                            fn()
                    g.replay()
                    torch.cuda.synchronize()
                    dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    iters = 10 if msg <= (8 << 20) else 3
                    e0.record()
                    for _ in range(iters):
                        g.replay()
                    e1.record()
                    torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / (iters * reps)
                t = torch.tensor([us], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                rec = {"what": "timing", "N": N, "H": H, "bytes": msg, "variant": name, "us_per_call": round(float(t[0]), 2),
                       "world": world}
                lines.append(rec)
                if rank == 0:
                    print(json.dumps(rec), flush=True)
                del g
    summary = {"what": "summary", "world": world, "nvls": ok_nvls, "status": ctx.nvls_status(),
               "parity_cases": sum(1 for l in lines if l["what"] == "parity"), "parity_failures": fails}
    lines.append(summary)
    if rank == 0:
        print(json.dumps(summary), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                for l in lines:
                    f.write(json.dumps(l) + "\n")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
