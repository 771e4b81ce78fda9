#!/usr/bin/env python
"""Benchmark of the Qwen3 forward hot path (BASELINE.json metric: decode tok/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload (N = 1): BASELINE.json configs[1] - Qwen3-0.6B bf16, batch 64, 512-token prompts,
decode with ignore_eos, paged KV cache block size 256, random-init (synthetic seeded) weights.
A "step" is ONE decode step of the whole batch.  N > 1 runs the same batch tensor-parallel
(TP = N, one process per GPU, NCCL all-reduce after each row-parallel linear): strong scaling.

One JSON line on stdout (rank 0):
  value     decode tok/s with the step's inputs already resident in HBM (CUDA events on the
            runner's stream around K replays of the step, max over ranks)
  e2e       decode tok/s through LLMEngine.step(): host scheduler -> pinned H2D of ids / positions /
            slots / context lens / block tables -> step -> D2H of the sampled tokens, K timed steps
  roofline  the dominant kernel (paged decode attention) against the measured HBM peak
  cpu_baseline  the oracle's restatement of the reference algorithm on the host cores (bounded sample)

--impl reference times that CPU restatement alone (the reference crate is Rust + candle and cannot
be built on this image).
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODEL = "qwen3-0.6b"
BATCH = 64
PROMPT_LEN = 512
BLOCK_SIZE = 256
CPU_SAMPLE_SEQS = 1          # sequences of the batch the CPU legs process per step


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the paged decode kernel on this
    workload, from the committed `ncu --set full` capture (profiles/paged_decode_ncu_traffic.json,
    written by tools/ncu_summarize.py); None if no capture has been committed."""
    p = os.path.join(ROOT, "profiles", "paged_decode_ncu_traffic.json")
    try:
        return float(json.load(open(p))["dram_bytes_per_launch"])
    except Exception:
        return None


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True).start()
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# CPU restatement of the reference (oracle/) - the ONLY place bench.py touches oracle/
# --------------------------------------------------------------------------------------------
def cpu_reference_decode(steps: int, warmup: int, sample_seqs: int):
    """The reference's algorithm for one decode step (src/engine/llm_engine.rs:146-189): right-padded
    [B, T] batch, FULL forward of every sequence, logits at all positions, host argmax; fp32, all host
    threads.  Returns (tok/s, threads, seconds per step)."""
    import torch
    from oracle import qwen3_oracle as O
    from oracle.engine_oracle import OracleModelRunner, SamplingParams, Sequence
    synth = importlib.import_module("nano-vllm-candle_b200.synth")
    cfg = O.MODEL_CONFIGS[MODEL]
    w = {k: v.float() for k, v in synth.iter_weights(cfg, 0)}
    model = O.Qwen3Oracle(cfg, w, "fp32")
    runner = OracleModelRunner(model, cfg.eos_token_id)
    prompts = synth.make_prompts(BATCH, PROMPT_LEN, cfg.vocab_size, 2)[:sample_seqs]
    seqs = [Sequence(p + [p[-1]], SamplingParams(1.0, 10 ** 6, True)) for p in prompts]   # ctx 513 = first decode step
    with torch.no_grad():
        for _ in range(warmup):
            for s, t in zip(seqs, runner.run(seqs, False)):
                s.append_token(t)
        t0 = time.perf_counter()
        for _ in range(steps):
            for s, t in zip(seqs, runner.run(seqs, False)):
                s.append_token(t)
        dt = time.perf_counter() - t0
    return sample_seqs * steps / dt, torch.get_num_threads(), dt / steps


def cpu_kv_cached_decode(steps: int):
    """The same fp32 arithmetic on the same host cores WITH a KV cache and last-token logits
    (oracle.Qwen3OracleKVCached - not the reference's algorithm): one sequence, 512-token prompt prefilled once
    (untimed), `steps` timed single-token decode steps.  Separates the algorithmic speed-up (what a KV cache buys
    the reference on its own hardware) from the hardware speed-up.  Returns tok/s."""
    import torch
    from oracle import qwen3_oracle as O
    synth = importlib.import_module("nano-vllm-candle_b200.synth")
    cfg = O.MODEL_CONFIGS[MODEL]
    model = O.Qwen3Oracle(cfg, {k: v.float() for k, v in synth.iter_weights(cfg, 0)}, "fp32")
    prompt = synth.make_prompts(BATCH, PROMPT_LEN, cfg.vocab_size, 2)[0]
    with torch.no_grad():
        inc = O.Qwen3OracleKVCached(model, PROMPT_LEN + steps + 2)
        tok = O.argmax_last(inc.step(prompt).numpy())
        tok = O.argmax_last(inc.step([tok]).numpy())          # warm-up step
        t0 = time.perf_counter()
        for _ in range(steps):
            tok = O.argmax_last(inc.step([tok]).numpy())
        dt = time.perf_counter() - t0
    return steps / dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps)
    toks, threads, sec = cpu_reference_decode(steps, args.warmup, CPU_SAMPLE_SEQS)
    sample = (f"{steps} decode steps (after {args.warmup} warm-up) of {CPU_SAMPLE_SEQS} of the {BATCH} sequences at "
              f"context {PROMPT_LEN + 1}+, reference algorithm (full recompute, dense mask, all-position logits), fp32")
    line = {
        "impl": "reference", "metric": "decode tok/s", "value": toks, "unit": "tok/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config(args.gpus),
        "cpu_baseline": {"value": toks, "unit": "tok/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": toks, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference crate (Rust + candle 0.9.1) cannot be built on this image; timed: oracle/ restatement",
    }
    _emit(line)


def _config(n_gpus: int) -> dict:
    return {"workload": f"{MODEL} bf16 decode, batch {BATCH}, {PROMPT_LEN}-token prompts, ignore_eos, paged KV block "
                        f"{BLOCK_SIZE} (BASELINE.json configs[1])",
            "parallelism": f"tp{n_gpus}", "global_batch": BATCH, "prompt_len": PROMPT_LEN,
            "l2": "per-step working set (weights 1.5 GB + KV 4+ GB) is far larger than the 126 MB L2"}


# --------------------------------------------------------------------------------------------
# the product arm
# --------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run --nproc-per-node N")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ops = importlib.import_module("nano-vllm-candle_b200.ops")
    eng = importlib.import_module("nano-vllm-candle_b200.engine")
    synth = importlib.import_module("nano-vllm-candle_b200.synth")

    ctx = ops.Context(local, rank, world)
    if world > 1:
        ctx.init_tp(dist)                  # NCCL communicator + NVLink peer buffers
    cfg = eng.MODEL_CONFIGS[MODEL]
    K, W = args.steps, args.warmup
    total_len = PROMPT_LEN + W + K + 16
    blocks_per_seq = -(-total_len // BLOCK_SIZE)
    nblocks = BATCH * blocks_per_seq + 8
    t0 = time.perf_counter()
    runner = eng.Qwen3ModelRunner.from_weights(
        ctx, cfg, synth.iter_weights(cfg, 0), num_kvcache_blocks=nblocks, kvcache_block_size=BLOCK_SIZE,
        max_num_seqs=BATCH, max_num_batched_tokens=BATCH * PROMPT_LEN, max_model_len=blocks_per_seq * BLOCK_SIZE)
    t_load = time.perf_counter() - t0
    engine = eng.LLMEngine(eng.Scheduler(eng.SchedulerConfig(BATCH, BATCH * PROMPT_LEN, cfg.eos_token_id, nblocks,
                                                             BLOCK_SIZE)), runner)
    prompts = synth.make_prompts(BATCH, PROMPT_LEN, cfg.vocab_size, 2)
    # untimed warm-up: two short sequences through prefill + one decode step (module load, NCCL
    # connection setup, first-launch costs), finished and freed before the timed batch is queued
    for p in prompts[:2]:
        engine.add_request(p[:64], eng.SamplingParams(1.0, 2, True))
    while not engine.is_finished():
        engine.step()
    sp = eng.SamplingParams(1.0, W + K + 8, True)
    for p in prompts:
        engine.add_request(p, sp)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- prefill (one step over 64 x 512 tokens), reported as an extra
    barrier()
    t0 = time.perf_counter()
    s = engine.step()
    barrier()
    t_prefill = time.perf_counter() - t0
    assert s.is_prefill
    # ---- warm-up decode steps (also captures the CUDA graph)
    for _ in range(max(W, 3)):
        s = engine.step()
        assert not s.is_prefill
    # ---- e2e: K decode steps through the public API, host buffers, H2D + D2H inside
    sampler = ClockSampler(local)
    l0 = ctx.l.nvc_ctx_launch_count(ctx.h)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    for _ in range(K):
        engine.step()
    barrier()
    t_e2e = time.perf_counter() - t0
    # ---- device-resident: replay the step K times with inputs already in HBM (CUDA events)
    ms_step, launches_per_step = runner.time_decode(K)
    barrier()
    clocks = sampler.stop()
    if dist is not None:
        t = torch.tensor([t_e2e, ms_step], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e, ms_step = float(t[0]), float(t[1])

    # ---- per-kernel timing (each family over all 28 layers' own weights / KV pools)
    ctx_len = PROMPT_LEN + max(W, 3) + K
    nh_l, kv_l = cfg.num_attention_heads // world, cfg.num_key_value_heads // world
    breakdown = {}
    # gemm_gate_up includes the fused SiLU-mul epilogue (K3 + K6 in one kernel)
    ops_list = ["paged_decode", "gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down", "rmsnorm", "qk_norm_rope",
                "lm_head", "argmax", "noop"] + (["allreduce"] if world > 1 else [])
    for op in ops_list:
        ms, _ = runner.time_op(op, 5)
        breakdown[op + "_us"] = round(ms * 1e3, 2)
    attn_bytes = BATCH * ctx_len * 2 * kv_l * 128 * 2 + 2 * BATCH * nh_l * 128 * 2     # K+V once per kv head, q in, out
    peak, peak_src = _peaks()
    attn_s = breakdown["paged_decode_us"] * 1e-6
    achieved = attn_bytes / attn_s / 1e9
    L = cfg.num_hidden_layers
    per_layer = sum(breakdown[k] for k in ("paged_decode_us", "gemm_qkv_us", "gemm_o_us", "gemm_gate_up_us",
                                           "gemm_down_us", "qk_norm_rope_us")) + 2 * breakdown["rmsnorm_us"]
    if world > 1:
        per_layer += 2 * breakdown["allreduce_us"]
        breakdown["allreduce_exposed_ms_per_step"] = round(2 * L * breakdown["allreduce_us"] / 1e3, 3)
    breakdown["sum_of_kernels_ms"] = round((per_layer * L + breakdown["lm_head_us"] + breakdown["argmax_us"]) / 1e3, 3)
    breakdown["paged_decode_share_of_step"] = round(breakdown["paged_decode_us"] * L / 1e3 / ms_step, 3)
    step_bytes = runner.weight_bytes() - 2 * cfg.vocab_size * cfg.hidden_size + L * attn_bytes   # embed table is gathered, not streamed

    cap = blocks_per_seq
    line = {
        "metric": "decode tok/s", "value": BATCH / (ms_step * 1e-3), "unit": "tok/s", "n_gpus": world, "steps": K,
        "warmup": max(W, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": _config(world), "clocks": clocks,
        "e2e": {"value": BATCH * K / t_e2e, "unit": "tok/s", "ms_per_step": t_e2e / K * 1e3,
                "h2d_bytes_per_step": BATCH * 4 * 4 + BATCH * cap * 4, "d2h_bytes_per_step": BATCH * 4},
        "gpu_launches": launches_per_step * K,
        "roofline": {"kernel": "pdmma::paged_decode_mma_kernel (K5)", "bound": "hbm", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": _ncu_traffic(), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": attn_bytes, "us_per_launch": breakdown["paged_decode_us"],
                     "whole_step_hbm_frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak},
        "breakdown": breakdown,
        "prefill": {"tokens": BATCH * PROMPT_LEN, "tok_per_s": BATCH * PROMPT_LEN / t_prefill, "ms": t_prefill * 1e3},
        "context_len_at_timing": ctx_len, "kernels_per_step": launches_per_step, "weights_load_s": round(t_load, 1),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        runner.close()
        toks, threads, sec = cpu_reference_decode(2, 1, CPU_SAMPLE_SEQS)
        kv_toks = cpu_kv_cached_decode(16)
        line["cpu_baseline_kv_cached"] = {
            "value": kv_toks, "unit": "tok/s", "cores": threads, "kind": "port",
            "sample": "16 single-token decode steps of 1 sequence at context 514+, same fp32 arithmetic with a KV cache "
                      "and last-token logits (NOT the reference's algorithm; separates algorithmic from hardware speed-up)"}
        line["cpu_baseline"] = {
            "value": toks, "unit": "tok/s", "cores": threads, "kind": "port",
            "sample": f"2 decode steps (after 1 warm-up) of {CPU_SAMPLE_SEQS} of the {BATCH} sequences at context "
                      f"{PROMPT_LEN + 1}+, reference algorithm (full recompute per step), fp32, {sec:.1f} s/step"}
    if rank == 0:
        _emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


_OUT = None


def _emit(line) -> None:
    """The one JSON line, on the process's ORIGINAL stdout."""
    out = _OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    # stdout carries exactly one JSON line: everything else written to file descriptor 1 by libraries (NCCL
    # prints its "NCCL version ..." banner there at NCCL_DEBUG=VERSION and above, WARN included) goes to stderr
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
