"""Throughput sweep for the BASELINE.md table: prefill tok/s and decode tok/s at several batch
sizes for one model (synthetic device-generated weights), with the HBM / tensor roofline
fractions.  One process per GPU (TP = WORLD_SIZE) when launched with torch.distributed.run.

    python tools/sweep.py --model qwen3-8b --batches 1 64 256 --prompt-len 512 --decode-ctx 1024
"""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-8b")
    ap.add_argument("--batches", type=int, nargs="+", default=[1, 64, 256])
    ap.add_argument("--prompt-len", type=int, default=512)
    ap.add_argument("--decode-steps", type=int, default=16)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
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
    cfg = eng.MODEL_CONFIGS[a.model]
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else \
        {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0, "bf16_tflops": 1590.0}
    bmax = max(a.batches)
    T = a.prompt_len
    total_len = T + a.decode_steps + 24
    bps = -(-total_len // 256)
    nblocks = bmax * bps + 8
    max_tokens_batched = min(bmax * T, 32768)
    runner = eng.Qwen3ModelRunner.from_weights(ctx, cfg, synth.iter_weights_device(cfg, 0, device=f"cuda:{local}"),
                                               num_kvcache_blocks=nblocks, kvcache_block_size=256, max_num_seqs=bmax,
                                               max_num_batched_tokens=max_tokens_batched, max_model_len=bps * 256)
    torch.cuda.empty_cache()
    L, H, I, V = cfg.num_hidden_layers, cfg.hidden_size, cfg.intermediate_size, cfg.vocab_size
    nh, kv = cfg.num_attention_heads, cfg.num_key_value_heads
    layer_params = H * (nh + 2 * kv) * 128 + nh * 128 * H + 3 * H * I
    rows = []

    def sync():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for B in a.batches:
        e = eng.LLMEngine(eng.Scheduler(eng.SchedulerConfig(B, max_tokens_batched, cfg.eos_token_id, nblocks, 256)), runner, greedy=True)
        prompts = synth.make_prompts(B, T, cfg.vocab_size, 10 + B)
        if B == a.batches[0]:           # warm-up
            for p in prompts[:1]:
                e.add_request(p[:64], eng.SamplingParams(1.0, 2, True))
            while not e.is_finished():
                e.step()
        for p in prompts:
            e.add_request(p, eng.SamplingParams(1.0, a.decode_steps + 8, True))
        sync()
        t0 = time.perf_counter()
        n_prefill_steps = 0
        while True:                      # prefill steps (max_num_batched_tokens may split the batch)
            s = e.step()
            n_prefill_steps += 1
            if e.num_waiting() == 0:
                break
        sync()
        t_prefill = time.perf_counter() - t0
        for _ in range(4):
            e.step()
        sync()
        t0 = time.perf_counter()
        for _ in range(a.decode_steps):
            e.step()
        sync()
        t_e2e = (time.perf_counter() - t0) / a.decode_steps
        ms_dev, kernels = runner.time_decode(a.decode_steps)
        if dist is not None:
            t = torch.tensor([t_prefill, t_e2e, ms_dev], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_prefill, t_e2e, ms_dev = [float(x) for x in t]
        ctx_len = T + n_prefill_steps - 1 + 4 + a.decode_steps
        ops_us = {}
        for op in ["paged_decode", "gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down", "lm_head"] + (["allreduce"] if world > 1 else []):
            ops_us[op] = round(runner.time_op(op, 3)[0] * 1e3, 2)
        kv_l, nh_l = kv // world, nh // world
        attn_bytes = B * ctx_len * 2 * kv_l * 128 * 2 + 2 * B * nh_l * 128 * 2
        step_bytes = (L * layer_params + V * H) * 2 / world + L * attn_bytes          # weights (+lm_head) + KV, per rank
        gemm_flops = 2.0 * B * (L * layer_params + V * H) / world
        prefill_flops = 2.0 * B * T * L * layer_params / world + L * 4.0 * nh_l * 128 * B * T * T / 2
        gemm_us = ops_us["gemm_qkv"] + ops_us["gemm_o"] + ops_us["gemm_gate_up"] + ops_us["gemm_down"]
        row = dict(model=a.model, tp=world, batch=B, prompt_len=T, ctx_len=ctx_len,
                   prefill_tok_s=round(B * T / t_prefill), prefill_ms=round(t_prefill * 1e3, 1),
                   prefill_frac_of_bf16_sustained=round(prefill_flops / t_prefill / 1e12 / peaks["bf16_tflops_sustained"], 3),
                   decode_tok_s=round(B / (ms_dev * 1e-3)), decode_ms_per_step=round(ms_dev, 3),
                   decode_tok_s_e2e=round(B / t_e2e), kernels_per_step=kernels,
                   step_hbm_frac=round(step_bytes / (ms_dev * 1e-3) / 1e9 / peaks["hbm_gbs"], 3),
                   paged_attn_gbs=round(attn_bytes / ops_us["paged_decode"] / 1e3, 1),
                   paged_attn_frac=round(attn_bytes / ops_us["paged_decode"] / 1e3 / peaks["hbm_gbs"], 3),
                   layer_gemm_tflops=round(2.0 * B * layer_params / world / gemm_us / 1e6, 1),
                   layer_gemm_frac_of_bf16_sustained=round(2.0 * B * layer_params / world / gemm_us / 1e6 / peaks["bf16_tflops_sustained"], 3),
                   layer_gemm_weight_gbs=round(layer_params * 2 / world / gemm_us / 1e3, 1),
                   ops_us=ops_us)
        if world > 1:
            row["allreduce_exposed_ms_per_step"] = round(2 * L * ops_us["allreduce"] / 1e3, 3)
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        while not e.is_finished():       # drain so the blocks are free for the next batch size
            e.step()
        e.close()
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)
    runner.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
