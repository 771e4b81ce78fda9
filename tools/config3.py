"""BASELINE config 3: Qwen3-8B bf16 on 1 x B200, batch 256 mixed prefill / decode, paged KV cache
block size 256 (half of the sequences arrive at step 0, half staggered so
that prefill and decode work interleave; prompt lengths uniform [256, 1024] seed 3; 256 decode
tokens, ignore-eos).  Runs the whole job through LLMEngine.step() twice:

  reference scheduling   a step is a whole-prompt prefill step OR a decode step (scheduler.rs:106-111)
  chunked prefill        mixed steps under a token budget

and reports wall-clock job throughput plus the time between consecutive tokens of a sequence
(the stall a running sequence sees while prompts are prefilled).

    python tools/config3.py [--model qwen3-8b] [--batch 256] [--budget 2048] [--out profiles/x.json]
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


def run(eng, runner, cfg, prompts, arrivals, decode_tokens, nblocks, budget, chunked):
    sc = eng.SchedulerConfig(len(prompts), budget, cfg.eos_token_id, nblocks, 256, enable_chunked_prefill=chunked)
    e = eng.LLMEngine(eng.Scheduler(sc), runner, greedy=True)
    sp = eng.SamplingParams(1.0, decode_tokens, True)
    added, step = 0, 0
    last_tok_time, gaps = {}, []
    n_prefill = n_decode = n_mixed = n_steps = 0
    prefill_tokens = decode_tokens_done = 0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    while added < len(prompts) or not e.is_finished():
        while added < len(prompts) and arrivals[added] <= step:
            e.add_request(prompts[added], sp)
            added += 1
        step += 1
        if e.is_finished():
            continue
        # schedule + run + post_process by hand to know which sequences got a token in this step
        si = e.schedule()
        if chunked and si.num_decode_seqs < len(si.seq_ids):
            toks = runner.step_mixed(si.ids, si.positions, si.cu_seqlens, si.slot_mapping, si.block_tables,
                                     si.context_lens, si.num_decode_seqs)
        elif si.is_prefill:
            toks = runner.prefill(si.ids, si.positions, si.cu_seqlens, si.slot_mapping)
        else:
            toks = runner.decode(si.ids, si.positions, si.slot_mapping, si.block_tables, si.context_lens)
        e.post_process(toks)
        now = time.perf_counter()
        nd = si.num_decode_seqs
        nq = len(si.ids)
        n_steps += 1
        n_mixed += int(0 < nd < len(si.seq_ids))
        n_prefill += int(nd == 0)
        n_decode += int(nd == len(si.seq_ids))
        prefill_tokens += nq - nd
        decode_tokens_done += nd
        for sid, m in zip(si.seq_ids.tolist(), si.emit.tolist()):
            if m:
                if sid in last_tok_time:
                    gaps.append(now - last_tok_time[sid])
                last_tok_time[sid] = now
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    e.close()
    g = np.asarray(gaps) * 1e3
    return dict(scheduler="chunked" if chunked else "reference", token_budget=budget, wall_s=round(wall, 3), steps=n_steps,
                prefill_steps=n_prefill, decode_steps=n_decode, mixed_steps=n_mixed,
                prefill_tokens=prefill_tokens, decode_tokens=decode_tokens_done,
                job_tok_s=round((prefill_tokens + decode_tokens_done) / wall),
                generated_tok_s=round(len(prompts) * decode_tokens / wall),
                inter_token_ms=dict(mean=round(float(g.mean()), 2), p50=round(float(np.percentile(g, 50)), 2),
                                    p99=round(float(np.percentile(g, 99)), 2), max=round(float(g.max()), 2)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-8b")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--decode-tokens", type=int, default=256)
    ap.add_argument("--budget", type=int, nargs="+", default=[2048])
    ap.add_argument("--min-prompt", type=int, default=256)
    ap.add_argument("--max-prompt", type=int, default=1024)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    ops = importlib.import_module("nano-vllm-candle_b200.ops")
    eng = importlib.import_module("nano-vllm-candle_b200.engine")
    synth = importlib.import_module("nano-vllm-candle_b200.synth")
    ctx = ops.Context(0, 0, 1)
    cfg = eng.MODEL_CONFIGS[a.model]
    rng = np.random.default_rng(3)
    B = a.batch
    lens = rng.integers(a.min_prompt, a.max_prompt + 1, size=B)
    prompts = [rng.integers(0, cfg.vocab_size, size=int(n)).tolist() for n in lens]
    # half at step 0, the other half one every second engine step
    arrivals = [0] * (B // 2) + [2 * (i + 1) for i in range(B - B // 2)]
    bps = -(-(a.max_prompt + a.decode_tokens + 8) // 256)
    nblocks = B * bps + 8
    ref_budget = 32768
    runner = eng.Qwen3ModelRunner.from_weights(ctx, cfg, synth.iter_weights_device(cfg, 0, device="cuda:0"),
                                               num_kvcache_blocks=nblocks, kvcache_block_size=256, max_num_seqs=B,
                                               max_num_batched_tokens=ref_budget, max_model_len=bps * 256)
    torch.cuda.empty_cache()
    # warm-up: graphs for the batch sizes come and go with the arrivals; one short job compiles nothing
    # but touches every kernel family
    run(eng, runner, cfg, [p[:64] for p in prompts[:8]], [0] * 8, 4, nblocks, ref_budget, False)
    run(eng, runner, cfg, [p[:64] for p in prompts[:8]], [0] * 4 + [1, 2, 3, 4], 4, nblocks, 64, True)
    rows = [run(eng, runner, cfg, prompts, arrivals, a.decode_tokens, nblocks, ref_budget, False)]
    for b in a.budget:
        rows.append(run(eng, runner, cfg, prompts, arrivals, a.decode_tokens, nblocks, max(b, B), True))
    out = dict(config="BASELINE configs[2]: %s bf16, 1 x B200, batch %d mixed prefill/decode, block size 256" % (a.model, B),
               prompt_lens="uniform[%d, %d] seed 3 (sum %d)" % (a.min_prompt, a.max_prompt, int(lens.sum())),
               arrivals="half at step 0, then one sequence every second step", decode_tokens_per_seq=a.decode_tokens,
               timing="wall clock around the whole job through LLMEngine schedule / runner step / post_process", runs=rows)
    print(json.dumps(out, indent=1))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)
    runner.close()


if __name__ == "__main__":
    main()
