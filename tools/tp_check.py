"""Tensor-parallel parity check, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/tp_check.py [--model qwen3-tiny]

Every rank builds the same synthetic checkpoint, slices its shard inside the library
(q/kv heads, MLP columns, vocab rows), runs the SAME replicated scheduler, and all-reduces
after o_proj / down_proj through NCCL.  Rank 0 checks the greedy tokens teacher-forced against
the fp32 oracle (oracle/ is test infrastructure; this script is a test driver) and that every
rank produced identical tokens."""
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
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-tiny")
    ap.add_argument("--max-tokens", type=int, default=24)
    ap.add_argument("--chunked", type=int, default=0, metavar="BUDGET",
                    help="chunked prefill with this token budget (mixed steps; all-reduce through NCCL)")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ops = importlib.import_module("nano-vllm-candle_b200.ops")
    eng = importlib.import_module("nano-vllm-candle_b200.engine")
    synth = importlib.import_module("nano-vllm-candle_b200.synth")
    ctx = ops.Context(local, rank, world)
    ctx.init_tp(dist)                      # NCCL communicator + NVLink peer buffers
    cfg = eng.MODEL_CONFIGS[a.model]
    tiny = a.model == "qwen3-tiny"
    bs = 16 if tiny else 256
    w = synth.make_weights(cfg, 0, norm_jitter=0.1 if tiny else 0.0)
    runner = eng.Qwen3ModelRunner.from_weights(ctx, cfg, w.items(), num_kvcache_blocks=64, kvcache_block_size=bs,
                                               max_num_seqs=8, max_num_batched_tokens=2048, max_model_len=64 * bs // 8)
    prompts = [synth.make_prompts(1, n, cfg.vocab_size, 300 + n)[0] for n in (9, 40, 33, 17)]
    sc = eng.SchedulerConfig(8, a.chunked or 2048, cfg.eos_token_id, 64, bs, enable_chunked_prefill=a.chunked > 0)
    e = eng.LLMEngine(eng.Scheduler(sc), runner, greedy=True)
    t0 = time.perf_counter()
    mixed = 0
    if a.chunked:           # step by step to count the mixed steps (decode tokens next to prompt chunks)
        for i, p in enumerate(prompts):
            e.add_request(p, eng.SamplingParams(1.0, a.max_tokens, True))
            if i == 1:      # two sequences are already decoding when the others arrive
                for _ in range(3):
                    e.step()
        while not e.is_finished():
            s = e.step()
            mixed += int(0 < s.num_decode_seqs < s.num_query_tokens)
        outs = [e.output_of(i)[0] for i in range(len(prompts))]
    else:
        outs = [o.token_ids for o in e.generate(prompts, eng.SamplingParams(1.0, a.max_tokens, True))]
    dt = time.perf_counter() - t0
    # all ranks must agree token for token (replicated scheduler + identical argmax merge)
    mine = torch.tensor(outs, dtype=torch.int64, device="cuda")
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    same = all(torch.equal(g, mine) for g in gathered)
    result = {"model": a.model, "tp": world, "ranks_agree": bool(same), "seconds": dt,
              "nvls": ctx.nvls_ready(), "nvls_status": ctx.nvls_status()}
    if a.chunked:
        result.update(chunked_budget=a.chunked, mixed_steps=mixed)
        same = same and mixed > 0
    ok = same
    if rank == 0:
        from oracle import qwen3_oracle as O
        model = O.Qwen3Oracle(O.MODEL_CONFIGS[a.model], {k: v.float() for k, v in w.items()}, "fp32")
        worst, exact, total = 0.0, 0, 0
        for p, g in zip(prompts, outs):
            ids = torch.tensor([p + g[:-1]])
            logits = model.compute_logits(model.forward(ids))[0]
            for i, t in enumerate(g):
                row = logits[len(p) - 1 + i]
                worst = max(worst, float(row.max() - row[t]))
                exact += int(O.argmax_last(row.numpy()) == t)
                total += 1
        result.update(worst_logit_deficit=worst, exact_argmax=exact, tokens=total)
        ok = ok and worst <= 4e-2
        result["ok"] = bool(ok)
        print(json.dumps(result), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(result, f)
    runner.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
