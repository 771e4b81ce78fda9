"""Cross-check the fp32 restatement against an independent implementation:
HuggingFace transformers' Qwen3ForCausalLM on CPU, fp32, same synthetic checkpoint.
(TEST INFRASTRUCTURE; run by tests/test_oracle_crosscheck.py and by hand.)

    python -m oracle.crosscheck_hf
"""
from __future__ import annotations

import sys
import tempfile

import torch

from .qwen3_oracle import MODEL_CONFIGS, Qwen3Oracle
from .synth import make_weights, make_prompts, write_hf_dir


def crosscheck(cfg_name: str = "qwen3-tiny", seed: int = 0, batch: int = 2, seqlen: int = 24):
    from transformers import AutoModelForCausalLM
    cfg = MODEL_CONFIGS[cfg_name]
    w = make_weights(cfg, seed, norm_jitter=0.1)
    with tempfile.TemporaryDirectory() as d:
        write_hf_dir(d, cfg, w)
        hf = AutoModelForCausalLM.from_pretrained(d, torch_dtype=torch.float32, attn_implementation="eager")
    hf.eval()
    ids = torch.tensor(make_prompts(batch, seqlen, cfg.vocab_size, seed + 1))
    with torch.no_grad():
        ref = hf(input_ids=ids).logits.float()
    orc = Qwen3Oracle(cfg, {k: v.float() for k, v in w.items()})
    got = orc.compute_logits(orc.forward(ids))
    err = (got - ref).abs().max().item()
    cos = torch.nn.functional.cosine_similarity(got.flatten(), ref.flatten(), dim=0).item()
    return err, cos, ref.abs().max().item()


if __name__ == "__main__":
    e, c, m = crosscheck()
    print(f"max-abs diff {e:.3e}  cosine {c:.8f}  (|logit| max {m:.3f})")
    sys.exit(0 if e < 1e-3 and c > 0.99999 else 1)
