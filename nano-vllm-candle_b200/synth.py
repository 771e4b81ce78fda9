"""Seeded synthetic Qwen3 checkpoints and prompts (data generation only: numpy on the host,
no model math; shared by bench.py, the tests and the oracle's golden-vector generator).

No Qwen3 weights exist on this box or the GPU box, and the reference's own random-init
constructors are broken, so "random-init" means: an HF-format
checkpoint with seeded tensors, loaded through the from_hf_dir path
(src/models/qwen3.rs:515-536 tensor names and shapes).

Values: matrices N(0, 0.02^2) (Tensor::randn(0, 0.02), src/models/qwen3.rs:122,291-292,413,511),
norm weights 1 (+ optional jitter for tests; Tensor::ones, src/layers/layernorm.rs:29),
then ROUNDED TO bf16 - the checkpoint dtype of the published Qwen3 models.  The reference
loads any checkpoint as F32 (src/models/qwen3.rs:519), i.e. it computes with exactly these
bf16-representable values widened to f32; so does the oracle; the CUDA path keeps them bf16.

Each tensor has its own PCG64 stream keyed by (seed, crc32(name)) so any subset (a TP
shard, one layer) can be generated without the rest.
"""
from __future__ import annotations

import json
import os
import zlib
from typing import Dict, Iterator, List, Tuple

import numpy as np
import torch



def tensor_specs(cfg) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(name, shape, kind) for every tensor from_hf_dir reads; kind in {"matrix","norm"}."""
    H, hd = cfg.hidden_size, cfg.head_dim
    nh, kv, I = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.intermediate_size
    specs = [("model.embed_tokens.weight", (cfg.vocab_size, H), "matrix")]
    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}"
        specs += [
            (f"{p}.self_attn.q_proj.weight", (nh * hd, H), "matrix"),
            (f"{p}.self_attn.k_proj.weight", (kv * hd, H), "matrix"),
            (f"{p}.self_attn.v_proj.weight", (kv * hd, H), "matrix"),
            (f"{p}.self_attn.o_proj.weight", (H, nh * hd), "matrix"),
            (f"{p}.self_attn.q_norm.weight", (hd,), "norm"),
            (f"{p}.self_attn.k_norm.weight", (hd,), "norm"),
            (f"{p}.mlp.gate_proj.weight", (I, H), "matrix"),
            (f"{p}.mlp.up_proj.weight", (I, H), "matrix"),
            (f"{p}.mlp.down_proj.weight", (H, I), "matrix"),
            (f"{p}.input_layernorm.weight", (H,), "norm"),
            (f"{p}.post_attention_layernorm.weight", (H,), "norm"),
        ]
    specs += [("model.norm.weight", (H,), "norm"), ("lm_head.weight", (cfg.vocab_size, H), "matrix")]
    return specs


def make_tensor(name: str, shape, kind: str, seed: int, std: float = 0.02,
                norm_jitter: float = 0.0) -> torch.Tensor:
    """One bf16 tensor of the synthetic checkpoint."""
    rng = np.random.Generator(np.random.PCG64([seed, zlib.crc32(name.encode())]))
    if kind == "norm":
        a = np.ones(shape, dtype=np.float32)
        if norm_jitter:
            a += norm_jitter * rng.standard_normal(shape, dtype=np.float32)
    else:
        a = rng.standard_normal(shape, dtype=np.float32)
        a *= np.float32(std)
    return torch.from_numpy(a).to(torch.bfloat16)


def iter_weights(cfg, seed: int = 0, std: float = 0.02,
                 norm_jitter: float = 0.0) -> Iterator[Tuple[str, torch.Tensor]]:
    for name, shape, kind in tensor_specs(cfg):
        yield name, make_tensor(name, shape, kind, seed, std, norm_jitter)


def make_weights(cfg, seed: int = 0, std: float = 0.02,
                 norm_jitter: float = 0.0) -> Dict[str, torch.Tensor]:
    return dict(iter_weights(cfg, seed, std, norm_jitter))


def hf_config_json(cfg) -> dict:
    """config.json fields read by Qwen3Config::from_hf_dir (src/models/qwen3.rs:36-52,77-101)."""
    names = ("vocab_size", "hidden_size", "head_dim", "num_hidden_layers", "num_attention_heads",
             "num_key_value_heads", "intermediate_size", "max_position_embeddings", "rms_norm_eps", "hidden_act",
             "rope_theta", "bos_token_id", "eos_token_id")
    d = {n: getattr(cfg, n) for n in names}
    d.update(architectures=["Qwen3ForCausalLM"], model_type="qwen3", tie_word_embeddings=False,
             torch_dtype="bfloat16", attention_bias=False)
    return d


def write_hf_dir(path: str, cfg, weights: Dict[str, torch.Tensor]) -> None:
    """config.json + single un-sharded model.safetensors (src/models/qwen3.rs:516-521)."""
    from safetensors.torch import save_file
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(hf_config_json(cfg), f, indent=1)
    save_file({k: v.contiguous() for k, v in weights.items()}, os.path.join(path, "model.safetensors"))


def make_prompts(num: int, length, vocab: int, seed: int) -> List[List[int]]:
    """Uniform random token ids in [0, vocab).  `length` is an int
    or an inclusive (lo, hi) range."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = []
    for _ in range(num):
        n = length if isinstance(length, int) else int(rng.integers(length[0], length[1] + 1))
        out.append(rng.integers(0, vocab, size=n).astype(np.int64).tolist())
    return out


def iter_weights_device(cfg, seed: int = 0, std: float = 0.02, device="cuda"):
    """Same tensor set generated on the GPU with torch's Philox generator: NOT the numpy streams of
    make_tensor (no oracle parity), only for throughput sweeps of the large models where host-side
    generation of 8-32 B normals would dominate the run."""
    g = torch.Generator(device=device)
    for i, (name, shape, kind) in enumerate(tensor_specs(cfg)):
        if kind == "norm":
            yield name, torch.ones(shape, dtype=torch.bfloat16, device=device)
        else:
            g.manual_seed(seed * 1000003 + i)
            t = torch.empty(shape, dtype=torch.bfloat16, device=device)
            t.normal_(0.0, std, generator=g)
            yield name, t
