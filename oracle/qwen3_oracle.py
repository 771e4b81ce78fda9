"""fp32 CPU restatement of the reference's Qwen3 forward (TEST INFRASTRUCTURE).

Every function cites the reference file:line it follows (paths relative to the
reference crate root).  Arithmetic is torch CPU float32, mirroring candle's
F32 CPU backend (weights are loaded as DType::F32, src/models/qwen3.rs:519).

Parity unpinned for attention / full forward (see oracle/__init__.py); pinned
for Linear / RMSNorm / SiluAndMul by the reference's unit-test vectors.

Two modes:
  * mode="fp32"  - the reference's algorithm exactly as written (full-sequence
                   recompute, dense -1e9 additive mask, unfused ops).
  * mode="bf16"  - same algorithm, but tensors are rounded to bf16 at the points
                   where the CUDA path stores bf16 (GEMM outputs, norm outputs,
                   q/k after rope, attention output, activation output).  Used
                   only to separate rounding-placement effects from kernel bugs.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, asdict
from typing import Dict, List, Optional, Sequence as Seq, Tuple

import numpy as np
import torch


# --------------------------------------------------------------------------- config
@dataclass
class Qwen3Config:
    """src/models/qwen3.rs:19-34 (fields), :77-101 (from_hf_dir)."""
    vocab_size: int = 151936
    hidden_size: int = 1024
    head_dim: int = 128
    num_hidden_layers: int = 28
    num_attention_heads: int = 16
    num_key_value_heads: int = 8
    intermediate_size: int = 3072
    max_position_embeddings: int = 40960
    rms_norm_eps: float = 1e-6
    hidden_act: str = "silu"
    rope_theta: float = 1e6
    bos_token_id: int = 151643
    eos_token_id: int = 151645

    @staticmethod
    def from_hf_dir(model_dir: str) -> "Qwen3Config":
        with open(os.path.join(model_dir, "config.json")) as f:
            d = json.load(f)
        hd = d.get("head_dim") or d["hidden_size"] // d["num_attention_heads"]
        return Qwen3Config(
            vocab_size=d["vocab_size"], hidden_size=d["hidden_size"], head_dim=hd,
            num_hidden_layers=d["num_hidden_layers"],
            num_attention_heads=d["num_attention_heads"],
            num_key_value_heads=d["num_key_value_heads"],
            intermediate_size=d["intermediate_size"],
            max_position_embeddings=d["max_position_embeddings"],
            rms_norm_eps=d["rms_norm_eps"], hidden_act=d["hidden_act"],
            rope_theta=d.get("rope_theta") or 1e6,
            bos_token_id=d["bos_token_id"], eos_token_id=d["eos_token_id"])

    def to_hf_json(self) -> dict:
        d = asdict(self)
        d["architectures"] = ["Qwen3ForCausalLM"]
        d["model_type"] = "qwen3"
        d["tie_word_embeddings"] = False
        d["torch_dtype"] = "bfloat16"
        d["attention_bias"] = False
        return d


# Published Qwen3 dense configs.
MODEL_CONFIGS: Dict[str, Qwen3Config] = {
    "qwen3-0.6b": Qwen3Config(),
    "qwen3-8b": Qwen3Config(hidden_size=4096, num_hidden_layers=36, num_attention_heads=32,
                            num_key_value_heads=8, intermediate_size=12288),
    "qwen3-14b": Qwen3Config(hidden_size=5120, num_hidden_layers=40, num_attention_heads=40,
                             num_key_value_heads=8, intermediate_size=17408),
    "qwen3-32b": Qwen3Config(hidden_size=5120, num_hidden_layers=64, num_attention_heads=64,
                             num_key_value_heads=8, intermediate_size=25600),
    # toy config for CPU-speed tests (same structure: GQA 2:1, head_dim 128)
    "qwen3-tiny": Qwen3Config(vocab_size=2048, hidden_size=256, num_hidden_layers=3,
                              num_attention_heads=4, num_key_value_heads=2,
                              intermediate_size=512, max_position_embeddings=4096,
                              bos_token_id=1, eos_token_id=2),
}


# --------------------------------------------------------------------------- rounding helper
def _r(x: torch.Tensor, mode: str) -> torch.Tensor:
    """Round-to-bf16-and-back when emulating the CUDA path's storage points."""
    if mode == "bf16":
        return x.to(torch.bfloat16).to(torch.float32)
    return x


# --------------------------------------------------------------------------- layers
def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """candle_nn::Linear as used by src/layers/linear.rs:33-38,72-77,99-104,184-198:
    y = x @ W^T (+ b); W is [out, in]."""
    y = x @ weight.t()
    if bias is not None:
        y = y + bias
    return y


def row_parallel_linear(x, weight, bias, tp_rank: int):
    """src/layers/linear.rs:184-198: bias only on rank 0, and NO all-reduce."""
    return linear(x, weight, bias if tp_rank == 0 else None)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float,
             residual: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """src/layers/layernorm.rs:44-60.
    sum = f32(x) + f32(res); var = mean(sum^2, -1); out = sum * (1/sqrt(var+eps)) -> orig dtype, * weight."""
    orig = x.dtype
    if residual is not None:
        s = x.to(torch.float32) + residual.to(torch.float32)
        new_res = s
    else:
        s = x.to(torch.float32)
        new_res = None
    var = s.square().mean(dim=-1, keepdim=True)
    x_norm = s * (var + eps).sqrt().reciprocal()
    out = x_norm.to(orig) * weight
    return out, new_res


def build_cos_sin(t: int, head_size: int, base: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """src/layers/rotary_embedding.rs:56-80.  All arithmetic in f32:
    inv_freq_j = 1 / base.powf(2j/dim); ang = (pos as f32) * inv_freq_j; cos/sin in f32.

    Rust's f32::powf / cos / sin lower to the platform libm (glibc powf/cosf/sinf), which are
    correctly rounded (computed in double, rounded once).  numpy's float32 ufuncs use SIMD
    approximations that can be 1 ulp off, and a 1-ulp inv_freq error grows to ~2e-4 rad at
    position 4096; so each libm call is restated as "evaluate in float64 on the f32 inputs,
    round once to f32"."""
    half = head_size // 2
    j = np.arange(half, dtype=np.float32)
    exponent = ((np.float32(2.0) * j) / np.float32(head_size)).astype(np.float32)
    powed = np.power(np.float64(np.float32(base)), exponent.astype(np.float64)).astype(np.float32)
    inv_freq = (np.float32(1.0) / powed).astype(np.float32)
    pos = np.arange(t, dtype=np.float32)
    ang = (pos[:, None] * inv_freq[None, :]).astype(np.float32)
    cos = np.cos(ang.astype(np.float64)).astype(np.float32)
    sin = np.sin(ang.astype(np.float64)).astype(np.float32)
    return torch.from_numpy(cos), torch.from_numpy(sin)


def apply_rotary(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """src/layers/rotary_embedding.rs:82-91 (half-split / neox style)."""
    x1, x2 = x.chunk(2, dim=-1)
    y1 = x1 * cos - x2 * sin
    y2 = x2 * cos + x1 * sin
    return torch.cat([y1, y2], dim=-1)


def rope_apply(q: torch.Tensor, k: torch.Tensor, base: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """src/layers/rotary_embedding.rs:93-107.  q [B,nh,T,hd], k [B,kv,T,hd]; positions 0..T."""
    t, hd = q.shape[2], q.shape[3]
    cos, sin = build_cos_sin(t, hd, base)
    cos = cos.reshape(1, 1, t, hd // 2)
    sin = sin.reshape(1, 1, t, hd // 2)
    return apply_rotary(q, cos, sin), apply_rotary(k, cos, sin)


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    """src/layers/activation.rs:13-18: silu(x[..., :d]) * x[..., d:]."""
    a, b = x.chunk(2, dim=-1)
    return torch.nn.functional.silu(a) * b


def gqa_expand(x: torch.Tensor, repeat: int) -> torch.Tensor:
    """src/models/qwen3.rs:240-256: unsqueeze -> expand -> reshape (numpy.repeat order,
    pinned by tests/debug_layer_test.rs:10-71)."""
    b, kv, t, hd = x.shape
    return x.unsqueeze(2).expand(b, kv, repeat, t, hd).reshape(b, kv * repeat, t, hd)


# --------------------------------------------------------------------------- model
class Qwen3Oracle:
    """Qwen3ForCausalLM restated: src/models/qwen3.rs:104-551.

    `weights` maps HF tensor names (src/models/qwen3.rs:146-187,301-314,351-363,
    431-447,525-528) to float32 torch tensors.
    """

    def __init__(self, cfg: Qwen3Config, weights: Dict[str, torch.Tensor], mode: str = "fp32"):
        assert mode in ("fp32", "bf16")
        self.cfg = cfg
        self.mode = mode
        self.embed = weights["model.embed_tokens.weight"].float()
        self.layers = []
        for i in range(cfg.num_hidden_layers):
            p = f"model.layers.{i}"
            w_qkv = torch.cat([weights[f"{p}.self_attn.q_proj.weight"],
                               weights[f"{p}.self_attn.k_proj.weight"],
                               weights[f"{p}.self_attn.v_proj.weight"]], dim=0).float()   # :171
            w_gu = torch.cat([weights[f"{p}.mlp.gate_proj.weight"],
                              weights[f"{p}.mlp.up_proj.weight"]], dim=0).float()         # :310
            self.layers.append(dict(
                w_qkv=w_qkv,
                w_o=weights[f"{p}.self_attn.o_proj.weight"].float(),
                q_norm=weights[f"{p}.self_attn.q_norm.weight"].float().reshape(1, -1),   # :180
                k_norm=weights[f"{p}.self_attn.k_norm.weight"].float().reshape(1, -1),   # :186
                w_gu=w_gu,
                w_down=weights[f"{p}.mlp.down_proj.weight"].float(),
                ln1=weights[f"{p}.input_layernorm.weight"].float().reshape(1, -1),        # :356
                ln2=weights[f"{p}.post_attention_layernorm.weight"].float().reshape(1, -1),
            ))
        self.norm = weights["model.norm.weight"].float().reshape(1, -1)
        # :525-528: lm_head.weight [V,H] transposed to [H,V]; no tied-embedding fallback.
        self.lm_head = weights["lm_head.weight"].float().t()

    # -- src/models/qwen3.rs:202-281
    def attention(self, x: torch.Tensor, L: dict) -> torch.Tensor:
        cfg, m = self.cfg, self.mode
        b, t, h = x.shape
        nh, kv, hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        x2d = x.reshape(b * t, h)
        qkv = _r(linear(x2d, L["w_qkv"]), m)                       # :205
        q_size, kv_size = nh * hd, kv * hd
        q2d = qkv.narrow(1, 0, q_size)
        k2d = qkv.narrow(1, q_size, kv_size)
        v2d = qkv.narrow(1, q_size + kv_size, kv_size)
        q = q2d.reshape(b, t, nh, hd).transpose(1, 2)              # [B,nh,T,hd]  :216
        k = k2d.reshape(b, t, kv, hd).transpose(1, 2)
        v = v2d.reshape(b, t, kv, hd).transpose(1, 2)
        q_n, _ = rms_norm(q.reshape(b * nh * t, hd), L["q_norm"], cfg.rms_norm_eps)   # :229
        k_n, _ = rms_norm(k.reshape(b * kv * t, hd), L["k_norm"], cfg.rms_norm_eps)
        q = q_n.reshape(b, nh, t, hd)
        k = k_n.reshape(b, kv, t, hd)
        q, k = rope_apply(q, k, float(np.float32(cfg.rope_theta)))  # :234 (theta cast `as f32`, :135)
        q, k = _r(q, m), _r(k, m)
        if kv != nh:                                               # :240-256
            k = gqa_expand(k, nh // kv)
            v = gqa_expand(v, nh // kv)
        scale = float(np.float32(hd) ** np.float32(-0.5))          # :132 (f32 powf)
        scores = (q @ k.transpose(2, 3)) * scale                   # :258-259
        idx = torch.arange(t)
        mask = torch.where(idx[None, :] > idx[:, None], torch.tensor(-1e9), torch.tensor(0.0))  # :260-269
        scores = scores + mask.reshape(1, 1, t, t).to(torch.float32)
        probs = torch.softmax(scores.to(torch.float32), dim=-1)    # :271-273
        ctx = probs @ v                                            # :275
        ctx2d = _r(ctx.transpose(1, 2).reshape(b * t, nh * hd), m) # :276-277
        out2d = _r(row_parallel_linear(ctx2d, L["w_o"], None, 0), m)  # :278
        return out2d.reshape(b, t, h)

    # -- src/models/qwen3.rs:323-327
    def mlp(self, x: torch.Tensor, L: dict) -> torch.Tensor:
        m = self.mode
        gu = _r(linear(x, L["w_gu"]), m)
        a = _r(silu_and_mul(gu), m)
        return _r(linear(a, L["w_down"]), m)

    # -- src/models/qwen3.rs:374-399
    def decoder_layer(self, x, residual, L):
        cfg, m = self.cfg, self.mode
        if residual is not None:
            x_norm, residual = rms_norm(x, L["ln1"], cfg.rms_norm_eps, residual)
        else:
            x_norm, _ = rms_norm(x, L["ln1"], cfg.rms_norm_eps, None)
            residual = x.clone()
        attn_out = self.attention(_r(x_norm, m), L)
        h_norm, residual = rms_norm(attn_out, L["ln2"], cfg.rms_norm_eps, residual)
        h = self.mlp(_r(h_norm, m), L)
        return h, residual

    # -- src/models/qwen3.rs:458-499
    def forward(self, input_ids: torch.Tensor, return_layers: bool = False):
        b, t = input_ids.shape
        h = self.embed[input_ids.reshape(-1).long()].reshape(b, t, self.cfg.hidden_size)
        residual = None
        per_layer = []
        for L in self.layers:
            h, residual = self.decoder_layer(h, residual, L)
            if return_layers:
                per_layer.append((h.clone(), residual.clone()))
        h, _ = rms_norm(h, self.norm, self.cfg.rms_norm_eps, residual)   # :497
        h = _r(h, self.mode)
        return (h, per_layer) if return_layers else h

    # -- src/models/qwen3.rs:542-550 (logits for ALL positions)
    def compute_logits(self, hidden: torch.Tensor) -> torch.Tensor:
        b, t, h = hidden.shape
        return (hidden.reshape(b * t, h) @ self.lm_head).reshape(b, t, self.cfg.vocab_size)


class Qwen3OracleKVCached:
    """The SAME arithmetic as Qwen3Oracle (fp32 mode), evaluated incrementally: K / V of earlier
    positions are kept per layer, a step computes only the new positions and logits only for the
    last one.  Not the reference's algorithm (it recomputes everything each step,
    src/engine/llm_engine.rs:146-189) - it exists so that bench.py can separate the ALGORITHMIC
    speed-up (KV cache, last-token logits) from the hardware speed-up, and
    tests/test_oracle_goldens.py checks it against the full-recompute oracle."""

    def __init__(self, model: Qwen3Oracle, max_len: int):
        assert model.mode == "fp32"
        self.m = model
        cfg = model.cfg
        self.cos, self.sin = build_cos_sin(max_len, cfg.head_dim, float(np.float32(cfg.rope_theta)))
        self.k = [None] * cfg.num_hidden_layers      # [kv, T, hd] per layer
        self.v = [None] * cfg.num_hidden_layers
        self.len = 0

    def step(self, new_ids) -> torch.Tensor:
        """Appends the tokens `new_ids` (positions len .. len + n) and returns the fp32 logits [V]
        of the last one."""
        M, cfg = self.m, self.m.cfg
        nh, kv, hd, eps = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim, cfg.rms_norm_eps
        ids = torch.as_tensor(new_ids, dtype=torch.long)
        n, p0 = len(ids), self.len
        cos = self.cos[p0:p0 + n].reshape(1, n, hd // 2)
        sin = self.sin[p0:p0 + n].reshape(1, n, hd // 2)
        scale = float(np.float32(hd) ** np.float32(-0.5))
        h = M.embed[ids]
        residual = None
        for li, L in enumerate(M.layers):
            if residual is not None:
                x, residual = rms_norm(h, L["ln1"], eps, residual)
            else:
                x, _ = rms_norm(h, L["ln1"], eps, None)
                residual = h.clone()
            qkv = linear(x, L["w_qkv"])
            q = qkv[:, :nh * hd].reshape(n, nh, hd).transpose(0, 1)                 # [nh, n, hd]
            k = qkv[:, nh * hd:(nh + kv) * hd].reshape(n, kv, hd).transpose(0, 1)
            v = qkv[:, (nh + kv) * hd:].reshape(n, kv, hd).transpose(0, 1)
            q = rms_norm(q.reshape(nh * n, hd), L["q_norm"], eps)[0].reshape(nh, n, hd)
            k = rms_norm(k.reshape(kv * n, hd), L["k_norm"], eps)[0].reshape(kv, n, hd)
            q, k = apply_rotary(q, cos, sin), apply_rotary(k, cos, sin)
            self.k[li] = k if self.k[li] is None else torch.cat([self.k[li], k], dim=1)
            self.v[li] = v if self.v[li] is None else torch.cat([self.v[li], v], dim=1)
            T = p0 + n
            g = nh // kv
            kk = self.k[li].unsqueeze(1).expand(kv, g, T, hd).reshape(nh, T, hd)     # gqa_expand order
            vv = self.v[li].unsqueeze(1).expand(kv, g, T, hd).reshape(nh, T, hd)
            scores = (q @ kk.transpose(1, 2)) * scale
            qpos = torch.arange(p0, T).reshape(n, 1)
            scores = scores + torch.where(torch.arange(T).reshape(1, T) > qpos, torch.tensor(-1e9), torch.tensor(0.0))
            ctx = torch.softmax(scores, dim=-1) @ vv                                 # [nh, n, hd]
            attn = linear(ctx.transpose(0, 1).reshape(n, nh * hd), L["w_o"])
            x, residual = rms_norm(attn, L["ln2"], eps, residual)
            h = linear(silu_and_mul(linear(x, L["w_gu"])), L["w_down"])
        self.len += n
        last, _ = rms_norm(h[-1:], M.norm, eps, residual[-1:])
        return (last @ M.lm_head)[0]


def argmax_last(logits_row) -> int:
    """src/engine/llm_engine.rs:135-142: Iterator::max_by returns the LAST maximal element."""
    a = np.asarray(logits_row, dtype=np.float32)
    mx = a.max()
    return int(np.nonzero(a == mx)[0][-1])
