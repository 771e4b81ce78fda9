"""The reference's layer API (src/layers/*.rs) on the B200 kernels.  Same class names, constructor
arguments and method meaning; tensors are bf16 CUDA tensors used as device-memory handles, every
computation is a call through the C ABI (ops.Context).  No CPU path: a CPU tensor raises.

  RMSNorm            src/layers/layernorm.rs:3-61       forward(x, residual) -> (out, new_residual)
  RotaryEmbedding    src/layers/rotary_embedding.rs     apply(q, k) on [B, heads, T, hd], positions 0..T
  SiluAndMul         src/layers/activation.rs:4-18
  *ParallelLinear    src/layers/linear.rs:4-223         load_weights(weight, bias) / forward(x)
  Attention          src/layers/attention.rs:4-16       forward(q, k, v) + per-step Context
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from .tp import TPConfig, get_tp

BF16 = torch.bfloat16


class RMSNorm:
    def __init__(self, ctx, hidden_size: int, eps: Optional[float] = None, device="cuda"):
        self.ctx = ctx
        self.weight = torch.ones(hidden_size, dtype=BF16, device=device)      # layernorm.rs:29 (ones)
        self.eps = 1e-6 if eps is None else eps

    @classmethod
    def from_weight(cls, ctx, weight: torch.Tensor, eps: float) -> "RMSNorm":
        self = cls.__new__(cls)
        self.ctx, self.weight, self.eps = ctx, weight.reshape(-1).to(BF16).contiguous(), eps
        return self

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None):
        """(x + residual) normalised; returns (out bf16, new_residual f32 or None) like
        layernorm.rs:44-60 (the sum is kept in f32)."""
        shape = x.shape
        x2 = x.reshape(-1, shape[-1]).contiguous()
        if residual is None:
            return self.ctx.fused_add_rmsnorm(x2, self.weight, self.eps).reshape(shape), None
        r2 = residual.reshape(-1, shape[-1]).to(torch.float32).contiguous()
        new_res = torch.empty_like(r2)
        out = self.ctx.fused_add_rmsnorm(x2, self.weight, self.eps, residual_in=r2, residual_out=new_res)
        return out.reshape(shape), new_res.reshape(shape)


class RotaryEmbedding:
    def __init__(self, ctx, head_size: int, max_position: int, base: float):
        if head_size != 128:
            raise ValueError("the B200 kernels support head_size 128 (all Qwen3 dense models)")
        self.ctx, self.head_size, self.max_position, self.base = ctx, head_size, max_position, base
        self.table = ctx.rope_table(base, head_size, max_position)      # built once (reference: per call)

    def apply(self, q: torch.Tensor, k: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """q [B, nh, T, hd], k [B, kv, T, hd]; position of token t is t (rotary_embedding.rs:93-107)."""
        B, nh, T, hd = q.shape
        kv = k.shape[1]
        if T > self.max_position:
            raise ValueError("sequence longer than max_position")
        # plumbing only: token-major fused rows [q heads | k heads | (unused v heads)]
        rows = torch.cat([q.permute(0, 2, 1, 3).reshape(B * T, nh * hd), k.permute(0, 2, 1, 3).reshape(B * T, kv * hd),
                          torch.zeros(B * T, kv * hd, dtype=BF16, device=q.device)], dim=1).contiguous()
        pos = torch.arange(T, dtype=torch.int32, device=q.device).repeat(B)
        self.ctx.qk_norm_rope_store(rows, None, None, 0.0, self.table, pos, nh, kv, hd)
        qo = rows[:, :nh * hd].reshape(B, T, nh, hd).permute(0, 2, 1, 3)
        ko = rows[:, nh * hd:(nh + kv) * hd].reshape(B, T, kv, hd).permute(0, 2, 1, 3)
        return qo, ko


class SiluAndMul:
    def __init__(self, ctx):
        self.ctx = ctx

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        shape = x.shape
        out = self.ctx.silu_and_mul(x.reshape(-1, shape[-1]).contiguous())
        return out.reshape(*shape[:-1], shape[-1] // 2)


# ----------------------------------------------------------------------------- linear family
class LinearBase:
    """weight [out, in] (the reference's layout, linear.rs:16)."""

    def __init__(self, ctx, tp_dim: int, input_size: int, output_size: int, tp: Optional[TPConfig] = None):
        self.ctx = ctx
        tp = tp or TPConfig(ctx.tp_size, ctx.tp_rank, tp_dim)
        self.tp_dim, self.tp_rank, self.tp_size = tp_dim, tp.rank, tp.size
        self.input_size, self.output_size = input_size, output_size
        self.weight: Optional[torch.Tensor] = None

    def _set(self, w: torch.Tensor, bias):
        if bias is not None:
            raise NotImplementedError("bias: the Qwen3 path has none (every call site passes None, qwen3.rs:121,123)")
        self.weight = w.to(BF16).contiguous()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.weight is None:
            raise RuntimeError("load_weights has not been called")
        shape = x.shape
        y = self.ctx.gemm(x.reshape(-1, shape[-1]), self.weight)
        return y.reshape(*shape[:-1], self.weight.shape[0])


class ReplicatedLinear(LinearBase):
    def __init__(self, ctx, input_size, output_size):
        super().__init__(ctx, 0, input_size, output_size)

    def load_weights(self, weight, bias=None):
        self._set(weight, bias)


class ColumnParallelLinear(LinearBase):
    """Output features sharded: rank r keeps rows [r*out/tp, (r+1)*out/tp) of the FULL weight.
    (The reference computes shard_size from the already-full dim, linear.rs:81-84, which only
    works at tp_size 1.)"""

    def __init__(self, ctx, input_size, output_size, tp: Optional[TPConfig] = None):
        super().__init__(ctx, 0, input_size, output_size, tp)

    def load_weights(self, weight, bias=None):
        n = weight.shape[0] // self.tp_size
        self._set(weight[self.tp_rank * n:(self.tp_rank + 1) * n], bias)


class MergedColumnParallelLinear(LinearBase):
    """Several column-parallel matrices fused along the output dim (gate_up): each is sharded
    separately, linear.rs:106-119."""

    def __init__(self, ctx, input_size, output_sizes: List[int], tp: Optional[TPConfig] = None):
        super().__init__(ctx, 0, input_size, sum(output_sizes), tp)
        self.output_sizes = output_sizes
        self._parts: List[Optional[torch.Tensor]] = [None] * len(output_sizes)

    def load_weights(self, weight, bias=None, loaded_shard_id: int = 0):
        n = self.output_sizes[loaded_shard_id] // self.tp_size
        self._parts[loaded_shard_id] = weight[self.tp_rank * n:(self.tp_rank + 1) * n]
        if all(p is not None for p in self._parts):
            self._set(torch.cat(self._parts, dim=0), bias)


class QKVParallelLinear(LinearBase):
    """Fused q/k/v rows, sharded by head group (linear.rs:121-176)."""

    def __init__(self, ctx, hidden_size, head_size, num_heads, num_kv_heads, tp: Optional[TPConfig] = None):
        super().__init__(ctx, 0, hidden_size, (num_heads + 2 * num_kv_heads) * head_size, tp)
        self.head_size, self.num_heads, self.num_kv_heads = head_size, num_heads, num_kv_heads
        self._parts = {}

    def load_qkv_weight(self, w, shard_id: str):
        heads = self.num_heads if shard_id == "q" else self.num_kv_heads
        n = heads // self.tp_size * self.head_size
        self._parts[shard_id] = w[self.tp_rank * n:(self.tp_rank + 1) * n]
        if len(self._parts) == 3:
            self._set(torch.cat([self._parts["q"], self._parts["k"], self._parts["v"]], dim=0), None)

    def load_weights(self, weight, bias=None):
        """Full fused [q; k; v] weight (qwen3.rs:171-173)."""
        hs = self.head_size
        q, k, v = weight.split([self.num_heads * hs, self.num_kv_heads * hs, self.num_kv_heads * hs], dim=0)
        for name, w in (("q", q), ("k", k), ("v", v)):
            self.load_qkv_weight(w, name)


class RowParallelLinear(LinearBase):
    """Input features sharded (columns of the [out, in] weight); partial outputs are summed with
    an all-reduce - the reference stops before the reduce (linear.rs:184-198) and shards dim 0
    (:217), which is wrong for a row-parallel layer."""

    def __init__(self, ctx, input_size, output_size, tp: Optional[TPConfig] = None):
        super().__init__(ctx, 1, input_size, output_size, tp)

    def load_weights(self, weight, bias=None):
        n = weight.shape[1] // self.tp_size
        self._set(weight[:, self.tp_rank * n:(self.tp_rank + 1) * n], bias)

    def forward(self, x):
        y = super().forward(x)
        if self.tp_size > 1:
            self.ctx.allreduce(y)
        return y


# ----------------------------------------------------------------------------- attention
@dataclass
class Context:
    """Per-step attention inputs (upstream nano-vllm's Context): the integer contract N1/N2."""
    is_prefill: bool = False
    cu_seqlens: Optional[torch.Tensor] = None      # i32 [B+1]
    max_seqlen: int = 0
    slot_mapping: Optional[torch.Tensor] = None    # i32 [N]
    context_lens: Optional[torch.Tensor] = None    # i32 [B]
    block_tables: Optional[torch.Tensor] = None    # i32 [B, max_blocks]


_CONTEXT = Context()


def set_context(**kw) -> None:
    global _CONTEXT
    _CONTEXT = Context(**kw)


def get_context() -> Context:
    return _CONTEXT


class Attention:
    """Attention{num_heads, head_dim, scale, softcapping} (attention.rs:4-9) + num_kv_heads and the
    paged KV cache.  forward(q, k, v): q [N, nh*hd], k / v [N, kv*hd] token-major (views into a
    fused qkv buffer are fine); writes k, v to the cache at slot_mapping, then varlen causal flash
    (prefill) or paged attention over block_tables (decode)."""

    def __init__(self, ctx, num_heads: int, head_dim: int, scale: float, num_kv_heads: int, softcapping: float = 1.0,
                 block_size: int = 256):
        if softcapping != 1.0:
            raise NotImplementedError("softcapping is unused by Qwen3")
        self.ctx, self.num_heads, self.head_dim, self.scale = ctx, num_heads, head_dim, scale
        self.num_kv_heads, self.block_size = num_kv_heads, block_size
        self.k_cache: Optional[torch.Tensor] = None      # [num_blocks, block_size, kv, hd]
        self.v_cache: Optional[torch.Tensor] = None

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        c = get_context()
        nh, kv, hd = self.num_heads, self.num_kv_heads, self.head_dim
        if self.k_cache is not None and c.slot_mapping is not None:
            self.ctx.store_kvcache(k, v, self.k_cache, self.v_cache, c.slot_mapping, kv, hd)
        if c.is_prefill:
            return self.ctx.prefill_attn(q, k, v, c.cu_seqlens, c.max_seqlen, nh, kv, hd, self.scale)
        if self.k_cache is None:
            raise RuntimeError("decode needs a KV cache")
        return self.ctx.paged_decode_attn(q, self.k_cache, self.v_cache, c.block_tables, c.context_lens, nh, kv, hd,
                                          self.block_size, self.scale)
