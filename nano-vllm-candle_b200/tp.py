"""Tensor-parallel configuration and sharding rules.

TPConfig mirrors src/tp.rs:3-70 (size / rank / dim from TP_SIZE / TP_RANK, shard_size,
shard_offset).  The reference stops there: its linears record rank / size but never shard
correctly and never communicate (src/layers/linear.rs:81-84,184-198,215-217).  The rules below are
the ones the C++ runner applies when it slices a full checkpoint (csrc/runner.cu,
nvc_runner_load_weight); they are kept here in executable form for the CPU tests."""
from __future__ import annotations

import os
from dataclasses import dataclass, replace
from typing import Dict, Mapping, Sequence, Tuple


@dataclass(frozen=True)
class TPConfig:
    size: int = 1
    rank: int = 0
    dim: int = 0

    @staticmethod
    def from_env(env: Mapping[str, str] = None) -> "TPConfig":
        env = os.environ if env is None else env

        def _p(k):
            try:
                return int(env[k])
            except (KeyError, ValueError):
                return None
        size = _p("TP_SIZE")
        size = 1 if size is None else size
        rank = _p("TP_RANK")
        rank = 0 if rank is None or rank >= size else rank
        return TPConfig(size, rank, 0)

    @staticmethod
    def single() -> "TPConfig":
        return TPConfig(1, 0, 0)

    def with_size(self, size):
        return replace(self, size=size)

    def with_rank(self, rank):
        return replace(self, rank=rank)

    def with_dim(self, dim):
        return replace(self, dim=dim)

    def is_distributed(self) -> bool:
        return self.size > 1

    def shard_size(self, total: int) -> int:
        return total // self.size

    def shard_offset(self, total: int) -> int:
        return self.rank * self.shard_size(total)


def get_tp() -> TPConfig:
    return TPConfig.from_env()


def _rows(c: TPConfig, w):
    n = c.shard_size(w.shape[0])
    return w[c.rank * n:(c.rank + 1) * n]


def _cols(c: TPConfig, w):
    n = c.shard_size(w.shape[1])
    return w[:, c.rank * n:(c.rank + 1) * n]


def shard_qwen3_layer(c: TPConfig, w: Dict[str, "torch.Tensor"], nh: int, kv: int, hd: int) -> Dict[str, "torch.Tensor"]:
    """Column-parallel q/k/v/gate/up: contiguous head (resp. column) groups, i.e. rows of the
    [out, in] weight; row-parallel o_proj / down_proj: the matching INPUT columns."""
    assert nh % c.size == 0 and kv % c.size == 0
    return dict(q=_rows(c, w["q"]), k=_rows(c, w["k"]), v=_rows(c, w["v"]), o=_cols(c, w["o"]),
                gate=_rows(c, w["gate"]), up=_rows(c, w["up"]), down=_cols(c, w["down"]))


def merge_argmax(pairs: Sequence[Tuple[float, int]]) -> int:
    """(max, idx) of every vocab shard -> global argmax, ties -> the larger index (the
    reference's last-maximum rule, src/engine/llm_engine.rs:135-142)."""
    best_v, best_i = pairs[0]
    for v, i in pairs[1:]:
        if v > best_v or (v == best_v and i > best_i):
            best_v, best_i = v, i
    return best_i
