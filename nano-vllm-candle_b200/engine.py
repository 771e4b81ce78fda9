"""Python face of the host side (C++ behind include/nvc_b200_engine.h), with the reference's
names and argument meaning so callers of nano-vllm-candle can switch over:

  SamplingParams      src/sampling_params.rs:1-46
  SchedulerConfig     src/engine/scheduler.rs:10-56     (builder methods kept)
  Scheduler           src/engine/scheduler.rs:58-250    (state lives in the C++ engine)
  DummyModelRunner    src/engine/llm_engine.rs:20-33
  Qwen3ModelRunner    src/engine/llm_engine.rs:35-189   (from_hf_dir)
  LLMEngine           src/engine/llm_engine.rs:209-326  (add_request / step / is_finished / generate)
  GenerationOutput    src/engine/llm_engine.rs:192-207
  StepOutput          src/engine/llm_engine.rs:328-333
"""
from __future__ import annotations

import ctypes as C
import json
import os
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence as Seq, Tuple

import numpy as np

from . import _lib
from ._engine_sigs import (Qwen3ConfigC, RunnerOptionsC, ScheduleInfoC, SchedulerConfigC, StepOutputC)
from ._lib import NvcError, check


# ----------------------------------------------------------------------------- params / config
@dataclass
class SamplingParams:
    temperature: float = 1.0
    max_tokens: int = 64
    ignore_eos: bool = False

    def __post_init__(self):
        if not self.temperature > 1e-10:          # sampling_params.rs:20,30
            raise ValueError("greedy sampling is not permitted")

    def with_temperature(self, t: float) -> "SamplingParams":
        return SamplingParams(t, self.max_tokens, self.ignore_eos)

    def with_max_tokens(self, n: int) -> "SamplingParams":
        return SamplingParams(self.temperature, n, self.ignore_eos)

    def with_ignore_eos(self, b: bool) -> "SamplingParams":
        return SamplingParams(self.temperature, self.max_tokens, b)


@dataclass
class SchedulerConfig:
    max_num_seqs: int = 1
    max_num_batched_tokens: int = 4096
    eos: int = 0
    num_kvcache_blocks: int = 0
    kvcache_block_size: int = 256
    enable_prefix_caching: bool = False      # beyond the reference; off = reference behaviour

    enable_chunked_prefill: bool = False     # beyond the reference: mixed prefill / decode steps
    max_model_len: int = 0                   # 0 = the runner's: a sequence that reaches it finishes, a longer prompt is refused

    def with_chunked_prefill(self, on: bool = True):
        self.enable_chunked_prefill = on
        return self

    def with_prefix_caching(self, on: bool = True):
        self.enable_prefix_caching = on
        return self

    def with_max_num_seqs(self, n):
        self.max_num_seqs = n
        return self

    def with_max_num_batched_tokens(self, n):
        self.max_num_batched_tokens = n
        return self

    def with_eos(self, eos):
        self.eos = eos
        return self

    def with_kvcache(self, num_blocks, block_size):
        self.num_kvcache_blocks, self.kvcache_block_size = num_blocks, block_size
        return self


class Scheduler:
    """Scheduler::new(config).  The queues and the block manager live in the C++ engine that
    LLMEngine creates from this config."""

    def __init__(self, config: SchedulerConfig):
        self.config = config


@dataclass
class Qwen3Config:
    """Qwen3Config (src/models/qwen3.rs:19-34) / from_hf_dir (:77-101)."""
    vocab_size: int
    hidden_size: int
    head_dim: int
    num_hidden_layers: int
    num_attention_heads: int
    num_key_value_heads: int
    intermediate_size: int
    max_position_embeddings: int
    rms_norm_eps: float = 1e-6
    hidden_act: str = "silu"
    rope_theta: float = 1e6
    bos_token_id: int = 151643
    eos_token_id: int = 151645
    tie_word_embeddings: bool = False

    @staticmethod
    def from_hf_dir(model_dir: str) -> "Qwen3Config":
        """Qwen3Config::from_hf_dir (qwen3.rs:77-101), parsed by the native loader."""
        l = _lib.lib()
        c, tie = Qwen3ConfigC(), C.c_int()
        st = l.nvc_qwen3_config_from_hf_dir(os.fsencode(model_dir), C.byref(c), C.byref(tie))
        if st != 0:
            raise NvcError(st, "nvc_qwen3_config_from_hf_dir", l.nvc_hf_last_error().decode())
        cfg = Qwen3Config(**{n: getattr(c, n) for n in Qwen3Config.__dataclass_fields__ if n not in ("hidden_act", "tie_word_embeddings")})
        cfg.tie_word_embeddings = bool(tie.value)
        return cfg

    @staticmethod
    def from_any(cfg) -> "Qwen3Config":
        names = [f for f in Qwen3Config.__dataclass_fields__]
        return Qwen3Config(**{n: getattr(cfg, n) for n in names if hasattr(cfg, n)})

    def to_c(self) -> Qwen3ConfigC:
        return Qwen3ConfigC(self.vocab_size, self.hidden_size, self.head_dim, self.num_hidden_layers,
                            self.num_attention_heads, self.num_key_value_heads, self.intermediate_size,
                            self.max_position_embeddings, self.rms_norm_eps, self.rope_theta,
                            self.bos_token_id, self.eos_token_id)


def _cfg(H, L, nh, kv, I, V=151936, maxpos=40960, bos=151643, eos=151645) -> Qwen3Config:
    return Qwen3Config(vocab_size=V, hidden_size=H, head_dim=128, num_hidden_layers=L, num_attention_heads=nh,
                       num_key_value_heads=kv, intermediate_size=I, max_position_embeddings=maxpos,
                       bos_token_id=bos, eos_token_id=eos)


# Published Qwen3 dense shapes + a toy shape for fast tests.
MODEL_CONFIGS: Dict[str, Qwen3Config] = {
    "qwen3-0.6b": _cfg(1024, 28, 16, 8, 3072),
    "qwen3-8b": _cfg(4096, 36, 32, 8, 12288),
    "qwen3-14b": _cfg(5120, 40, 40, 8, 17408),
    "qwen3-32b": _cfg(5120, 64, 64, 8, 25600),
    "qwen3-tiny": _cfg(256, 3, 4, 2, 512, V=2048, maxpos=4096, bos=1, eos=2),
}


# ----------------------------------------------------------------------------- runners
class DummyModelRunner:
    """next token = last_token + 1 (implemented inside the C++ engine when no runner is given)."""
    handle = None


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Qwen3ModelRunner:
    """Qwen3ModelRunner (src/engine/llm_engine.rs:35-189) on the B200 step executor.

    Greedy by default: a step returns argmax with the reference's tie rule (last maximum).  The
    reference's temperature path draws from an unseeded RNG (llm_engine.rs:128) and cannot be
    reproduced; `set_sampling` / `LLMEngine.set_sampling` arm the seeded on-device sampler instead
    (Gumbel-max on Philox noise).  Steps: `prefill` (whole prompts), `decode` (one token
    per sequence, CUDA graph), `step_mixed` (chunked prefill: decode rows + prompt chunks)."""

    def __init__(self, ctx, cfg: Qwen3Config, num_kvcache_blocks: int, kvcache_block_size: int = 256,
                 max_num_seqs: int = 64, max_num_batched_tokens: int = 4096, max_model_len: Optional[int] = None,
                 use_cuda_graph: bool = True):
        self.l = _lib.lib()
        self.ctx = ctx
        self.cfg = cfg
        self.eos_id = cfg.eos_token_id
        self.block_size = kvcache_block_size
        self.num_kvcache_blocks = num_kvcache_blocks
        max_model_len = max_model_len or min(cfg.max_position_embeddings, num_kvcache_blocks * kvcache_block_size)
        self.max_model_len = max_model_len
        self.max_num_seqs, self.max_num_batched_tokens = max_num_seqs, max_num_batched_tokens
        opt = RunnerOptionsC(num_kvcache_blocks, kvcache_block_size, max_num_seqs, max_num_batched_tokens,
                             max_model_len, 1 if use_cuda_graph else 0, 0)
        c = cfg.to_c()
        h = C.c_void_p()
        st = self.l.nvc_runner_create(ctx.h, C.byref(c), C.byref(opt), C.byref(h))
        self.handle = h
        if st != 0:
            msg = self.l.nvc_last_error(ctx.h).decode()
            self.close()
            raise NvcError(st, "nvc_runner_create", msg)
        self._final = False

    # -- construction -------------------------------------------------------------------------
    @classmethod
    def from_weights(cls, ctx, cfg, weights: Iterable[Tuple[str, "torch.Tensor"]], **kw) -> "Qwen3ModelRunner":
        """weights: (HF name, CPU tensor) pairs of the FULL model; this rank's shard is sliced
        inside the library."""
        r = cls(ctx, Qwen3Config.from_any(cfg), **kw)
        for name, t in weights:
            r.load_weight(name, t)
        r.finalize()
        return r

    @classmethod
    def from_hf_dir(cls, ctx, model_dir: str, **kw) -> "Qwen3ModelRunner":
        """Qwen3ModelRunner::from_hf_dir (llm_engine.rs:43-58): config.json + model.safetensors
        (qwen3.rs:515-536), read by the native loader (mmap, sharded index, bf16 / f16 / f32, tied
        embeddings) - no Python tensor library on this path."""
        cfg = Qwen3Config.from_hf_dir(model_dir)
        r = cls(ctx, cfg, **kw)
        st = r.l.nvc_runner_load_hf_dir(r.handle, os.fsencode(model_dir))
        if st != 0:
            msg = r.l.nvc_hf_last_error().decode() + " | " + r.l.nvc_last_error(ctx.h).decode()
            r.close()
            raise NvcError(st, "nvc_runner_load_hf_dir", msg)
        r._final = True
        return r

    def load_weight(self, name: str, t) -> None:
        import torch
        t = t.detach().to(torch.bfloat16).contiguous()          # host or device memory, both accepted
        rows = t.shape[0]
        cols = t.shape[1] if t.dim() == 2 else 1
        st = self.l.nvc_runner_load_weight(self.handle, name.encode(), C.c_void_p(t.data_ptr()), rows, cols)
        check(st, f"nvc_runner_load_weight({name})", self.ctx.h)

    def finalize(self) -> None:
        check(self.l.nvc_runner_finalize(self.handle), "nvc_runner_finalize", self.ctx.h)
        self._final = True

    def close(self):
        if getattr(self, "handle", None):
            self.l.nvc_runner_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- steps ----------------------------------------------------------------------------------
    def prefill(self, ids, positions, cu_seqlens, slot_mapping) -> np.ndarray:
        ids, positions, cu, slots = _i32(ids), _i32(positions), _i32(cu_seqlens), _i32(slot_mapping)
        B, N = len(cu) - 1, len(ids)
        max_seqlen = int(np.max(np.diff(cu))) if B else 0
        out = np.empty(B, dtype=np.int32)
        st = self.l.nvc_runner_prefill(self.handle, _p(ids), _p(positions), _p(cu), _p(slots), B, N, max_seqlen, _p(out))
        check(st, "nvc_runner_prefill", self.ctx.h)
        return out

    def decode(self, ids, positions, slot_mapping, block_tables, context_lens) -> np.ndarray:
        ids, positions, slots, ctxl = _i32(ids), _i32(positions), _i32(slot_mapping), _i32(context_lens)
        bt = _i32(block_tables)
        B = len(ids)
        out = np.empty(B, dtype=np.int32)
        st = self.l.nvc_runner_decode(self.handle, _p(ids), _p(positions), _p(slots), _p(bt), bt.shape[1], _p(ctxl), B, _p(out))
        check(st, "nvc_runner_decode", self.ctx.h)
        return out

    def step_mixed(self, ids, positions, cu_seqlens, slot_mapping, block_tables, context_lens, num_decode: int) -> np.ndarray:
        """One chunked-prefill step: the first num_decode sequences have one query token (paged
        decode attention), the others are prompt chunks (paged prefill attention)."""
        ids, positions, cu, slots, ctxl = _i32(ids), _i32(positions), _i32(cu_seqlens), _i32(slot_mapping), _i32(context_lens)
        bt = _i32(block_tables)
        B, N = len(cu) - 1, len(ids)
        out = np.empty(B, dtype=np.int32)
        st = self.l.nvc_runner_step_mixed(self.handle, _p(ids), _p(positions), _p(cu), _p(slots), _p(bt), bt.shape[1],
                                          _p(ctxl), num_decode, B, N, int(np.max(np.diff(cu))) if B else 0, _p(out))
        check(st, "nvc_runner_step_mixed", self.ctx.h)
        return out

    def set_sampling(self, temperatures, streams, seed: int) -> None:
        """Arms on-device temperature sampling for the next step (None disarms -> greedy)."""
        if temperatures is None:
            check(self.l.nvc_runner_set_sampling(self.handle, None, None, 0, 0), "nvc_runner_set_sampling", self.ctx.h)
            return
        t = np.ascontiguousarray(np.asarray(temperatures, dtype=np.float32))
        s = np.ascontiguousarray(np.asarray(streams, dtype=np.uint64))
        check(self.l.nvc_runner_set_sampling(self.handle, _p(t), _p(s), len(t), C.c_uint64(seed & (2 ** 64 - 1))),
              "nvc_runner_set_sampling", self.ctx.h)

    def last_logits(self, B: int) -> np.ndarray:
        """fp32 logits [B, vocab / tp_size] of this rank's vocab shard for the last step."""
        v_l = self.cfg.vocab_size // self.ctx.tp_size
        out = np.empty((B, v_l), dtype=np.float32)
        check(self.l.nvc_runner_last_logits(self.handle, _p(out), B), "nvc_runner_last_logits", self.ctx.h)
        return out

    def time_decode(self, iters: int) -> Tuple[float, int]:
        ms, n = C.c_float(), C.c_int()
        check(self.l.nvc_runner_time_decode(self.handle, iters, C.byref(ms), C.byref(n)), "nvc_runner_time_decode", self.ctx.h)
        return ms.value, n.value

    OPS = {"paged_decode": 0, "gemm_qkv": 1, "gemm_o": 2, "gemm_gate_up": 3, "gemm_down": 4, "lm_head": 5,
           "rmsnorm": 6, "qk_norm_rope": 7, "silu_mul": 8, "argmax": 9, "noop": 10, "allreduce": 11}

    def time_op(self, op: str, iters: int) -> Tuple[float, int]:
        """(average ms per call, kernels launched) of one kernel family over all layers."""
        ms, n = C.c_float(), C.c_int()
        check(self.l.nvc_runner_time_op(self.handle, self.OPS[op], iters, C.byref(ms), C.byref(n)),
              "nvc_runner_time_op", self.ctx.h)
        return ms.value, n.value

    def weight_bytes(self) -> int:
        return self.l.nvc_runner_weight_bytes(self.handle)

    def h2d_bytes_last_step(self) -> int:
        """Host-to-device bytes the last step copied (counted where the copies are issued)."""
        return self.l.nvc_runner_h2d_bytes_last_step(self.handle)

    def kv_bytes_per_token(self) -> int:
        return self.l.nvc_runner_kv_bytes_per_token(self.handle)


# ----------------------------------------------------------------------------- engine
@dataclass
class GenerationOutput:
    seq_id: int
    token_ids: List[int]
    text: Optional[str] = None

    def with_text(self, text: str) -> "GenerationOutput":
        self.text = text
        return self


@dataclass
class StepOutput:
    outputs: List[Tuple[int, List[int]]]
    is_prefill: bool
    num_tokens: int
    num_query_tokens: int = 0      # tokens computed in the step
    num_decode_seqs: int = 0       # one-token sequences of the step
    num_sampled: int = 0           # sequences that received a token


@dataclass
class StepInputs:
    is_prefill: bool
    seq_ids: np.ndarray
    ids: np.ndarray
    positions: np.ndarray
    cu_seqlens: np.ndarray
    slot_mapping: np.ndarray
    block_tables: np.ndarray
    context_lens: np.ndarray
    num_decode_seqs: int = 0                  # chunked mode: leading one-token sequences
    emit: Optional[np.ndarray] = None         # chunked mode: which sequences take the sampled token


class LLMEngine:
    """LLMEngine<R: ModelRunner> (src/engine/llm_engine.rs:209-326).

    Like the reference, a step DRAWS the next token from softmax(logits / temperature) with each
    request's SamplingParams.temperature (llm_engine.rs:97-133) - here on the device and seeded
    (`seed`), so a run can be reproduced, which the reference's unseeded RNG cannot.  `greedy=True` is the
    explicit opt-in to argmax decoding (ties -> last index, llm_engine.rs:135-142), the mode the parity
    tests and the benchmark use; temperatures are then ignored by construction, not silently."""

    def __init__(self, scheduler: Scheduler, model_runner, greedy: bool = False, seed: int = 0):
        self.l = _lib.lib()
        self.scheduler = scheduler
        self.model_runner = model_runner
        cfg = scheduler.config
        c = SchedulerConfigC(cfg.max_num_seqs, cfg.max_num_batched_tokens, cfg.eos, cfg.num_kvcache_blocks,
                             cfg.kvcache_block_size, int(bool(cfg.enable_prefix_caching)),
                             int(bool(cfg.enable_chunked_prefill)), int(cfg.max_model_len))
        self.chunked = bool(cfg.enable_chunked_prefill)
        h = C.c_void_p()
        rh = getattr(model_runner, "handle", None)
        st = self.l.nvc_engine_create(C.byref(c), rh, C.byref(h))
        if st != 0:
            raise NvcError(st, "nvc_engine_create", "scheduler config refused (sizes must be positive, block size >= 2; with a "
                           "runner: kvcache_block_size equal to the runner's and num_kvcache_blocks <= the runner's pool)")
        self.h = h
        self.greedy = bool(greedy)
        if rh is not None and not greedy:
            self.set_sampling(True, seed)
        self._live: List[int] = []          # unfinished seq ids, to report outputs per step
        self.prefill_tokens = 0
        self.decode_tokens = 0

    def close(self):
        if getattr(self, "h", None):
            self.l.nvc_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, st, where):
        if st != 0:
            raise NvcError(st, where, self.l.nvc_engine_last_error(self.h).decode() or
                           (self.l.nvc_last_error(self.model_runner.ctx.h).decode()
                            if getattr(self.model_runner, "ctx", None) else ""))

    def add_request(self, token_ids: Seq[int], sampling_params: SamplingParams) -> int:
        a = _i32(token_ids)
        sid = self.l.nvc_engine_add_request(self.h, _p(a), len(a), float(sampling_params.temperature),
                                            int(sampling_params.max_tokens), int(sampling_params.ignore_eos))
        if sid < 0:
            self._ck(int(sid), "nvc_engine_add_request")
        self._live.append(int(sid))
        return int(sid)

    def is_finished(self) -> bool:
        return bool(self.l.nvc_engine_is_finished(self.h))

    def set_sampling(self, enabled: bool, seed: int = 0) -> None:
        """On-device temperature sampling with each request's SamplingParams.temperature
        (seeded, reproducible per sequence); off = greedy argmax (see the class docstring)."""
        self._ck(self.l.nvc_engine_set_sampling(self.h, int(enabled), C.c_uint64(seed & (2 ** 64 - 1))),
                 "nvc_engine_set_sampling")

    def output_of(self, seq_id: int) -> Tuple[List[int], bool]:
        n = C.c_int()
        self.l.nvc_engine_get_output(self.h, seq_id, None, 0, C.byref(n))
        buf = np.empty(max(n.value, 1), dtype=np.int32)
        st = self.l.nvc_engine_get_output(self.h, seq_id, _p(buf), len(buf), C.byref(n))
        return buf[:n.value].tolist(), st == 0

    def step(self) -> StepOutput:
        so = StepOutputC()
        self._ck(self.l.nvc_engine_step(self.h, C.byref(so)), "nvc_engine_step")
        outputs = []
        if so.num_finished:
            still = []
            for sid in self._live:
                toks, fin = self.output_of(sid)
                if fin:
                    outputs.append((sid, toks))
                else:
                    still.append(sid)
            self._live = still
        return StepOutput(outputs, bool(so.is_prefill), so.num_tokens, so.num_query_tokens, so.num_decode_seqs,
                          so.num_sampled)

    def generate(self, prompts: Seq[Seq[int]], sampling_params: SamplingParams) -> List[GenerationOutput]:
        for p in prompts:
            self.add_request(p, sampling_params)
        outs: Dict[int, List[int]] = {}
        while not self.is_finished():
            s = self.step()
            if self.chunked:     # a mixed step: prompt tokens computed + decode tokens, counted exactly
                self.prefill_tokens += s.num_query_tokens - s.num_decode_seqs
                self.decode_tokens += s.num_decode_seqs
            elif s.is_prefill:
                self.prefill_tokens += s.num_tokens
            else:
                self.decode_tokens += -s.num_tokens
            for sid, toks in s.outputs:
                outs[sid] = toks
        return [GenerationOutput(sid, outs[sid]) for sid in sorted(outs)]

    # -- pieces of step(), exposed for the contract tests and the benchmark ----------------------
    def schedule(self) -> StepInputs:
        info = ScheduleInfoC()
        self._ck(self.l.nvc_engine_schedule(self.h, C.byref(info)), "nvc_engine_schedule")
        ptrs = [C.POINTER(C.c_int32)() for _ in range(6)]
        sp = C.POINTER(C.c_int64)()
        self._ck(self.l.nvc_engine_step_inputs(self.h, *[C.byref(p) for p in ptrs], C.byref(sp)), "nvc_engine_step_inputs")
        B, N = info.num_seqs, info.num_tokens

        def arr(p, n):
            return np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.empty(0, np.int32)
        ep = C.POINTER(C.c_int32)()
        self._ck(self.l.nvc_engine_step_emit_mask(self.h, C.byref(ep)), "nvc_engine_step_emit_mask")
        return StepInputs(
            num_decode_seqs=info.num_decode_seqs, emit=arr(ep, B),
            is_prefill=bool(info.is_prefill), seq_ids=np.ctypeslib.as_array(sp, shape=(B,)).copy(),
            ids=arr(ptrs[0], N), positions=arr(ptrs[1], N),
            cu_seqlens=arr(ptrs[2], B + 1) if (info.is_prefill or self.chunked) else np.empty(0, np.int32),
            slot_mapping=arr(ptrs[3], N), block_tables=arr(ptrs[4], B * info.max_blocks).reshape(B, info.max_blocks),
            context_lens=arr(ptrs[5], B))

    def post_process(self, token_ids) -> None:
        a = _i32(token_ids)
        self._ck(self.l.nvc_engine_post_process(self.h, _p(a), len(a)), "nvc_engine_post_process")

    def preempt(self, seq_id: int) -> None:
        self._ck(self.l.nvc_engine_preempt(self.h, seq_id), "nvc_engine_preempt")

    def block_table_of(self, seq_id: int) -> List[int]:
        n = C.c_int()
        self.l.nvc_engine_seq_block_table(self.h, seq_id, None, 0, C.byref(n))
        buf = np.empty(max(n.value, 1), dtype=np.int32)
        self._ck(self.l.nvc_engine_seq_block_table(self.h, seq_id, _p(buf), len(buf), C.byref(n)), "nvc_engine_seq_block_table")
        return buf[:n.value].tolist()

    def num_free_blocks(self) -> int:
        return self.l.nvc_engine_num_free_blocks(self.h)

    def prefix_cache_stats(self) -> Tuple[int, int]:
        """(full blocks found in the prefix cache, full blocks looked up) since the engine was created."""
        hit, q = C.c_int64(), C.c_int64()
        self._ck(self.l.nvc_engine_prefix_cache_stats(self.h, C.byref(hit), C.byref(q)), "nvc_engine_prefix_cache_stats")
        return hit.value, q.value

    def num_waiting(self) -> int:
        return self.l.nvc_engine_num_waiting(self.h)

    def num_running(self) -> int:
        return self.l.nvc_engine_num_running(self.h)
