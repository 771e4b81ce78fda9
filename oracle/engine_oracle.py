"""CPU restatement of the reference's engine book-keeping and of the integer
contract (slot_mapping / block_tables / store_kvcache) -- TEST INFRASTRUCTURE.

Reference files followed:
  src/engine/sequence.rs:5-97, src/engine/block_manager.rs:3-99,
  src/engine/scheduler.rs:10-250, src/engine/llm_engine.rs:16-333,
  src/sampling_params.rs:1-46, src/tp.rs:3-70.

The reference BlockManager is an admitted stub (block_manager.rs:24-29: identity
block_table, always-true capacity).  `BlockManagerStub` restates it as written;
`BlockManager` is the real allocator the north_star contract needs (free list,
disjoint ids), following upstream nano-vllm's rule which the reference says its API
mirrors (block_manager.rs:26-29); `BlockManagerPrefix` adds upstream's hash-chained
prefix cache.
The slot_mapping rule (N1) is expressed only with the Sequence helpers the reference
exposes (sequence.rs:73-91).
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass
from typing import Callable, Deque, Dict, List, Optional, Tuple

import numpy as np


# ----------------------------------------------------------------- sampling params / tp
@dataclass
class SamplingParams:
    """src/sampling_params.rs:1-46."""
    temperature: float = 1.0
    max_tokens: int = 64
    ignore_eos: bool = False

    def __post_init__(self):
        assert self.temperature > 1e-10, "greedy sampling is not permitted"   # :20,30


@dataclass(frozen=True)
class TPConfig:
    """src/tp.rs:3-70."""
    size: int = 1
    rank: int = 0
    dim: int = 0

    @staticmethod
    def from_env(env) -> "TPConfig":
        def _p(k):
            try:
                return int(env[k])
            except Exception:
                return None
        size = _p("TP_SIZE")
        size = 1 if size is None else size                            # :22
        rank = _p("TP_RANK")
        rank = 0 if rank is None or rank >= size else rank          # :24-30
        return TPConfig(size, rank, 0)

    def is_distributed(self) -> bool:
        return self.size > 1

    def shard_size(self, total: int) -> int:
        return total // self.size                                   # :59-61

    def shard_offset(self, total: int) -> int:
        return self.rank * self.shard_size(total)                   # :63-65


# ----------------------------------------------------------------- sequence
WAITING, RUNNING, FINISHED = 0, 1, 2


class Sequence:
    """src/engine/sequence.rs:16-97.  block_size is hard-coded 256 in the reference (:35);
    here it is a constructor argument defaulting to 256 so the contract can be tested at
    other sizes."""
    _counter = 0

    def __init__(self, token_ids: List[int], sp: SamplingParams, block_size: int = 256):
        self.seq_id = Sequence._counter
        Sequence._counter += 1
        self.block_size = block_size
        self.block_table: List[int] = []
        self.status = WAITING
        self.token_ids = list(token_ids)
        self.last_token = token_ids[-1] if token_ids else 0
        self.num_tokens = len(token_ids)
        self.num_prompt_tokens = len(token_ids)
        self.num_cached_tokens = 0
        self.num_computed_tokens = 0      # chunked prefill only (SchedulerChunked): tokens whose K / V are stored
        self.temperature = sp.temperature
        self.max_tokens = sp.max_tokens
        self.ignore_eos = sp.ignore_eos

    def __len__(self):
        return self.num_tokens

    def is_finished(self):
        return self.status == FINISHED

    def num_completion_tokens(self):
        return self.num_tokens - self.num_prompt_tokens

    def prompt_token_ids(self):
        return self.token_ids[:self.num_prompt_tokens]

    def completion_token_ids(self):
        return self.token_ids[self.num_prompt_tokens:]

    def num_cached_blocks(self):
        return self.num_cached_tokens // self.block_size                 # :73-75

    def num_blocks(self):
        return -(-self.num_tokens // self.block_size)                    # :77-79

    def last_block_num_tokens(self):
        return self.num_tokens - (self.num_blocks() - 1) * self.block_size   # :81-83

    def block(self, i):
        assert i < self.num_blocks()
        return self.token_ids[i * self.block_size:min((i + 1) * self.block_size, self.num_tokens)]

    def append_token(self, token_id: int):
        self.token_ids.append(token_id)
        self.last_token = token_id
        self.num_tokens += 1


# ----------------------------------------------------------------- block managers
class BlockManagerStub:
    """src/engine/block_manager.rs:31-99 exactly as written."""

    def __init__(self, num_blocks=0, block_size=256):
        self.num_blocks_, self.block_size_ = num_blocks, block_size

    def block_size(self):
        return self.block_size_

    def num_blocks(self):
        return self.num_blocks_

    def can_allocate(self, seq):
        return True

    def allocate(self, seq):
        seq.block_table = list(range(seq.num_blocks()))       # :72-75
        seq.num_cached_tokens = len(seq)                      # :76

    def deallocate(self, seq):
        seq.block_table = []
        seq.num_cached_tokens = 0

    def can_append(self, seq):
        return True

    def may_append(self, seq):
        pass


class BlockManager:
    """Real allocator behind the same API.  Free list is a FIFO of block ids,
    initially 0..num_blocks-1; allocate pops from the front; deallocate pushes the
    sequence's blocks back in reverse table order (upstream nano-vllm's order)."""

    def __init__(self, num_blocks: int, block_size: int = 256):
        self.num_blocks_, self.block_size_ = num_blocks, block_size
        self.free: Deque[int] = deque(range(num_blocks))

    def block_size(self):
        return self.block_size_

    def num_blocks(self):
        return self.num_blocks_

    def num_free_blocks(self):
        return len(self.free)

    def can_allocate(self, seq):
        return len(self.free) >= seq.num_blocks()

    def allocate(self, seq):
        assert not seq.block_table
        seq.block_table = [self.free.popleft() for _ in range(seq.num_blocks())]
        seq.num_cached_tokens = 0          # no prefix cache: nothing is cached yet

    def deallocate(self, seq):
        for b in reversed(seq.block_table):
            self.free.append(b)
        seq.block_table = []
        seq.num_cached_tokens = 0

    def can_append(self, seq):
        # a block is missing only when the newest token opens one and the table does not hold it yet
        need = len(seq) % self.block_size_ == 1 and len(seq.block_table) < seq.num_blocks()
        return len(self.free) >= (1 if need else 0)

    def may_append(self, seq):
        if len(seq) % self.block_size_ == 1 and len(seq.block_table) < seq.num_blocks():
            seq.block_table.append(self.free.popleft())


class BlockManagerPrefix(BlockManager):
    """Upstream nano-vllm's BlockManager with prefix caching (the "block deduplication" that
    block_manager.rs:24-29 says the reference leaves out and will plug in behind the same API):
    ref-counted blocks, a chained hash per FULL block, hash -> block id map, a hit confirmed
    by comparing token ids; freed blocks keep their hash until they are handed out again.
    One rule is added: a block is taken from the cache only if (i + 1) * block_size < len(seq),
    so at least one token is left to compute (upstream would schedule an empty prefill).
    The hash key here is Python's hash of (prefix key, tokens): only equality is observable."""

    def __init__(self, num_blocks: int, block_size: int = 256):
        super().__init__(num_blocks, block_size)
        self.ref = [0] * num_blocks
        self.key: List[Optional[int]] = [None] * num_blocks
        self.tokens: List[List[int]] = [[] for _ in range(num_blocks)]
        self.key_to_block: Dict[int, int] = {}
        self.chunked = False                 # chunked prefill: a hit needs a block whose K / V are computed
        self.ready = [False] * num_blocks
        self.hits = 0
        self.queries = 0

    def _take(self, b):
        assert self.ref[b] == 0
        self.ref[b] = 1
        self.key[b] = None
        self.ready[b] = False
        self.tokens[b] = []
        self.free.remove(b)

    def _set_key(self, b, key, toks):
        self.key[b] = key
        self.tokens[b] = list(toks)
        self.key_to_block[key] = b

    def allocate(self, seq):
        assert not seq.block_table
        seq.num_cached_tokens = 0
        bs = self.block_size_
        key, miss = None, False
        for i in range(seq.num_blocks()):
            toks = seq.block(i)
            key = hash((key, tuple(toks))) if len(toks) == bs else None
            b = -1
            if key is not None:
                self.queries += 1
                b = self.key_to_block.get(key, -1)
            if b == -1 or self.key[b] != key or self.tokens[b] != toks or (i + 1) * bs >= len(seq) \
                    or (self.chunked and not self.ready[b]):
                miss = True
            if miss:
                b = self.free[0]
                self._take(b)
            else:
                self.hits += 1
                seq.num_cached_tokens += bs
                if self.ref[b] == 0:
                    was_ready = self.ready[b]      # a revived cached block still holds its K / V
                    self._take(b)
                    self.ready[b] = was_ready
                else:
                    self.ref[b] += 1
            if key is not None:
                self._set_key(b, key, toks)
            seq.block_table.append(b)

    def deallocate(self, seq):
        for b in reversed(seq.block_table):
            self.ref[b] -= 1
            if self.ref[b] == 0:
                self.free.append(b)
        seq.block_table = []
        seq.num_cached_tokens = 0

    def may_append(self, seq):
        bs = self.block_size_
        if len(seq) % bs == 1 and len(seq.block_table) < seq.num_blocks():
            b = self.free[0]
            self._take(b)
            seq.block_table.append(b)
        elif len(seq) % bs == 0 and seq.block_table:
            b = seq.block_table[-1]
            if self.key[b] is None:
                nb = len(seq.block_table)
                prefix = self.key[seq.block_table[-2]] if nb > 1 else None
                if nb == 1 or prefix is not None:
                    toks = seq.block(nb - 1)
                    self._set_key(b, hash((prefix, tuple(toks))), toks)


# ----------------------------------------------------------------- integer contract N1/N2/N3
def prefill_slot_mapping(seq: Sequence) -> List[int]:
    """N1 prefill: for block i in num_cached_blocks..num_blocks:
    slots block_table[i]*BS + [0,BS) (last block: [0,last_block_num_tokens))."""
    bs = seq.block_size
    out: List[int] = []
    for i in range(seq.num_cached_blocks(), seq.num_blocks()):
        start = seq.block_table[i] * bs
        n = bs if i != seq.num_blocks() - 1 else seq.last_block_num_tokens()
        out.extend(range(start, start + n))
    return out


def decode_slot(seq: Sequence) -> int:
    """N1 decode: block_table[last]*BS + last_block_num_tokens - 1."""
    return seq.block_table[-1] * seq.block_size + seq.last_block_num_tokens() - 1


def build_block_tables(seqs: List[Sequence]) -> np.ndarray:
    """N2: [B, max_blocks] int32, row i = seq.block_table padded with -1."""
    mb = max(len(s.block_table) for s in seqs)
    bt = np.full((len(seqs), mb), -1, dtype=np.int32)
    for i, s in enumerate(seqs):
        bt[i, :len(s.block_table)] = s.block_table
    return bt


def prepare_prefill(seqs: List[Sequence]):
    """Flat varlen step inputs (replacement for build_batch, llm_engine.rs:60-95)."""
    ids, pos, cu, slots = [], [], [0], []
    for s in seqs:
        start = s.num_cached_tokens
        ids.extend(s.token_ids[start:])
        pos.extend(range(start, len(s)))
        cu.append(cu[-1] + len(s) - start)
        slots.extend(prefill_slot_mapping(s))
    return (np.asarray(ids, np.int32), np.asarray(pos, np.int32),
            np.asarray(cu, np.int32), np.asarray(slots, np.int32))


def prepare_decode(seqs: List[Sequence]):
    ids = np.asarray([s.last_token for s in seqs], np.int32)
    pos = np.asarray([len(s) - 1 for s in seqs], np.int32)
    slots = np.asarray([decode_slot(s) for s in seqs], np.int32)
    ctx = np.asarray([len(s) for s in seqs], np.int32)
    return ids, pos, slots, ctx, build_block_tables(seqs)


def store_kvcache(k: np.ndarray, v: np.ndarray, k_cache: np.ndarray, v_cache: np.ndarray,
                  slot_mapping: np.ndarray) -> None:
    """N3: cache viewed as [num_slots, kv, hd]; k_cache[slot] = k[n]; slot -1 = skip."""
    kc = k_cache.reshape(-1, *k.shape[1:])
    vc = v_cache.reshape(-1, *v.shape[1:])
    for n, s in enumerate(slot_mapping.tolist()):
        if s < 0:
            continue
        kc[s] = k[n]
        vc[s] = v[n]


# ----------------------------------------------------------------- scheduler
@dataclass
class SchedulerConfig:
    """src/engine/scheduler.rs:10-56."""
    max_num_seqs: int = 1
    max_num_batched_tokens: int = 4096
    eos: int = 0
    num_kvcache_blocks: int = 0
    kvcache_block_size: int = 256


class Scheduler:
    """src/engine/scheduler.rs:58-250, statement by statement (including its quirks:
    max_num_seqs bounds one round only, :118; the budget adds len - num_cached_tokens, :134-138)."""

    def __init__(self, cfg: SchedulerConfig, block_manager=None):
        self.max_num_seqs = cfg.max_num_seqs
        self.max_num_batched_tokens = cfg.max_num_batched_tokens
        self.block_manager = block_manager if block_manager is not None else \
            BlockManagerStub(cfg.num_kvcache_blocks, cfg.kvcache_block_size)
        self.waiting: Deque[Sequence] = deque()
        self.running: Deque[Sequence] = deque()
        self.eos = cfg.eos

    def is_finished(self):
        return not self.waiting and not self.running

    def add(self, seq):
        self.waiting.append(seq)

    def schedule(self) -> Tuple[List[Sequence], bool]:
        r = self._try_schedule_prefill()
        if r is not None:
            return r
        return self._schedule_decode()

    def _try_schedule_prefill(self):                                   # :113-157
        scheduled, num_seqs, num_batched = [], 0, 0
        while self.waiting and num_seqs < self.max_num_seqs:
            seq = self.waiting[0]
            ok = (num_batched + len(seq) <= self.max_num_batched_tokens
                  and self.block_manager.can_allocate(seq))
            if not ok:
                break
            num_seqs += 1
            self.block_manager.allocate(seq)
            num_batched += len(seq) - seq.num_cached_tokens
            seq.status = RUNNING
            self.waiting.popleft()
            self.running.append(seq)
            scheduled.append(seq)
        if not scheduled:
            return None
        return scheduled, True

    def _schedule_decode(self):                                        # :159-186
        scheduled, num_seqs = [], 0
        while self.running and num_seqs < self.max_num_seqs:
            seq = self.running.popleft()
            if self._ensure_can_append(seq):
                num_seqs += 1
                self.block_manager.may_append(seq)
                scheduled.append(seq)
        assert scheduled, "decode stage should schedule at least one sequence"
        for seq in reversed(scheduled):
            self.running.appendleft(seq)
        return scheduled, False

    def _ensure_can_append(self, seq):                                 # :188-198
        while not self.block_manager.can_append(seq):
            if self.running:
                self.preempt(self.running.pop())
            else:
                self.preempt(seq)
                return False
        return True

    def preempt(self, seq):                                            # :200-211
        seq.status = WAITING
        self.block_manager.deallocate(seq)
        self.waiting.appendleft(seq)

    def post_process(self, seqs, token_ids):                           # :213-249
        for seq, tok in zip(seqs, token_ids):
            seq.append_token(tok)
            is_eos = tok == self.eos
            reached_max = seq.num_completion_tokens() >= seq.max_tokens
            if (not seq.ignore_eos and is_eos) or reached_max:
                seq.status = FINISHED
                self.block_manager.deallocate(seq)
                self.running = deque(s for s in self.running if s.seq_id != seq.seq_id)


class SchedulerChunked(Scheduler):
    """Chunked prefill with mixed prefill / decode steps.  The reference
    has no such mode (scheduler.rs:106-111 is prefill-step-or-decode-step); this restates the
    rule the C++ engine implements, written independently from the same specification:

      1. visit the running queue in order (as scheduler.rs:159-186 does).  A sequence with ONE
         pending token appends exactly as in decode (:166-170, victims preempted from the back of
         the queue, :188-198).  A sequence in the middle of its prompt reserves one budget token.
      2. the remaining budget goes to those prompt chunks in order,
      3. then to waiting sequences while len(running) < max_num_seqs (bound over ALL running
         sequences - the fix of :118) and blocks can be allocated; the budget counts the tokens
         actually computed (the fix of :134-138).  No admission in a step that preempted.
    Batch order: one-token sequences first, then chunks.  A sequence receives the sampled token
    only in the step that computes its last pending token."""

    def __init__(self, cfg: SchedulerConfig, block_manager):
        assert cfg.max_num_batched_tokens >= cfg.max_num_seqs
        super().__init__(cfg, block_manager)
        if hasattr(block_manager, "chunked"):
            block_manager.chunked = True

    def preempt(self, seq):
        super().preempt(seq)
        seq.num_computed_tokens = 0

    def schedule_chunked(self):
        """-> (seqs, q_lens, num_decode)."""
        ones, chunks, visited = [], [], []
        budget = self.max_num_batched_tokens
        preempted = False
        while self.running and len(visited) < self.max_num_seqs and budget > 0:
            seq = self.running.popleft()
            pending = len(seq) - getattr(seq, "num_computed_tokens", 0)
            if pending == 1:
                before = len(self.waiting)
                ok = self._ensure_can_append(seq)
                preempted |= len(self.waiting) != before
                if not ok:
                    continue
                self.block_manager.may_append(seq)
                ones.append(seq)
            else:
                chunks.append(seq)
            budget -= 1
            visited.append(seq)
        for seq in reversed(visited):
            self.running.appendleft(seq)
        chunk_q = []
        for seq in chunks:
            extra = min(len(seq) - seq.num_computed_tokens - 1, budget)
            chunk_q.append(1 + extra)
            budget -= extra
        while not preempted and self.waiting and len(self.running) < self.max_num_seqs and budget > 0:
            seq = self.waiting[0]
            if not self.block_manager.can_allocate(seq):
                break
            self.block_manager.allocate(seq)
            seq.num_computed_tokens = seq.num_cached_tokens
            q = min(len(seq) - seq.num_computed_tokens, budget)
            budget -= q
            seq.status = RUNNING
            self.waiting.popleft()
            self.running.append(seq)
            chunks.append(seq)
            chunk_q.append(q)
        seqs = ones + chunks
        assert seqs, "no sequence could be scheduled"
        return seqs, [1] * len(ones) + chunk_q, len(ones)

    def post_process_chunked(self, seqs, q_lens, token_ids):
        bm, bs = self.block_manager, self.block_manager.block_size_
        for seq, q, tok in zip(seqs, q_lens, token_ids):
            b0 = seq.num_computed_tokens // bs
            seq.num_computed_tokens += q
            if hasattr(bm, "ready"):
                for i in range(b0, min(seq.num_computed_tokens // bs, len(seq.block_table))):
                    bm.ready[seq.block_table[i]] = True
            if seq.num_computed_tokens == len(seq):
                self.post_process([seq], [tok])


def prepare_mixed(seqs: List[Sequence], q_lens: List[int]):
    """Step inputs of a chunked step: tokens [num_computed, num_computed + q) of every sequence;
    the slot of token t is N1's rule per token, block_table[t // BS] * BS + t % BS."""
    ids, pos, cu, slots, ctx, emit = [], [], [0], [], [], []
    for s, q in zip(seqs, q_lens):
        a = s.num_computed_tokens
        ids.extend(s.token_ids[a:a + q])
        pos.extend(range(a, a + q))
        slots.extend(s.block_table[t // s.block_size] * s.block_size + t % s.block_size for t in range(a, a + q))
        cu.append(cu[-1] + q)
        ctx.append(a + q)
        emit.append(int(a + q == len(s)))
    i32 = lambda x: np.asarray(x, np.int32)
    return i32(ids), i32(pos), i32(cu), i32(slots), i32(ctx), build_block_tables(seqs), i32(emit)


# ----------------------------------------------------------------- engine
class DummyModelRunner:
    """src/engine/llm_engine.rs:20-33: next token = last_token + 1."""

    def run(self, seqs, is_prefill):
        return [s.last_token + 1 for s in seqs]


class LLMEngine:
    """src/engine/llm_engine.rs:209-326."""

    def __init__(self, scheduler: Scheduler, model_runner, block_size: int = 256):
        self.scheduler = scheduler
        self.model_runner = model_runner
        self.block_size = block_size

    def add_request(self, token_ids, sp):
        self.scheduler.add(Sequence(token_ids, sp, self.block_size))

    def step(self):                                                    # :239-264
        seqs, is_prefill = self.scheduler.schedule()
        toks = self.model_runner.run(seqs, is_prefill)
        self.scheduler.post_process(seqs, toks)
        outputs = [(s.seq_id, s.completion_token_ids()) for s in seqs if s.is_finished()]
        num_tokens = sum(len(s) for s in seqs) if is_prefill else -len(seqs)
        return outputs, is_prefill, num_tokens

    def is_finished(self):
        return self.scheduler.is_finished()

    def generate(self, prompts, sp):                                   # :270-325
        for p in prompts:
            self.add_request(list(p), sp)
        outs: Dict[int, List[int]] = {}
        while not self.is_finished():
            o, _, _ = self.step()
            for sid, toks in o:
                outs[sid] = toks
        return [(sid, outs[sid]) for sid in sorted(outs)]


class OracleModelRunner:
    """Qwen3ModelRunner::run restated (src/engine/llm_engine.rs:60-95,146-189):
    right-pad with eos to [B,max_len], FULL forward of every sequence every step,
    logits at ALL positions, pick position len-1, host argmax (tie -> last index).
    Sampling with temperature is unseeded in the reference (llm_engine.rs:128) and thus
    irreproducible; the oracle implements the greedy/argmax branch only."""

    def __init__(self, model, eos_id: int):
        self.model = model
        self.eos_id = eos_id
        self.last_logits = None

    def build_batch(self, seqs):                                       # :60-95
        import torch
        all_ids = [s.prompt_token_ids() + s.completion_token_ids() for s in seqs]
        lens = [len(i) for i in all_ids]
        max_len = max(lens) if lens else 0
        flat = [ids + [self.eos_id] * (max_len - len(ids)) for ids in all_ids]
        return torch.tensor(flat, dtype=torch.long), lens

    def run(self, seqs, is_prefill):
        from .qwen3_oracle import argmax_last
        if not seqs:
            return []
        input_ids, lens = self.build_batch(seqs)
        hidden = self.model.forward(input_ids)
        logits = self.model.compute_logits(hidden)                     # all positions, :169
        self.last_logits = [logits[i, l - 1].clone() for i, l in enumerate(lens)]
        return [argmax_last(logits[i, l - 1].numpy()) for i, l in enumerate(lens)]
