"""CPU restatement of the on-device temperature sampling (K8s) -- TEST INFRASTRUCTURE.

The reference samples on the host: softmax((l - max) / max(T, 1e-6)) followed by
rand::distr::WeightedIndex with an UNSEEDED rand::rng() (src/engine/llm_engine.rs:97-133),
so its draws cannot be reproduced.  What can be pinned is the distribution it draws from
(`reference_probs`) and, for the replacement, the exact counter-based noise: Philox4x32-10
(Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11; known-answer vectors of
the Random123 distribution are checked in tests/test_sampling_oracle.py) feeding a Gumbel-max
draw, which samples the same categorical distribution.
"""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(counter: np.ndarray, key) -> np.ndarray:
    """counter: uint32 [..., 4]; key: (k0, k1).  Returns uint32 [..., 4]."""
    c = [counter[..., i].astype(np.uint64) for i in range(4)]
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return np.stack(c, axis=-1).astype(np.uint32)


def uniforms(V: int, seed: int, stream: int, index_offset: int = 0) -> np.ndarray:
    """u[i] for global vocab indices index_offset .. index_offset + V - 1 (both multiples of 4)."""
    assert V % 4 == 0 and index_offset % 4 == 0
    blocks = np.arange(index_offset // 4, (index_offset + V) // 4, dtype=np.uint32)
    ctr = np.zeros((len(blocks), 4), dtype=np.uint32)
    ctr[:, 0] = blocks
    ctr[:, 1] = stream & 0xFFFFFFFF
    ctr[:, 2] = (stream >> 32) & 0xFFFFFFFF
    x = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)).reshape(-1)
    # (2 * (x >> 9) + 1) * 2^-24: odd 24-bit integer over 2^24, exact in float32, never 0 or 1
    return ((((x >> np.uint32(9)) << np.uint32(1)) | np.uint32(1)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)


def gumbel_scores(logits: np.ndarray, temperature: float, seed: int, stream: int, index_offset: int = 0) -> np.ndarray:
    """float64 perturbed scores l_i / max(T, 1e-6) + g_i of one row (float32 division as on device)."""
    T = np.float32(max(np.float32(temperature), np.float32(1e-6)))
    u = uniforms(logits.shape[0], seed, stream, index_offset).astype(np.float64)
    g = -np.log(-np.log(u))
    return (logits.astype(np.float32) / T).astype(np.float64) + g


def sample(logits: np.ndarray, temperature: float, seed: int, stream: int, index_offset: int = 0) -> int:
    """argmax of the perturbed scores, ties -> last index (the kernel's rule)."""
    s = gumbel_scores(logits, temperature, seed, stream, index_offset)
    return int(len(s) - 1 - np.argmax(s[::-1])) + index_offset


def reference_probs(logits: np.ndarray, temperature: float) -> np.ndarray:
    """The distribution Qwen3ModelRunner::sample_token draws from (llm_engine.rs:100-125)."""
    t = max(float(temperature), 1e-6)
    z = (logits.astype(np.float64) - float(logits.max())) / t
    e = np.exp(z)
    return e / e.sum()


def stream_id(seq_id: int, step: int) -> int:
    """One noise stream per (sequence, number of completion tokens so far)."""
    return ((seq_id & 0xFFFFFFFF) << 32) | (step & 0xFFFFFFFF)
