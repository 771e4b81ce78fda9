"""CPU tests of the sampling oracle: the Philox4x32-10 restatement against the Random123
known-answer vectors (kat_vectors of the Random123 distribution, Salmon et al. SC'11), and the
Gumbel-max draw against the distribution the reference samples from
(src/engine/llm_engine.rs:100-125)."""
import numpy as np

from oracle import sampling_oracle as S

KAT = [
    ([0, 0, 0, 0], (0, 0), [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, (0xffffffff, 0xffffffff), [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], (0xa4093822, 0x299f31d0),
     [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


def _chi2(counts, expected):
    """Pearson statistic with the bins of expectation < 5 lumped into one."""
    small = expected < 5
    c = np.append(counts[~small], counts[small].sum())
    e = np.append(expected[~small], expected[small].sum())
    keep = e > 0
    return float((((c - e) ** 2)[keep] / e[keep]).sum())


def test_philox4x32_10_known_answers():
    for ctr, key, want in KAT:
        got = S.philox4x32_10(np.array([ctr], dtype=np.uint32), key)[0].tolist()
        assert got == want


def test_uniforms_are_open_interval_and_depend_on_global_index_only():
    u = S.uniforms(4096, seed=1234, stream=S.stream_id(7, 3))
    assert u.dtype == np.float32 and (u > 0).all() and (u < 1).all()
    # the extreme noise words still give finite Gumbel values (a 24-bit draw + 2^-25 rounds to 1.0f)
    ext = ((((np.array([0, 0xFFFFFFFF], np.uint32) >> np.uint32(9)) << np.uint32(1)) | np.uint32(1)).astype(np.float32)
           * np.float32(2.0 ** -24))
    assert ext[0] == np.float32(2.0 ** -24) and ext[1] == np.float32(1 - 2.0 ** -24) and ext[1] < 1
    # a vocab shard sees the same noise as the full row (tensor-parallel invariance)
    shard = S.uniforms(1024, seed=1234, stream=S.stream_id(7, 3), index_offset=2048)
    assert np.array_equal(shard, u[2048:3072])
    assert not np.array_equal(u, S.uniforms(4096, seed=1234, stream=S.stream_id(7, 4)))
    assert not np.array_equal(u, S.uniforms(4096, seed=1235, stream=S.stream_id(7, 3)))


def test_gumbel_max_draws_follow_the_reference_distribution():
    rng = np.random.default_rng(0)
    logits = rng.normal(0, 2.0, size=16).astype(np.float32)
    for k, T in enumerate((0.7, 1.0, 2.5)):
        p = S.reference_probs(logits, T)
        n = 20000
        counts = np.bincount([S.sample(logits, T, seed=1, stream=S.stream_id(k, i)) for i in range(n)], minlength=16)
        chi2 = _chi2(counts, n * p)
        assert chi2 < 45.0, (T, chi2)          # <= 15 degrees of freedom: P(chi2 > 45) < 1e-4


def test_low_temperature_is_greedy_and_ties_go_to_the_last_index():
    logits = np.array([0.1, 3.0, -1.0, 2.9, 0.0, 0.0, 0.0, 0.0], np.float32)
    assert all(S.sample(logits, 1e-9, 5, s) == 1 for s in range(50))      # T clamps to 1e-6: noise is negligible
    z = np.zeros(8, np.float32)
    s = S.gumbel_scores(z, 1.0, 5, 0)
    assert S.sample(z, 1.0, 5, 0) == int(np.argmax(s))
