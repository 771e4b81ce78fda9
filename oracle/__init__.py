"""CPU oracle for the Qwen3 forward hot path of nano-vllm-candle.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the package
``nano-vllm-candle_b200`` and its CUDA library) may import, call or link
anything from this directory.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs use it, and
only as the checker / the timed CPU baseline.

Parity status (see DESIGN.md section "Oracle"):
  * Linear, RMSNorm, SiluAndMul, TP shard arithmetic, GQA interleave:
    pinned by the reference's own unit-test vectors (tests/test_oracle_goldens.py).
  * RoPE: pinned by the reference's norm-preservation properties only.
  * Attention / full forward / logits: PARITY UNPINNED by any reference test
    runnable here (the reference's integration tests need real Qwen3-0.6B
    weights at a Windows path; cargo/rustc are absent).  The restatement is
    cross-checked against an independent implementation (HF transformers
    Qwen3 on CPU, same synthetic weights) instead.
"""
