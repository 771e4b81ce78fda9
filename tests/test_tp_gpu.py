"""Tensor-parallel GPU parity (needs >= 2 GPUs on the box; skipped otherwise): runs
tools/tp_check.py under torch.distributed.run, one process per GPU, and checks the greedy tokens against the
fp32 oracle - once with the in-switch (NVLS multicast) all-reduce the runner uses by default and once with it
switched off (NVC_NVLS=0: peer-load kernel for decode, NCCL for prefill).  tools/nvls_check.py compares the NVLS
kernels themselves with a torch fp32 evaluation over 13 shapes x 4 modes and requires bit-identical results on
all ranks.  (bench.py repeats the oracle check at every N it is run at: that is the evidence on 4 and 8 GPUs.)"""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("tp", [2])
@pytest.mark.parametrize("model", ["qwen3-tiny"])
@pytest.mark.parametrize("chunked", [0, 24])
@pytest.mark.parametrize("nvls", [1, 0])
def test_tp_generation_matches_oracle(tp, model, chunked, nvls, tmp_path):
    if torch.cuda.device_count() < tp:
        pytest.skip(f"needs {tp} GPUs")
    out = tmp_path / "tp.json"
    env = dict(os.environ, NVC_NVLS=str(nvls))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={tp}",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tools", "tp_check.py"),
           "--model", model, "--out", str(out)] + (["--chunked", str(chunked)] if chunked else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    assert res["ranks_agree"] and res["ok"] and res["worst_logit_deficit"] <= 4e-2
    assert not chunked or res["mixed_steps"] > 0
    assert not nvls or res.get("nvls", True), res


@pytest.mark.parametrize("tp", [2])
def test_nvls_allreduce_kernels_match_torch(tp, tmp_path):
    """K9 on NVLS (csrc/nvls_allreduce.cu): sum over ranks + residual + RMSNorm vs torch fp32 on the all-gathered
    partial sums, every mode, ragged row counts back to back (the LL buffers are refilled by extent), NaN payloads
    that collide with the arrival sentinel; all ranks bit-identical.  A box without NVSwitch multicast reports
    why and the test is skipped (the runner then stays on the peer-load / NCCL path)."""
    if torch.cuda.device_count() < tp:
        pytest.skip(f"needs {tp} GPUs")
    out = tmp_path / "nvls.jsonl"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={tp}",
           "--master-addr", "127.0.0.1", "--master-port", "29518", os.path.join(ROOT, "tools", "nvls_check.py"),
           "--no-timing", "--out", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    lines = [json.loads(l) for l in open(out)]
    if not lines[0]["ok"]:
        pytest.skip("NVLS unavailable: " + lines[0]["status"])
    summary = lines[-1]
    assert summary["parity_cases"] >= 90 and summary["parity_failures"] == 0
