"""Tensor-parallel GPU parity (needs >= 2 GPUs on the box; skipped otherwise): runs
tools/tp_check.py under torch.distributed.run, one process per GPU, NCCL all-reduce after each
row-parallel linear, and checks the greedy tokens against the fp32 oracle."""
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
def test_tp_generation_matches_oracle(tp, model, chunked, tmp_path):
    if torch.cuda.device_count() < tp:
        pytest.skip(f"needs {tp} GPUs")
    out = tmp_path / "tp.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={tp}",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tools", "tp_check.py"),
           "--model", model, "--out", str(out)] + (["--chunked", str(chunked)] if chunked else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.load(open(out))
    assert res["ranks_agree"] and res["ok"] and res["worst_logit_deficit"] <= 4e-2
    assert not chunked or res["mixed_steps"] > 0
