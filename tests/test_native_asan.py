"""Host side under AddressSanitizer + UBSan: tests/native/engine_fuzz.cpp links csrc/engine.cpp and
csrc/hf_loader.cpp against a stub model runner that checks every step's inputs against the integer
contract (slot_mapping / block_tables / context_lens / cu_seqlens) the way csrc/runner.cu
validates them, and drives random call sequences + mutated checkpoint files.  CPU only; no oracle involved."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "nano-vllm-candle_b200", "csrc")


@pytest.fixture(scope="module")
def fuzz_binary(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("g++ not available")
    out = tmp_path_factory.mktemp("asan") / "engine_fuzz"
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
           "-fno-omit-frame-pointer", os.path.join(ROOT, "tests", "native", "engine_fuzz.cpp"),
           os.path.join(CSRC, "engine.cpp"), os.path.join(CSRC, "hf_loader.cpp"), "-o", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 and "asan" in r.stderr.lower() and "cannot find" in r.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert r.returncode == 0, r.stderr[-4000:]
    return str(out)


@pytest.mark.parametrize("seed", [1, 2])
def test_engine_and_loader_fuzz_under_sanitizers(fuzz_binary, tmp_path, seed):
    r = subprocess.run([fuzz_binary, str(seed), "300", str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-6000:]
    last = r.stdout.strip().splitlines()[-1].split()
    stats = {last[i]: int(last[i + 1]) for i in range(1, len(last) - 1) if last[i + 1].isdigit()}
    # the run reached every runner entry point and verified finished outputs, not just error paths
    assert stats["engines"] > 100 and stats["steps"] > 1000 and stats["outputs"] > 20, r.stdout
    assert min(stats["prefill"], stats["paged"], stats["mixed"], stats["decode"]) > 0, r.stdout
    assert "loader:" in r.stdout
