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
    assert stats["injected"] > 20, r.stdout            # runner failures were injected and rolled back
    assert "loader:" in r.stdout


MUTANTS = {
    # chunked prefill + prefix caching: a hit on a block whose K / V are not computed yet
    "hit_before_ready": ("(chunked && !blocks[id].ready))", "(false))"),
    # a shared block returned to the free list while another sequence still holds it
    "freed_while_shared": ("if (--b.ref_count == 0) free_.push_back", "if (--b.ref_count >= 0) free_.push_back"),
    # a step that failed in the runner leaves its batch scheduled (running, blocks allocated and hash-registered)
    "no_rollback": ("    rollback_failed_step(e);", "    // (mutant)"),
    # a re-admitted (preempted) sequence keeps its old progress although its blocks are new
    "stale_progress": ("      seq->num_computed_tokens = seq->num_cached_tokens;", "      // (mutant)",
                       "    seq->num_computed_tokens = 0;", "    // (mutant)"),
}


@pytest.mark.parametrize("name", sorted(MUTANTS))
def test_fuzz_harness_catches_seeded_scheduler_bugs(fuzz_binary, tmp_path, name):
    """The harness is only worth its green runs if it goes red on real bugs: each mutant is a one-line change to a
    COPY of engine.cpp (the product source is untouched) that the cache model / invariants must catch within a few seeds."""
    edits = MUTANTS[name]
    src = open(os.path.join(CSRC, "engine.cpp")).read()
    for old, new in zip(edits[::2], edits[1::2]):
        assert src.count(old) == 1, f"mutation site of {name} moved"
        src = src.replace(old, new)
    src = src.replace('#include "../../include/nvc_b200_engine.h"',
                                        f'#include "{os.path.join(ROOT, "include", "nvc_b200_engine.h")}"')
    mutant = tmp_path / "engine_mutant.cpp"
    mutant.write_text(src)
    out = tmp_path / "engine_fuzz_mutant"
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
                        os.path.join(ROOT, "tests", "native", "engine_fuzz.cpp"), str(mutant),
                        os.path.join(CSRC, "hf_loader.cpp"), "-o", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    for seed in range(1, 7):
        r = subprocess.run([str(out), str(seed), "800"], capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            assert "contract violated" in r.stderr or "Sanitizer" in r.stderr, r.stderr[-2000:]
            return
    pytest.fail(f"mutant {name} survived 6 seeds")
