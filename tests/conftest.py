import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def nvc():
    """The product package (directory name is not an identifier)."""
    return importlib.import_module("nano-vllm-candle_b200")


@pytest.fixture(scope="session")
def ctx(nvc):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    ops = importlib.import_module("nano-vllm-candle_b200.ops")
    c = ops.Context(0)
    yield c
    c.close()


def diag(msg: str) -> None:
    """Append a diagnostic line to gpurun_out/diag.log (travels back from the GPU box)."""
    d = os.path.join(ROOT, "gpurun_out")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "diag.log"), "a") as f:
        f.write(msg.rstrip() + "\n")
