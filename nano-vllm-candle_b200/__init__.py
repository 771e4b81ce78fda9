"""nano-vllm-candle_b200: B200-native (sm_100a) Qwen3 forward hot path behind
nano-vllm-candle's layer / runner API.  Import with
``importlib.import_module("nano-vllm-candle_b200")`` (the directory name is not a Python
identifier).  The CUDA library is a C-ABI shared object, see include/nvc_b200.h."""
from . import _lib  # noqa: F401
from ._lib import NvcError, lib, LIB_PATH  # noqa: F401
from .build import build as build_library  # noqa: F401

__all__ = ["NvcError", "lib", "LIB_PATH", "build_library"]
