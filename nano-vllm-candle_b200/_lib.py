"""ctypes binding of libnvc_b200.so (include/nvc_b200.h, include/nvc_b200_engine.h).

The product path has no CPU fallback: if the shared library is missing or a call fails,
an exception is raised with the library's own error string."""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Dict, List, Tuple

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnvc_b200.so")
INCLUDE_DIR = os.path.join(os.path.dirname(HERE), "include")

_lib = None


class NvcError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        super().__init__(f"{where} failed with status {status} ({status_string(status)}){': ' + detail if detail else ''}")


def lib() -> C.CDLL:
    """Loads the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python nano-vllm-candle_b200/build.py` "
                "(or __graft_entry__.build()). There is no CPU fallback on this path.")
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _declare(_lib)
    return _lib


def status_string(status: int) -> str:
    l = lib()
    return l.nvc_status_string(status).decode()


def check(status: int, where: str, ctx=None) -> None:
    if status != 0:
        detail = lib().nvc_last_error(ctx).decode() if True else ""
        raise NvcError(status, where, detail)


_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t


def _declare(l: C.CDLL) -> None:
    l.nvc_status_string.restype = C.c_char_p
    l.nvc_status_string.argtypes = [_i]
    l.nvc_last_error.restype = C.c_char_p
    l.nvc_last_error.argtypes = [_vp]
    l.nvc_abi_version.restype = _i
    sigs = {
        "nvc_ctx_create": [_i, _i, _i, C.POINTER(_vp)],
        "nvc_ctx_destroy": [_vp],
        "nvc_ctx_device": [_vp], "nvc_ctx_tp_rank": [_vp], "nvc_ctx_tp_size": [_vp], "nvc_ctx_num_sms": [_vp],
        "nvc_nccl_unique_id": [_vp],
        "nvc_ctx_init_nccl": [_vp, _vp],
        "nvc_allreduce_bf16": [_vp, _vp, _sz, _vp],
        "nvc_allreduce_f32": [_vp, _vp, _sz, _vp],
        "nvc_allgather_bytes": [_vp, _vp, _sz, _vp],
        "nvc_ctx_p2p_export": [_vp, _sz, _vp],
        "nvc_ctx_p2p_import": [_vp, _vp],
        "nvc_ctx_p2p_ready": [_vp],
        "nvc_p2p_allreduce_rmsnorm": [_vp, _i, _vp, _vp, _vp, _f, _vp, _i, _i, _vp],
        "nvc_ctx_nvls_init": [_vp, _sz, _sz, C.c_char_p],
        "nvc_ctx_nvls_ready": [_vp],
        "nvc_nvls_allreduce_rmsnorm": [_vp, _vp, _vp, _vp, _vp, _f, _vp, _i, _i, _i, _vp],
        "nvc_nvls_argmax_merge": [_vp, _vp, _vp, _vp, _vp, _i, _vp],
        "nvc_embed_gather": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp],
        "nvc_fused_add_rmsnorm": [_vp, _vp, _vp, _vp, _vp, _f, _vp, _i, _i, _vp],
        "nvc_gemm_bf16": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _vp],
        "nvc_gemm_bf16_ex": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp],
        "nvc_gemm_bf16_silu_mul": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp],
        "nvc_gemm_bf16_silu_mul_ex": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
        "nvc_gemm_bf16_trace": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp],
        "nvc_rope_table_init_sync": [_vp, _f, _i, _i, _vp],
        "nvc_qkv_proj_norm_rope": [_vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
        "nvc_qk_norm_rope_store": [_vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
        "nvc_store_kvcache": [_vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _vp],
        "nvc_prefill_attn_varlen": [_vp, _vp, _i, _vp, _i, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _f, _vp, _vp],
        "nvc_prefill_attn_paged": [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _vp],
        "nvc_paged_decode_attn": [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp, _sz, _vp],
        "nvc_paged_decode_attn_static": [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp],
        "nvc_silu_and_mul": [_vp, _vp, _vp, _i, _i, _vp],
        "nvc_gather_rows": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
        "nvc_argmax_last": [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp],
        "nvc_sample_gumbel": [_vp, _vp, _i, _vp, _vp, C.c_uint64, _vp, _vp, _i, _i, _i, _vp],
        "nvc_debug_set": [_i, _i],
    }
    for name, args in sigs.items():
        fn = getattr(l, name)
        fn.argtypes = args
        fn.restype = _i
    l.nvc_ctx_p2p_slot.restype = _vp
    l.nvc_ctx_p2p_slot.argtypes = [_vp, _i]
    l.nvc_ctx_p2p_slot_bytes.restype = _sz
    l.nvc_ctx_p2p_slot_bytes.argtypes = [_vp]
    l.nvc_ctx_nvls_slot.restype = _vp
    l.nvc_ctx_nvls_slot.argtypes = [_vp]
    l.nvc_ctx_nvls_slot_bytes.restype = _sz
    l.nvc_ctx_nvls_slot_bytes.argtypes = [_vp]
    l.nvc_ctx_nvls_ll_bytes.restype = _sz
    l.nvc_ctx_nvls_ll_bytes.argtypes = [_vp]
    l.nvc_ctx_nvls_status.restype = C.c_char_p
    l.nvc_ctx_nvls_status.argtypes = [_vp]
    l.nvc_paged_decode_workspace_bytes.restype = _sz
    l.nvc_paged_decode_workspace_bytes.argtypes = [_i, _i, _i, _i]
    l.nvc_paged_decode_static_scratch_bytes.restype = _sz
    l.nvc_paged_decode_static_scratch_bytes.argtypes = []
    from . import _engine_sigs
    _engine_sigs.declare(l)


def declared_symbols() -> List[str]:
    """Every function name declared in include/*.h (used by the not-gpu export test)."""
    names: List[str] = []
    for fn in sorted(os.listdir(INCLUDE_DIR)):
        if not fn.endswith(".h"):
            continue
        src = open(os.path.join(INCLUDE_DIR, fn)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(nvc_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))
