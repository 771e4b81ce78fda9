"""Kernel-level Python entry points: torch tensors are used ONLY as device-memory handles
(data_ptr) and for the current CUDA stream; all compute is in libnvc_b200.so through the C ABI
(include/nvc_b200.h).  There is no torch / CPU fallback: a missing library or a non-zero status
raises."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import check

BF16 = torch.bfloat16


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (no CPU fallback on this path)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")


class Context:
    """nvc_ctx: one per (process, device).  Mirrors TPConfig{size, rank} (src/tp.rs:3-70)."""

    def __init__(self, device: int = 0, tp_rank: int = 0, tp_size: int = 1):
        self.l = _lib.lib()
        h = C.c_void_p()
        check(self.l.nvc_ctx_create(device, tp_rank, tp_size, C.byref(h)), "nvc_ctx_create")
        self.h = h
        self.device = device
        self.tp_rank, self.tp_size = tp_rank, tp_size
        self.num_sms = self.l.nvc_ctx_num_sms(self.h)
        self._ws = None

    def close(self):
        if self.h:
            self.l.nvc_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, st, where):
        check(st, where, self.h)

    # ---- collectives
    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(_lib.lib().nvc_nccl_unique_id(buf), "nvc_nccl_unique_id")
        return buf.raw

    def init_nccl(self, unique_id: bytes):
        assert len(unique_id) == 128
        self._ck(self.l.nvc_ctx_init_nccl(self.h, C.c_char_p(unique_id)), "nvc_ctx_init_nccl")

    def p2p_export(self, slot_bytes: int) -> bytes:
        """Allocates this rank's peer buffer; returns the 64-byte CUDA IPC handle to all-gather."""
        buf = C.create_string_buffer(64)
        self._ck(self.l.nvc_ctx_p2p_export(self.h, slot_bytes, buf), "nvc_ctx_p2p_export")
        return buf.raw

    def p2p_import(self, handles) -> None:
        """handles: the tp_size 64-byte handles in rank order."""
        blob = b"".join(handles)
        assert len(blob) == 64 * self.tp_size
        self._ck(self.l.nvc_ctx_p2p_import(self.h, C.c_char_p(blob)), "nvc_ctx_p2p_import")

    def init_tp(self, dist, slot_bytes: int = 4 << 20) -> None:
        """NCCL communicator + NVLink peer buffers, rendezvous through torch.distributed."""
        box = [Context.nccl_unique_id() if self.tp_rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        self.init_nccl(box[0])
        mine = self.p2p_export(slot_bytes)
        allh = [None] * self.tp_size
        dist.all_gather_object(allh, mine)
        self.p2p_import(allh)

    def nvls_init(self, slot_bytes: int, ll_buf_bytes: int = 4 << 20, group_name: Optional[str] = None) -> bool:
        """COLLECTIVE: the NVSwitch multicast buffers of the TP group (nvc_ctx_nvls_init).  False (on every rank
        alike) when multicast is not available on this box; nvls_status() then says why."""
        st = self.l.nvc_ctx_nvls_init(self.h, slot_bytes, ll_buf_bytes, None if group_name is None else group_name.encode())
        if st == -8:
            return False
        self._ck(st, "nvc_ctx_nvls_init")
        return True

    def nvls_ready(self) -> bool:
        return bool(self.l.nvc_ctx_nvls_ready(self.h))

    def nvls_status(self) -> str:
        return self.l.nvc_ctx_nvls_status(self.h).decode()

    def nvls_slot(self, rows: int, cols: int) -> torch.Tensor:
        """This rank's NVLS slot viewed as a bf16 [rows, cols] tensor (memory owned by the context)."""
        nbytes = rows * cols * 2
        assert nbytes <= self.l.nvc_ctx_nvls_slot_bytes(self.h)
        ptr = self.l.nvc_ctx_nvls_slot(self.h)
        iface = {"shape": (rows, cols), "typestr": "<u2", "data": (ptr, False), "version": 3}
        holder = type("_Slot", (), {"__cuda_array_interface__": iface})()
        return torch.as_tensor(holder, device=f"cuda:{self.device}").view(BF16)

    def nvls_allreduce_rmsnorm(self, x_partial, w=None, eps=1e-6, residual_in=None, residual_out=None, out=None, mode=0):
        """sum over ranks of x_partial [N, H] (+ residual, RMSNorm when w is given) through the switch."""
        _req(x_partial, BF16, "x_partial")
        N, H = x_partial.shape
        out = torch.empty((N, H), dtype=BF16, device=x_partial.device) if out is None else out
        self._ck(self.l.nvc_nvls_allreduce_rmsnorm(self.h, _ptr(x_partial), _ptr(residual_in), _ptr(residual_out), _ptr(w),
                                                   float(eps), _ptr(out), N, H, mode, _stream()), "nvc_nvls_allreduce_rmsnorm")
        return out

    def allreduce(self, buf: torch.Tensor):
        if buf.dtype == BF16:
            self._ck(self.l.nvc_allreduce_bf16(self.h, _ptr(buf), buf.numel(), _stream()), "nvc_allreduce_bf16")
        elif buf.dtype == torch.float32:
            self._ck(self.l.nvc_allreduce_f32(self.h, _ptr(buf), buf.numel(), _stream()), "nvc_allreduce_f32")
        else:
            raise TypeError("allreduce: bf16 or f32")
        return buf

    # ---- kernels
    def embed_gather(self, ids: torch.Tensor, table: torch.Tensor, out: Optional[torch.Tensor] = None):
        _req(ids, torch.int32, "ids"); _req(table, BF16, "table")
        N, (V, H) = ids.numel(), table.shape
        out = torch.empty((N, H), dtype=BF16, device=table.device) if out is None else out
        self._ck(self.l.nvc_embed_gather(self.h, _ptr(ids), _ptr(table), _ptr(out), N, H, V, _stream()), "nvc_embed_gather")
        return out

    def fused_add_rmsnorm(self, x, w, eps, residual_in=None, residual_out=None, out=None):
        _req(x, BF16, "x"); _req(w, BF16, "w")
        N, H = x.shape
        out = torch.empty_like(x) if out is None else out
        self._ck(self.l.nvc_fused_add_rmsnorm(self.h, _ptr(x), _ptr(residual_in), _ptr(residual_out), _ptr(w),
                                              float(eps), _ptr(out), N, H, _stream()), "nvc_fused_add_rmsnorm")
        return out

    def gemm(self, a: torch.Tensor, w: torch.Tensor, out_dtype=BF16, out=None, force_bn=0, force_splits=0):
        """C = a @ w.T; a [M,K] (row stride allowed), w [N,K] contiguous."""
        _req(a, BF16, "a"); _req(w, BF16, "w")
        M, K = a.shape
        N = w.shape[0]
        assert w.shape[1] == K and w.is_contiguous() and a.stride(1) == 1
        if out is None:
            out = torch.empty((M, N), dtype=out_dtype, device=a.device)
        dt = 0 if out.dtype == BF16 else 1
        self._ck(self.l.nvc_gemm_bf16_ex(self.h, _ptr(a), a.stride(0), _ptr(w), _ptr(out), out.stride(0), dt,
                                         M, N, K, force_bn, force_splits, _stream()), "nvc_gemm_bf16")
        return out

    @staticmethod
    def interleave_gate_up(w_gu: torch.Tensor) -> torch.Tensor:
        """[gate; up] rows (reference layout, qwen3.rs:310) -> groups of 16 gate rows then 16 up rows,
        the layout nvc_gemm_bf16_silu_mul expects."""
        two_i, k = w_gu.shape
        i = two_i // 2
        assert i % 16 == 0
        return torch.stack([w_gu[:i].reshape(i // 16, 16, k), w_gu[i:].reshape(i // 16, 16, k)], dim=1).reshape(two_i, k).contiguous()

    def gemm_silu_mul(self, a: torch.Tensor, w_gu16: torch.Tensor, out=None, force_bn=0, force_splits=0):
        """act = silu(a @ Wgate.T) * (a @ Wup.T); w_gu16 = interleave_gate_up([gate; up])."""
        _req(a, BF16, "a"); _req(w_gu16, BF16, "w_gu16")
        M, K = a.shape
        I = w_gu16.shape[0] // 2
        assert w_gu16.shape[1] == K and w_gu16.is_contiguous() and a.stride(1) == 1
        if out is None:
            out = torch.empty((M, I), dtype=BF16, device=a.device)
        self._ck(self.l.nvc_gemm_bf16_silu_mul_ex(self.h, _ptr(a), a.stride(0), _ptr(w_gu16), _ptr(out), out.stride(0),
                                                  M, I, K, force_bn, force_splits, _stream()), "nvc_gemm_bf16_silu_mul")
        return out

    def rope_table(self, theta: float, head_dim: int, max_pos: int, device=None):
        dev = torch.device("cuda", self.device) if device is None else device
        tab = torch.empty((max_pos, head_dim), dtype=torch.float32, device=dev)
        self._ck(self.l.nvc_rope_table_init_sync(self.h, float(theta), head_dim, max_pos, _ptr(tab)), "nvc_rope_table_init_sync")
        return tab

    def qk_norm_rope_store(self, qkv, qw, kw, eps, cos_sin, positions, nh, kv, hd,
                           slot_mapping=None, k_cache=None, v_cache=None):
        _req(qkv, BF16, "qkv"); _req(positions, torch.int32, "positions")
        N = qkv.shape[0]
        assert qkv.is_contiguous() and qkv.shape[1] == (nh + 2 * kv) * hd
        self._ck(self.l.nvc_qk_norm_rope_store(self.h, _ptr(qkv), _ptr(qw), _ptr(kw), float(eps), _ptr(cos_sin),
                                               _ptr(positions), _ptr(slot_mapping), _ptr(k_cache), _ptr(v_cache),
                                               N, nh, kv, hd, _stream()), "nvc_qk_norm_rope_store")
        return qkv

    def qkv_proj_norm_rope(self, a, w_qkv, qw, kw, eps, cos_sin, positions, nh, kv, hd, slot_mapping=None, k_cache=None,
                           v_cache=None, out=None):
        """K3 + K2: qkv = a @ w_qkv.T with q / k heads normalised + rotated and k / v scattered to the caches (one kernel at
        prefill sizes)."""
        _req(a, BF16, "a"); _req(w_qkv, BF16, "w_qkv"); _req(positions, torch.int32, "positions")
        M, K = a.shape
        N = (nh + 2 * kv) * hd
        assert w_qkv.shape == (N, K) and w_qkv.is_contiguous() and a.stride(1) == 1
        out = torch.empty((M, N), dtype=BF16, device=a.device) if out is None else out
        self._ck(self.l.nvc_qkv_proj_norm_rope(self.h, _ptr(a), a.stride(0), _ptr(w_qkv), _ptr(out), M, K, _ptr(qw), _ptr(kw),
                                               float(eps), _ptr(cos_sin), _ptr(positions), _ptr(slot_mapping), _ptr(k_cache),
                                               _ptr(v_cache), nh, kv, hd, _stream()), "nvc_qkv_proj_norm_rope")
        return out

    def store_kvcache(self, k, v, k_cache, v_cache, slot_mapping, kv, hd):
        _req(k, BF16, "k"); _req(slot_mapping, torch.int32, "slot_mapping")
        N = k.shape[0]
        assert k.stride(0) == v.stride(0)
        self._ck(self.l.nvc_store_kvcache(self.h, _ptr(k), _ptr(v), k.stride(0), _ptr(k_cache), _ptr(v_cache),
                                          _ptr(slot_mapping), N, kv, hd, _stream()), "nvc_store_kvcache")

    def prefill_attn(self, q, k, v, cu_seqlens, max_seqlen, nh, kv, hd, scale, out=None):
        """q [N, nh*hd], k/v [N, kv*hd] views (row stride allowed)."""
        _req(q, BF16, "q"); _req(cu_seqlens, torch.int32, "cu_seqlens")
        N = q.shape[0]
        B = cu_seqlens.numel() - 1
        out = torch.empty((N, nh * hd), dtype=BF16, device=q.device) if out is None else out
        self._ck(self.l.nvc_prefill_attn_varlen(self.h, _ptr(q), q.stride(0), _ptr(k), k.stride(0), _ptr(v), v.stride(0),
                                                _ptr(cu_seqlens), B, N, max_seqlen, nh, kv, hd, float(scale),
                                                _ptr(out), _stream()), "nvc_prefill_attn_varlen")
        return out

    def prefill_attn_paged(self, q, k_cache, v_cache, block_tables, context_lens, cu_seqlens_q, max_seqlen_q,
                           nh, kv, hd, block_size, scale, out=None):
        """K4 over the paged cache (cached prefix): q [N, nh*hd] = the last q_len_b positions of each
        context; k_cache / v_cache [num_blocks, block_size, kv, hd] already hold the whole context."""
        _req(q, BF16, "q"); _req(cu_seqlens_q, torch.int32, "cu_seqlens_q")
        _req(block_tables, torch.int32, "block_tables"); _req(context_lens, torch.int32, "context_lens")
        N = q.shape[0]
        B = cu_seqlens_q.numel() - 1
        out = torch.empty((N, nh * hd), dtype=BF16, device=q.device) if out is None else out
        self._ck(self.l.nvc_prefill_attn_paged(self.h, _ptr(q), q.stride(0), _ptr(k_cache), _ptr(v_cache),
                                               _ptr(block_tables), block_tables.shape[1], _ptr(context_lens),
                                               _ptr(cu_seqlens_q), B, N, max_seqlen_q, nh, kv, hd, block_size,
                                               k_cache.shape[0], float(scale), _ptr(out), _stream()),
                 "nvc_prefill_attn_paged")
        return out

    def paged_decode_attn(self, q, k_cache, v_cache, block_tables, context_lens, nh, kv, hd, block_size, scale, out=None):
        _req(q, BF16, "q"); _req(block_tables, torch.int32, "block_tables"); _req(context_lens, torch.int32, "context_lens")
        B = q.shape[0]
        out = torch.empty((B, nh * hd), dtype=BF16, device=q.device) if out is None else out
        need = self.l.nvc_paged_decode_workspace_bytes(B, nh, hd, self.num_sms)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 1), dtype=torch.uint8, device=q.device)
        self._ck(self.l.nvc_paged_decode_attn(self.h, _ptr(q), q.stride(0), _ptr(k_cache), _ptr(v_cache),
                                              _ptr(block_tables), block_tables.shape[1], _ptr(context_lens), B, nh, kv,
                                              hd, block_size, float(scale), _ptr(out), _ptr(self._ws),
                                              self._ws.numel(), _stream()), "nvc_paged_decode_attn")
        return out

    def silu_and_mul(self, x, out=None):
        _req(x, BF16, "x")
        N, I2 = x.shape
        out = torch.empty((N, I2 // 2), dtype=BF16, device=x.device) if out is None else out
        self._ck(self.l.nvc_silu_and_mul(self.h, _ptr(x), _ptr(out), N, I2 // 2, _stream()), "nvc_silu_and_mul")
        return out

    def gather_rows(self, x, residual, rows):
        _req(x, BF16, "x"); _req(rows, torch.int32, "rows")
        B, H = rows.numel(), x.shape[1]
        xo = torch.empty((B, H), dtype=BF16, device=x.device)
        ro = None if residual is None else torch.empty((B, H), dtype=torch.float32, device=x.device)
        self._ck(self.l.nvc_gather_rows(self.h, _ptr(x), _ptr(residual), _ptr(rows), _ptr(xo), _ptr(ro), B, H, _stream()),
                 "nvc_gather_rows")
        return xo, ro

    def argmax_last(self, logits, index_offset=0):
        _req(logits, torch.float32, "logits")
        B, V = logits.shape
        idx = torch.empty((B,), dtype=torch.int32, device=logits.device)
        val = torch.empty((B,), dtype=torch.float32, device=logits.device)
        self._ck(self.l.nvc_argmax_last(self.h, _ptr(logits), logits.stride(0), _ptr(idx), _ptr(val), B, V,
                                        index_offset, _stream()), "nvc_argmax_last")
        return idx, val

    def sample_gumbel(self, logits, temperatures, streams, seed: int, index_offset=0):
        """K8s: seeded Gumbel-max draw per row; streams: int64 tensor holding the u64 stream ids."""
        _req(logits, torch.float32, "logits"); _req(temperatures, torch.float32, "temperatures")
        _req(streams, torch.int64, "streams")
        B, V = logits.shape
        idx = torch.empty((B,), dtype=torch.int32, device=logits.device)
        val = torch.empty((B,), dtype=torch.float32, device=logits.device)
        self._ck(self.l.nvc_sample_gumbel(self.h, _ptr(logits), logits.stride(0), _ptr(temperatures), _ptr(streams),
                                          C.c_uint64(seed & (2 ** 64 - 1)), _ptr(idx), _ptr(val), B, V, index_offset,
                                          _stream()), "nvc_sample_gumbel")
        return idx, val
