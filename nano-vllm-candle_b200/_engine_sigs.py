"""ctypes prototypes for include/nvc_b200_engine.h."""
import ctypes as C

_vp, _i, _f, _sz, _i64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int64
_pi32 = C.POINTER(C.c_int32)


class Qwen3ConfigC(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("vocab_size", "hidden_size", "head_dim", "num_hidden_layers",
                                         "num_attention_heads", "num_key_value_heads", "intermediate_size",
                                         "max_position_embeddings")] + \
               [("rms_norm_eps", C.c_float), ("rope_theta", C.c_float),
                ("bos_token_id", C.c_int32), ("eos_token_id", C.c_int32)]


class RunnerOptionsC(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("num_kvcache_blocks", "kvcache_block_size", "max_num_seqs",
                                         "max_num_batched_tokens", "max_model_len", "use_cuda_graph",
                                         "allreduce_f32")] + [("reserved", C.c_int32 * 4)]


class SchedulerConfigC(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("max_num_seqs", "max_num_batched_tokens", "eos",
                                         "num_kvcache_blocks", "kvcache_block_size",
                                         "enable_prefix_caching", "enable_chunked_prefill", "max_model_len")] + \
               [("reserved", C.c_int32 * 1)]


class ScheduleInfoC(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("is_prefill", "num_seqs", "num_tokens", "max_blocks", "max_seqlen",
                                         "num_decode_seqs")]


class StepOutputC(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("is_prefill", "num_tokens", "num_seqs", "num_finished",
                                         "num_query_tokens", "num_sampled", "num_decode_seqs")]


def declare(l: C.CDLL) -> None:
    sigs = {
        "nvc_runner_create": [_vp, C.POINTER(Qwen3ConfigC), C.POINTER(RunnerOptionsC), C.POINTER(_vp)],
        "nvc_runner_destroy": [_vp],
        "nvc_runner_load_weight": [_vp, C.c_char_p, _vp, _i64, _i64],
        "nvc_runner_finalize": [_vp],
        "nvc_qwen3_config_from_hf_dir": [C.c_char_p, C.POINTER(Qwen3ConfigC), C.POINTER(C.c_int)],
        "nvc_runner_load_hf_dir": [_vp, C.c_char_p],
        "nvc_safetensors_open": [C.c_char_p, C.POINTER(_vp)],
        "nvc_safetensors_close": [_vp],
        "nvc_safetensors_num_tensors": [_vp],
        "nvc_safetensors_tensor_info": [_vp, _i, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.POINTER(C.c_int64)],
        "nvc_safetensors_read_bf16": [_vp, C.c_char_p, _vp, _i64],
        "nvc_runner_prefill": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
        "nvc_runner_prefill_paged": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _vp],
        "nvc_runner_decode": [_vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp],
        "nvc_runner_step_mixed": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp],
        "nvc_runner_set_sampling": [_vp, _vp, _vp, _i, C.c_uint64],
        "nvc_runner_last_logits": [_vp, _vp, _i],
        "nvc_runner_time_decode": [_vp, _i, C.POINTER(C.c_float), C.POINTER(C.c_int)],
        "nvc_runner_time_op": [_vp, _i, _i, C.POINTER(C.c_float), C.POINTER(C.c_int)],
        "nvc_runner_get_options": [_vp, C.POINTER(RunnerOptionsC)],
        "nvc_engine_create": [C.POINTER(SchedulerConfigC), _vp, C.POINTER(_vp)],
        "nvc_engine_destroy": [_vp],
        "nvc_engine_schedule": [_vp, C.POINTER(ScheduleInfoC)],
        "nvc_engine_step_inputs": [_vp] + [C.POINTER(_pi32)] * 6 + [C.POINTER(C.POINTER(C.c_int64))],
        "nvc_engine_step_emit_mask": [_vp, C.POINTER(_pi32)],
        "nvc_engine_post_process": [_vp, _vp, _i],
        "nvc_engine_step": [_vp, C.POINTER(StepOutputC)],
        "nvc_engine_is_finished": [_vp],
        "nvc_engine_set_sampling": [_vp, _i, C.c_uint64],
        "nvc_engine_preempt": [_vp, _i64],
        "nvc_engine_num_finished": [_vp],
        "nvc_engine_get_output": [_vp, _i64, _vp, _i, C.POINTER(C.c_int)],
        "nvc_engine_seq_block_table": [_vp, _i64, _vp, _i, C.POINTER(C.c_int)],
        "nvc_engine_num_free_blocks": [_vp],
        "nvc_engine_prefix_cache_stats": [_vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)],
        "nvc_engine_num_waiting": [_vp],
        "nvc_engine_num_running": [_vp],
    }
    for name, args in sigs.items():
        fn = getattr(l, name)
        fn.argtypes = args
        fn.restype = _i
    l.nvc_engine_add_request.argtypes = [_vp, _vp, _i, C.c_double, _i, _i]
    l.nvc_engine_add_request.restype = _i64
    l.nvc_hf_last_error.argtypes = []
    l.nvc_hf_last_error.restype = C.c_char_p
    l.nvc_hf_last_warning.argtypes = []
    l.nvc_hf_last_warning.restype = C.c_char_p
    l.nvc_engine_last_error.argtypes = [_vp]
    l.nvc_engine_last_error.restype = C.c_char_p
    l.nvc_runner_weight_bytes.argtypes = [_vp]
    l.nvc_runner_weight_bytes.restype = _i64
    l.nvc_runner_h2d_bytes_last_step.argtypes = [_vp]
    l.nvc_runner_h2d_bytes_last_step.restype = _i64
    l.nvc_runner_kv_bytes_per_token.argtypes = [_vp]
    l.nvc_runner_kv_bytes_per_token.restype = _i64
    l.nvc_ctx_launch_count.argtypes = [_vp]
    l.nvc_ctx_launch_count.restype = C.c_uint64
