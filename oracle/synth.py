"""Synthetic checkpoints / prompts: re-exported from the package's data generator
(nano-vllm-candle_b200/synth.py), which contains no model math."""
import importlib as _il

_m = _il.import_module("nano-vllm-candle_b200.synth")
tensor_specs, make_tensor, iter_weights, make_weights = _m.tensor_specs, _m.make_tensor, _m.iter_weights, _m.make_weights
write_hf_dir, make_prompts, hf_config_json = _m.write_hf_dir, _m.make_prompts, _m.hf_config_json
