"""Native HF loader (csrc/hf_loader.cpp) against independent implementations: files written by
the `safetensors` package, dtype conversion against torch's .to(bfloat16) bit for bit, config
parsing against json + the defaults of Qwen3Config::from_hf_dir (src/models/qwen3.rs:77-101).
The checkpoint -> runner path needs a GPU and is in the gpu-marked test at the bottom."""
import ctypes as C
import importlib
import json
import os

import numpy as np
import pytest
import torch
from safetensors.torch import save_file

from oracle import qwen3_oracle as O
from oracle import synth

eng = importlib.import_module("nano-vllm-candle_b200.engine")
_lib = importlib.import_module("nano-vllm-candle_b200._lib")


def _open(path):
    l = _lib.lib()
    h = C.c_void_p()
    st = l.nvc_safetensors_open(os.fsencode(str(path)), C.byref(h))
    return l, h, st


def _infos(l, h):
    out = {}
    for i in range(l.nvc_safetensors_num_tensors(h)):
        name, dt, nd = C.c_char_p(), C.c_int(), C.c_int()
        shape = (C.c_int64 * 4)()
        assert l.nvc_safetensors_tensor_info(h, i, C.byref(name), C.byref(dt), C.byref(nd), shape) == 0
        out[name.value.decode()] = (dt.value, list(shape)[:nd.value])
    return out


def _read(l, h, name, n):
    buf = np.empty(n, dtype=np.uint16)
    st = l.nvc_safetensors_read_bf16(h, name.encode(), buf.ctypes.data_as(C.c_void_p), n)
    assert st == 0, l.nvc_hf_last_error().decode()
    return buf


def test_safetensors_header_and_dtype_conversion(tmp_path):
    g = torch.Generator().manual_seed(0)
    specials = torch.tensor([0.0, -0.0, 1.0, -1.0, float("inf"), float("-inf"), float("nan"), 1e-40, -1e-40,
                             65504.0, 6.1e-5, 5.96e-8, 3.3895314e38, 1.0039062, 1.0117188])
    tensors = {
        "a.bf16": torch.randn(7, 24, generator=g).to(torch.bfloat16),
        "b.f16": torch.cat([torch.randn(500, generator=g) * 10, specials.clamp(-65504, 65504)]).to(torch.float16),
        "c.f32": torch.cat([torch.randn(1000, generator=g) * 100, specials]),
        "d.vec": torch.randn(128, generator=g).to(torch.bfloat16),
        "e.i64": torch.arange(5),
    }
    path = tmp_path / "t.safetensors"
    save_file(tensors, str(path), metadata={"format": "pt", "note": 'quotes " and \\ unicode é'})
    l, h, st = _open(path)
    assert st == 0, l.nvc_hf_last_error().decode()
    try:
        info = _infos(l, h)
        assert info == {"a.bf16": (0, [7, 24]), "b.f16": (1, [515]), "c.f32": (2, [1015]), "d.vec": (0, [128]),
                        "e.i64": (3, [5])}
        for name in ("a.bf16", "b.f16", "c.f32", "d.vec"):
            t = tensors[name]
            want = t.to(torch.bfloat16).contiguous().view(torch.int16).numpy().view(np.uint16).reshape(-1)
            got = _read(l, h, name, t.numel())
            nan = torch.isnan(t.float().reshape(-1)).numpy()
            assert np.array_equal(got[~nan], want[~nan]), name
            assert all(((int(x) & 0x7f80) == 0x7f80 and (int(x) & 0x7f) != 0) for x in got[nan]), "NaN must stay NaN"
        buf = np.empty(5, dtype=np.uint16)
        assert l.nvc_safetensors_read_bf16(h, b"e.i64", buf.ctypes.data_as(C.c_void_p), 5) == -3      # NVC_ENOTIMPL
        assert l.nvc_safetensors_read_bf16(h, b"a.bf16", buf.ctypes.data_as(C.c_void_p), 5) == -5     # NVC_ENOMEM
        assert l.nvc_safetensors_read_bf16(h, b"missing", buf.ctypes.data_as(C.c_void_p), 5) == -1
        assert b"missing" in l.nvc_hf_last_error()
    finally:
        l.nvc_safetensors_close(h)


def test_safetensors_rejects_malformed_files(tmp_path):
    cases = {
        "short.safetensors": b"\x01\x02",
        "hugehdr.safetensors": (1 << 40).to_bytes(8, "little") + b"{}",
        "notjson.safetensors": (5).to_bytes(8, "little") + b"hello",
        "badoffsets.safetensors": None,
    }
    hdr = json.dumps({"x": {"dtype": "F32", "shape": [4], "data_offsets": [0, 64]}}).encode()
    cases["badoffsets.safetensors"] = len(hdr).to_bytes(8, "little") + hdr + b"\0" * 16
    hdr = json.dumps({"x": {"dtype": "F32", "shape": [3], "data_offsets": [0, 16]}}).encode()
    cases["sizemismatch.safetensors"] = len(hdr).to_bytes(8, "little") + hdr + b"\0" * 16
    # found by tests/native/engine_fuzz.cpp: the same tensor listed twice (tensor_info and read_bf16 disagreed)
    hdr = b'{"x":{"dtype":"F32","shape":[2],"data_offsets":[0,8]},"x":{"dtype":"I64","shape":[1],"data_offsets":[8,16]}}'
    cases["duplicate.safetensors"] = len(hdr).to_bytes(8, "little") + hdr + b"\0" * 16
    hdr = json.dumps({"x": {"dtype": "F32", "shape": [2 ** 40, 2 ** 40], "data_offsets": [0, 0]}}).encode()   # product wraps
    cases["overflow.safetensors"] = len(hdr).to_bytes(8, "little") + hdr + b"\0" * 16
    hdr = json.dumps({"x": {"dtype": "F32", "shape": [2.5], "data_offsets": [0, 10]}}).encode()
    cases["fractional.safetensors"] = len(hdr).to_bytes(8, "little") + hdr + b"\0" * 16
    hdr = json.dumps({"x": {"dtype": "F32", "shape": [3], "data_offsets": [0, 12]},          # two tensors sharing bytes
                      "y": {"dtype": "F32", "shape": [2], "data_offsets": [8, 16]}}).encode()
    cases["overlap.safetensors"] = len(hdr).to_bytes(8, "little") + hdr + b"\0" * 16
    for name, blob in cases.items():
        p = tmp_path / name
        p.write_bytes(blob)
        l, h, st = _open(p)
        assert st != 0 and not h.value, name
        assert l.nvc_hf_last_error().decode() != ""
    l, h, st = _open(tmp_path / "does_not_exist.safetensors")
    assert st != 0


def _write_config(d, cfg, **extra):
    c = dict(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_hidden_layers=cfg.num_hidden_layers,
             num_attention_heads=cfg.num_attention_heads, num_key_value_heads=cfg.num_key_value_heads,
             intermediate_size=cfg.intermediate_size, max_position_embeddings=cfg.max_position_embeddings,
             rms_norm_eps=cfg.rms_norm_eps, hidden_act="silu", bos_token_id=cfg.bos_token_id,
             eos_token_id=cfg.eos_token_id, architectures=["Qwen3ForCausalLM"], nested={"a": [1, 2.5e-3, None, True]})
    c.update(extra)
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(c, f, indent=2)


def test_config_from_hf_dir(tmp_path):
    cfg = O.MODEL_CONFIGS["qwen3-tiny"]
    _write_config(tmp_path, cfg, head_dim=128, rope_theta=1000000.0, tie_word_embeddings=True)
    got = eng.Qwen3Config.from_hf_dir(str(tmp_path))
    for f in ("vocab_size", "hidden_size", "head_dim", "num_hidden_layers", "num_attention_heads", "num_key_value_heads",
              "intermediate_size", "max_position_embeddings", "bos_token_id", "eos_token_id"):
        assert getattr(got, f) == getattr(cfg, f), f
    assert got.rms_norm_eps == pytest.approx(cfg.rms_norm_eps) and got.rope_theta == 1e6 and got.tie_word_embeddings
    # defaults of qwen3.rs:88-96: head_dim = hidden / heads, rope_theta = 1e6; absent tie flag = False
    _write_config(tmp_path, cfg)
    got = eng.Qwen3Config.from_hf_dir(str(tmp_path))
    assert got.head_dim == cfg.hidden_size // cfg.num_attention_heads and got.rope_theta == 1e6
    assert not got.tie_word_embeddings
    # a missing required field is an error naming the field, like serde's
    c = json.load(open(tmp_path / "config.json"))
    del c["intermediate_size"]
    json.dump(c, open(tmp_path / "config.json", "w"))
    with pytest.raises(eng.NvcError, match="intermediate_size"):
        eng.Qwen3Config.from_hf_dir(str(tmp_path))
    with pytest.raises(eng.NvcError):
        eng.Qwen3Config.from_hf_dir(str(tmp_path / "nope"))


def test_config_json_is_parsed_strictly(tmp_path):
    """The loader's JSON reader accepts what serde_json (the reference's parser, qwen3.rs:77-86) accepts: RFC 8259
    numbers, no control characters / malformed UTF-8 / lone surrogates in strings, finite numbers only, and
    non-negative integers where the reference deserialises into usize / u32."""
    cfg = O.MODEL_CONFIGS["qwen3-tiny"]
    _write_config(tmp_path, cfg)
    good = open(tmp_path / "config.json", "rb").read()
    eng.Qwen3Config.from_hf_dir(str(tmp_path))

    def bad(old, new, match=None):
        assert good.count(old) == 1, old
        (tmp_path / "config.json").write_bytes(good.replace(old, new))
        with pytest.raises(eng.NvcError, match=match):
            eng.Qwen3Config.from_hf_dir(str(tmp_path))

    h = str(cfg.hidden_size).encode()
    for lit in (b"0" + h, h + b".", h + b"e", b"+" + h, b"." + h, b"--" + h, b"NaN", b"Infinity", b"0x10"):
        bad(b'"hidden_size": ' + h, b'"hidden_size": ' + lit)
    for lit in (b"-1", b"0.5", b"4294967296", b"1e400", b'"256"', b"null"):
        bad(b'"hidden_size": ' + h, b'"hidden_size": ' + lit, "hidden_size|out of range")
    bad(b'"silu"', b'"si\x01lu"', "control character")
    bad(b'"silu"', b'"si\xfflu"', "UTF-8")
    bad(b'"silu"', b'"si\xc3lu"', "UTF-8")
    bad(b'"silu"', b'"si\\ud800lu"', "surrogate")
    bad(b'"silu"', b'"si\\udc00lu"', "surrogate")
    bad(b'"silu"', b'"gelu"', "not supported")
    # escapes and multi-byte UTF-8 that ARE valid still parse
    (tmp_path / "config.json").write_bytes(good.replace(b'"silu"', b'"s\\u0069lu"').replace(b'"Qwen3ForCausalLM"', '"Qwen3 \\ud83d\\ude00 é 漢"'.encode()))
    assert eng.Qwen3Config.from_hf_dir(str(tmp_path)).hidden_size == cfg.hidden_size


def test_loader_survives_mutated_headers(tmp_path):
    """Seeded byte-level mutations of a valid safetensors file and a valid config.json (flipped bytes, truncation,
    header-length field, replaced numbers, structural characters): every call returns a status - no crash, no
    out-of-bounds read (each accepted file is read back in full) - and failures carry a message."""
    import random
    import re
    l = _lib.lib()
    save_file({"a": torch.randn(3, 8).to(torch.bfloat16), "b.c": torch.randn(17), "e": torch.arange(5)},
              str(tmp_path / "ok.safetensors"), metadata={"format": "pt"})
    blob = (tmp_path / "ok.safetensors").read_bytes()
    hl = int.from_bytes(blob[:8], "little")
    cfg = O.MODEL_CONFIGS["qwen3-tiny"]
    _write_config(tmp_path, cfg, head_dim=64, rope_theta=1e6, tie_word_embeddings=False)
    text = (tmp_path / "config.json").read_bytes()
    numbers = ["0", "-1", "99999999999999999999999", "18446744073709551615", "9223372036854775807", "1e9", "4294967296",
               "2147483648", "0.5", "1e400", "-1e400", "01", "1.", "1e", "NaN"]
    rng = random.Random(1234)

    def mutate(b, lo, hi):
        b = bytearray(b)
        kind = rng.randrange(4)
        if kind == 0:
            for _ in range(rng.randrange(1, 6)):
                b[lo + rng.randrange(hi - lo)] = rng.randrange(256)
        elif kind == 1:
            b = b[:rng.randrange(len(b))]
        elif kind == 2:
            body = bytes(b[lo:hi]).decode("latin1")
            m = rng.choice(list(re.finditer(r"\d+", body)))
            body = (body[:m.start()] + rng.choice(numbers + [str(rng.randrange(1000))]) + body[m.end():]).encode("latin1")
            b = bytearray(bytes(b[:lo]) if lo == 0 else len(body).to_bytes(8, "little")) + body + b[hi:]
        else:
            for _ in range(rng.randrange(1, 4)):
                b[lo + rng.randrange(hi - lo)] = ord(rng.choice('{}[]",:\\ \x00'))
        return bytes(b)

    opened = 0
    for it in range(1500):
        if it % 10 == 9:
            b = bytearray(blob)
            b[:8] = rng.choice([0, 1, hl - 1, hl + 1, len(b), len(b) - 8, 2**31, 2**63, 2**64 - 1, rng.randrange(2**64)]).to_bytes(8, "little")
            b = bytes(b)
        else:
            b = mutate(blob, 8, 8 + hl)
        p = tmp_path / "m.safetensors"
        p.write_bytes(b)
        l2, h, st = _open(p)
        if st != 0:
            assert not h.value and l.nvc_hf_last_error() != b""
            continue
        opened += 1
        for name, (dt, shape) in _infos(l, h).items():
            n = int(np.prod(shape, dtype=np.float64))
            if 0 <= n <= 1 << 16:
                buf = np.empty(max(n, 1), dtype=np.uint16)
                l.nvc_safetensors_read_bf16(h, name.encode("utf-8", "surrogateescape"), buf.ctypes.data_as(C.c_void_p), n)
        l.nvc_safetensors_close(h)
    assert opened > 20                                      # the mutations are not all trivially rejected
    d = tmp_path / "cfg"
    d.mkdir()
    accepted = 0
    for it in range(1500):
        (d / "config.json").write_bytes(mutate(text, 0, len(text)))
        c, tie = eng.Qwen3ConfigC(), C.c_int(0)
        st = l.nvc_qwen3_config_from_hf_dir(os.fsencode(str(d)), C.byref(c), C.byref(tie))
        if st != 0:
            assert l.nvc_hf_last_error() != b""
            continue
        accepted += 1
        json.loads((d / "config.json").read_bytes().decode("utf-8"))        # whatever we accept is valid JSON
        assert min(c.vocab_size, c.hidden_size, c.num_hidden_layers, c.num_key_value_heads, c.intermediate_size) >= 0
        assert c.num_attention_heads > 0 and c.head_dim >= 0
    assert accepted > 20


def test_sharded_index_refuses_paths_outside_the_model_dir(tmp_path):
    """weight_map file names are joined onto the model directory: '../x', 'a/b' and absolute names are refused before
    any file is opened (no runner or device needed: the check precedes the first nvc_runner_load_weight)."""
    l = _lib.lib()
    d = tmp_path / "model"
    d.mkdir()
    save_file({"model.norm.weight": torch.ones(4)}, str(tmp_path / "outside.safetensors"))
    for bad in ("../outside.safetensors", "sub/x.safetensors", "/etc/passwd", "..", ""):
        json.dump({"weight_map": {"model.norm.weight": bad}}, open(d / "model.safetensors.index.json", "w"))
        st = l.nvc_runner_load_hf_dir(C.c_void_p(1), os.fsencode(str(d)))       # the handle is never dereferenced on this path
        assert st != 0 and "plain file name" in l.nvc_hf_last_error().decode(), bad


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["single_bf16", "sharded_f32_tied", "single_f16"])
def test_runner_from_hf_dir_matches_from_weights(ctx, tmp_path, layout):
    """Same synthetic checkpoint through the native loader (three on-disk layouts) and through
    load_weight: identical greedy tokens and bit-identical prefill logits."""
    cfg = O.MODEL_CONFIGS["qwen3-tiny"]
    w = synth.make_weights(cfg, seed=3, norm_jitter=0.1)
    tied = layout == "sharded_f32_tied"
    if tied:
        w["lm_head.weight"] = w["model.embed_tokens.weight"].clone()
    _write_config(tmp_path, cfg, head_dim=cfg.head_dim, tie_word_embeddings=tied)
    if layout == "single_bf16":
        d = {k: v.to(torch.bfloat16).contiguous() for k, v in w.items()}
        d["model.layers.0.mlp.some_export_buffer"] = torch.ones(3)            # not a model tensor: skipped with a warning
        save_file(d, str(tmp_path / "model.safetensors"))
    elif layout == "single_f16":
        w = {k: v.to(torch.float16) for k, v in w.items()}
        save_file({k: v.contiguous() for k, v in w.items()}, str(tmp_path / "model.safetensors"))
    else:
        names = [k for k in w if k != "lm_head.weight"]                      # tied: no lm_head.weight on disk
        shards = {"model-00001-of-00002.safetensors": names[::2], "model-00002-of-00002.safetensors": names[1::2]}
        wm = {}
        for fn, ks in shards.items():
            # bf16-representable values stored as f32, plus a legacy non-weight buffer that must be skipped
            d = {k: w[k].to(torch.bfloat16).float().contiguous() for k in ks}
            d["model.layers.0.self_attn.rotary_emb.inv_freq"] = torch.ones(4)
            save_file(d, str(tmp_path / fn))
            wm.update({k: fn for k in ks})
        json.dump({"metadata": {"total_size": 0}, "weight_map": wm}, open(tmp_path / "model.safetensors.index.json", "w"))
    kw = dict(num_kvcache_blocks=16, kvcache_block_size=16, max_num_seqs=2, max_num_batched_tokens=256, max_model_len=256)
    a = eng.Qwen3ModelRunner.from_hf_dir(ctx, str(tmp_path), **kw)
    warn = a.l.nvc_hf_last_warning().decode()
    assert ("some_export_buffer" in warn) == (layout == "single_bf16"), warn
    b = eng.Qwen3ModelRunner.from_weights(ctx, cfg, w.items(), **kw)
    try:
        prompt = synth.make_prompts(1, 37, cfg.vocab_size, 5)[0]
        outs = []
        for r in (a, b):
            e = eng.LLMEngine(eng.Scheduler(eng.SchedulerConfig(2, 256, cfg.eos_token_id, 16, 16)), r, greedy=True)
            e.add_request(prompt, eng.SamplingParams(1.0, 6, True))
            si = e.schedule()
            tok = r.prefill(si.ids, si.positions, si.cu_seqlens, si.slot_mapping)
            lg = r.last_logits(1).copy()
            e.post_process(tok)
            while not e.is_finished():
                e.step()
            outs.append((lg, e.output_of(0)[0]))
        assert np.array_equal(outs[0][0], outs[1][0]) and outs[0][1] == outs[1][1]
    finally:
        a.close()
        b.close()
