"""Model-directory reader of the engine (csrc/loader.cc) through its host-only C view: config.json of every supported
family parsed into the right description, safetensors shards in BF16 / F16 / F32 read back as the bf16 bits torch
produces (round to nearest even, inf / nan / subnormals included), row / column slices as the tensor-parallel loader
takes them, and clear errors.  CPU only."""
import ctypes as C
import json
import os

import numpy as np
import pytest
import torch
from safetensors.torch import save_file

from oracle import model as om


def open_dir(kge, d):
    h = kge.lib().kge_model_open(str(d).encode())
    assert h, kge.lib().kge_last_error().decode()
    return h


def spec(kge, h):
    buf = C.create_string_buffer(4096)
    assert kge.lib().kge_model_spec_json(h, buf, 4096) > 0
    return json.loads(buf.value)


def read(kge, h, name, r0=0, r1=-1, c0=0, c1=-1, cap=1 << 20):
    out = np.zeros(cap, dtype=np.uint16)
    n = kge.lib().kge_model_read_bf16(h, name.encode(), r0, r1, c0, c1, out.ctypes.data, cap)
    return None if n < 0 else out[:n]


def bits(t):
    return t.to(torch.bfloat16).contiguous().view(torch.int16).numpy().astype(np.uint16).ravel()


@pytest.mark.parametrize("name", ["tiny-opt", "tiny-llama", "tiny-llama-hd128", "tiny-qwen2", "tiny-qwen3", "llama-3.1-8b", "opt-125m"])
def test_config_json_of_every_family(kge, tmp_path, name):
    c = om.CONFIGS[name]
    (tmp_path / "config.json").write_text(json.dumps(c.to_hf()))
    h = open_dir(kge, tmp_path)
    s = spec(kge, h)
    kge.lib().kge_model_close(h)
    assert s["model_type"] == c.model_type and not s["has_weights"]
    assert (s["vocab"], s["d"], s["L"], s["H"], s["KV"], s["hd"], s["ffn"]) == (c.vocab_size, c.hidden_size, c.num_layers, c.num_heads,
                                                                                c.num_kv_heads, c.head_dim, c.ffn)
    assert s["tied"] == c.tied and s["qkv_bias"] == c.qkv_bias and s["qk_norm"] == c.qk_norm
    if c.model_type != "opt":
        assert s["rope_theta"] == c.rope_theta and abs(s["eps"] - c.eps) < 1e-9
        if c.rope_scaling:
            assert (s["rope_factor"], s["rope_low"], s["rope_high"], s["rope_orig_max"]) == (
                c.rope_scaling["factor"], c.rope_scaling["low_freq_factor"], c.rope_scaling["high_freq_factor"],
                c.rope_scaling["original_max_position_embeddings"])
        else:
            assert s["rope_factor"] == 0


def test_shards_dtypes_and_slices(kge, tmp_path):
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    g = torch.Generator().manual_seed(0)
    a = torch.randn(37, 24, generator=g)
    special = torch.tensor([0.0, -0.0, float("inf"), float("-inf"), float("nan"), 1e-40, -3e-39, 65504.0, 1.0 + 2 ** -8, 1.0 + 2 ** -8 + 2 ** -20,
                            3.3895314e38, 1.0 + 3 * 2 ** -9], dtype=torch.float32)
    save_file({"w.bf16": a.to(torch.bfloat16), "w.f32": a.clone(), "special.f32": special}, str(tmp_path / "model-00001-of-00002.safetensors"))
    save_file({"w.f16": a.to(torch.float16), "vec.f16": a[0].to(torch.float16), "special.f16": special.to(torch.float16)},
              str(tmp_path / "model-00002-of-00002.safetensors"))
    h = open_dir(kge, tmp_path)
    assert spec(kge, h)["has_weights"]
    assert np.array_equal(read(kge, h, "w.bf16"), bits(a.to(torch.bfloat16)))
    assert np.array_equal(read(kge, h, "w.f32"), bits(a))                                   # fp32 -> bf16, round to nearest even
    assert np.array_equal(read(kge, h, "w.f16"), bits(a.to(torch.float16).float()))         # fp16 -> fp32 (exact) -> bf16
    assert np.array_equal(read(kge, h, "vec.f16", 3, 11), bits(a[0, 3:11].to(torch.float16).float()))
    # the slices the tensor-parallel loader takes: a row block (column-parallel) and a column block (row-parallel)
    assert np.array_equal(read(kge, h, "w.f32", 8, 20), bits(a[8:20]))
    assert np.array_equal(read(kge, h, "w.f32", 0, -1, 6, 18), bits(a[:, 6:18]))
    assert np.array_equal(read(kge, h, "w.f16", 5, 9, 2, 3), bits(a[5:9, 2:3].to(torch.float16).float()))
    got, ref = read(kge, h, "special.f32"), bits(special)
    nan = np.isnan(special.numpy())
    assert np.array_equal(got[~nan], ref[~nan]) and all((int(x) & 0x7F80) == 0x7F80 and (int(x) & 0x7F) for x in got[nan])
    got16, ref16 = read(kge, h, "special.f16"), bits(special.to(torch.float16).float())
    nan16 = np.isnan(special.to(torch.float16).float().numpy())
    assert np.array_equal(got16[~nan16], ref16[~nan16])
    # errors are reported, not swallowed
    assert read(kge, h, "missing") is None and "missing tensor" in kge.lib().kge_last_error().decode()
    assert read(kge, h, "w.f32", 0, 99) is None and "out of range" in kge.lib().kge_last_error().decode()
    assert read(kge, h, "w.f32", cap=10) is None and "capacity" in kge.lib().kge_last_error().decode()
    kge.lib().kge_model_close(h)


def test_bad_directories(kge, tmp_path):
    L = kge.lib()
    assert not L.kge_model_open(str(tmp_path / "nope").encode()) and "config.json" in L.kge_last_error().decode()
    (tmp_path / "config.json").write_text("{not json")
    assert not L.kge_model_open(str(tmp_path).encode()) and "config.json" in L.kge_last_error().decode()
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "mamba", "vocab_size": 10}))
    assert not L.kge_model_open(str(tmp_path).encode())
    (tmp_path / "config.json").write_text(json.dumps(om.CONFIGS["tiny-llama"].to_hf()))
    (tmp_path / "model.safetensors").write_bytes(b"\xff" * 64)
    assert not L.kge_model_open(str(tmp_path).encode()) and "header" in L.kge_last_error().decode()


def test_float16_view_of_every_source_dtype(kge, tmp_path):
    """kge_model_read_f16: what a float16 engine uploads.  BF16, F16 and F32 tensors converted to IEEE half with round to
    nearest even (torch's .to(float16) is the reference), including values that overflow to inf, subnormals and ties."""
    from safetensors.torch import save_file
    c = om.CONFIGS["tiny-llama"]
    g = torch.Generator().manual_seed(5)
    base = torch.randn(64, 48, generator=g)
    special = torch.tensor([0.0, -0.0, 65504.0, 65519.9, 65520.0, 1e6, -1e6, 6.1e-5, 5.96e-8, 2.98e-8, 2.99e-8, 1e-9, 1.0009765625, 1.00048828125,
                            1.0014648438, float("inf"), -float("inf")])
    f32 = torch.cat([base.flatten(), special, base.flatten()[: 48 - special.numel() % 48]]).reshape(-1, 48)
    (tmp_path / "config.json").write_text(json.dumps(c.to_hf()))
    save_file({"a.f32": f32.contiguous(), "a.f16": f32.to(torch.float16).contiguous(), "a.bf16": f32.to(torch.bfloat16).contiguous()},
              str(tmp_path / "model.safetensors"))
    h = open_dir(kge, tmp_path)
    cap = f32.numel()
    for name, src in (("a.f32", f32), ("a.f16", f32.to(torch.float16)), ("a.bf16", f32.to(torch.bfloat16))):
        out = np.zeros(cap, dtype=np.uint16)
        n = kge.lib().kge_model_read_f16(h, name.encode(), 0, -1, 0, -1, out.ctypes.data, cap)
        assert n == cap
        ref = src.to(torch.float16).contiguous().view(torch.int16).numpy().astype(np.uint16).ravel()
        assert np.array_equal(out, ref), (name, np.nonzero(out != ref)[0][:5])
    kge.lib().kge_model_close(h)
