"""CPU-side checks of the C-ABI boundary: the library loads, exports every symbol include/kge.h
declares, and refuses compute without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from oracle import model as om

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "kge.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(kge_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_bindings_agree(kge):
    hs = set(header_symbols())
    assert hs == set(kge.SIGNATURES), hs ^ set(kge.SIGNATURES)


def test_library_exports_every_declared_symbol(kge):
    L = kge.lib()
    for name in header_symbols():
        assert hasattr(L, name), name
    assert L.kge_abi_version() == 1


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(kge, tmp_path):
    assert kge.lib().kge_device_sm_count() == 0
    with pytest.raises(kge.KgeError):
        kge.require_cuda()
    cfg = om.CONFIGS["tiny-llama"]
    (tmp_path / "config.json").write_text(__import__("json").dumps(cfg.to_hf()))
    h = C.c_void_p()
    ec = kge.EngineConfig(8, 64, 128, 16, 8, 0.9, 0, 1, 1, 0, 0)
    rc = kge.lib().kge_engine_create(C.byref(h), str(tmp_path).encode(), C.byref(ec), 0)
    assert rc == -2 and b"no CUDA device" in kge.lib().kge_last_error()


@pytest.mark.parametrize("name", ["llama-3.2-1b", "llama-3.1-8b"])
def test_rope_cache_host_matches_oracle(kge, name):
    """kge_rope_cache_host is host arithmetic, so it is checked here against the oracle (itself
    bit-equal to vLLM's Llama3RotaryEmbedding on the golden rows)."""
    c = om.CONFIGS[name]
    rs = c.rope_scaling
    n = 4096
    out = np.zeros((n, c.head_dim), dtype=np.uint16)
    rc = kge.lib().kge_rope_cache_host(out.ctypes.data, n, c.head_dim, c.rope_theta, rs["factor"],
                                       rs["low_freq_factor"], rs["high_freq_factor"],
                                       rs["original_max_position_embeddings"])
    assert rc == 0
    ref = om.rope_cos_sin_cache(c, n, torch.bfloat16).view(torch.int16).numpy().astype(np.int32) & 0xFFFF
    got = out.astype(np.int32)
    diff = np.abs(got - ref)
    # fp32 libm differences can move a value across a bf16 rounding boundary: allow 1 ulp on <0.1 %
    assert diff.max() <= 1, diff.max()
    assert (diff != 0).mean() < 1e-3, (diff != 0).mean()


def test_rope_cache_host_float16(kge):
    """The float16 twin rounds the fp32 cos/sin to IEEE half (base.py:58-61 casts the cache to the model dtype)."""
    c = om.CONFIGS["llama-3.1-8b"]
    rs = c.rope_scaling
    n = 2048
    out = np.zeros((n, c.head_dim), dtype=np.uint16)
    rc = kge.lib().kge_rope_cache_host_f16(out.ctypes.data, n, c.head_dim, c.rope_theta, rs["factor"], rs["low_freq_factor"],
                                           rs["high_freq_factor"], rs["original_max_position_embeddings"])
    assert rc == 0
    ref = om.rope_cos_sin_cache(c, n, torch.float16).view(torch.int16).numpy().astype(np.int32) & 0xFFFF
    diff = np.abs(out.astype(np.int32) - ref)
    # +-0 differ in the sign bit only; fp32 libm differences can cross a rounding boundary: 1 ulp on < 0.1 %
    diff[(out.astype(np.int32) & 0x7FFF) + (ref & 0x7FFF) == 0] = 0
    assert diff.max() <= 1 and (diff != 0).mean() < 1e-3, (diff.max(), (diff != 0).mean())
