"""The oracle against the committed golden vectors (tests/golden/make_golden.py):
vLLM's own pure-PyTorch ops and HF transformers models.  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import model as om
from oracle import sampler as osamp

BF = torch.bfloat16


def from_bits(a):
    return torch.from_numpy(a.copy()).view(BF)


@pytest.fixture(scope="module")
def ops(golden_dir):
    return np.load(os.path.join(golden_dir, "ops_vllm.npz"))


@pytest.fixture(scope="module")
def hf(golden_dir):
    return np.load(os.path.join(golden_dir, "hf_models.npz"))


@pytest.mark.parametrize("d", [768, 2048])
def test_rms_norm_bit_exact(ops, d):
    x, r, w = (from_bits(ops[f"rms_{n}_{d}"]) for n in "xrw")
    assert torch.equal(om.rms_norm(x, w, 1e-5), from_bits(ops[f"rms_y_{d}"]))
    y, res = om.fused_add_rms_norm(x, r, w, 1e-5)
    assert torch.equal(y, from_bits(ops[f"frms_y_{d}"]))
    assert torch.equal(res, from_bits(ops[f"frms_res_{d}"]))


def test_silu_mul_bit_exact(ops):
    assert torch.equal(om.silu_and_mul(from_bits(ops["silu_x"])), from_bits(ops["silu_y"]))


@pytest.mark.parametrize("name", ["llama-3.2-1b", "llama-3.1-8b"])
def test_rope_bit_exact(ops, name):
    c = om.CONFIGS[name]
    inv = om.rope_inv_freq(c)
    assert torch.equal(inv, torch.from_numpy(ops[f"rope_inv_{name}"]))
    pos = torch.from_numpy(ops[f"rope_pos_{name}"])
    cache = om.rope_cos_sin_cache(c, 4096, BF)
    rows = cache.index_select(0, pos)
    assert torch.equal(rows, from_bits(ops[f"rope_cache_rows_{name}"]))
    cos, sin = rows.chunk(2, dim=-1)
    hd = c.head_dim
    q = from_bits(ops[f"rope_q_{name}"]).view(7, -1, hd)
    k = from_bits(ops[f"rope_k_{name}"]).view(7, -1, hd)
    assert torch.equal(om.apply_rope_neox(q, cos, sin).reshape(7, -1), from_bits(ops[f"rope_qo_{name}"]))
    assert torch.equal(om.apply_rope_neox(k, cos, sin).reshape(7, -1), from_bits(ops[f"rope_ko_{name}"]))


def test_top_k_top_p_mask(ops):
    lg = torch.from_numpy(ops["tkp_logits"])
    got = osamp.apply_top_k_top_p(lg, torch.from_numpy(ops["tkp_k"]), torch.from_numpy(ops["tkp_p"]))
    assert torch.equal(got, torch.from_numpy(ops["tkp_masked"]))


@pytest.mark.parametrize("name", ["tiny-opt", "tiny-llama", "tiny-llama-hd128", "tiny-qwen2", "tiny-qwen3"])
def test_model_vs_hf_fp32(hf, name):
    """fp32 oracle == HF fp32 to float tolerance: pins graph structure, weight naming,
    fused-row order, OPT position offset and llama3 RoPE scaling."""
    c = om.CONFIGS[name]
    m = om.OracleModel(c, om.make_weights(c, 0, torch.float32), torch.float32)
    prompt = hf[f"{name}_prompt"].tolist()
    kv = m.new_kv(64)
    h = m.forward(torch.tensor(prompt), [kv], [0, len(prompt)])
    lg = m.logits(h)
    ref = torch.from_numpy(hf[f"{name}_logits"])
    assert torch.allclose(lg, ref, atol=2e-4, rtol=2e-4), (lg - ref).abs().max()
    ids = m.generate_greedy([prompt], 8)[0]
    assert ids == hf[f"{name}_greedy"].tolist()


@pytest.mark.parametrize("name", ["tiny-opt", "tiny-llama"])
def test_bf16_oracle_close_to_fp32(name):
    """bf16 tolerance used throughout: vLLM IR default atol 1e-3 / rtol 1.6e-2 per op
    (vllm:ir/tolerances.py:12-36); end-to-end logits are compared at atol 3e-2, rtol 5e-2."""
    c = om.CONFIGS[name]
    w = om.make_weights(c, 0, torch.float32)
    wb = {k: v.to(BF) for k, v in w.items()}
    m32 = om.OracleModel(c, {k: v.float() for k, v in wb.items()}, torch.float32)
    m16 = om.OracleModel(c, wb, BF)
    prompt = om.make_prompts(c, 1, 16, 5)[0]
    l32 = m32.logits(m32.forward(torch.tensor(prompt), [m32.new_kv(32)], [0, 16]))
    l16 = m16.logits(m16.forward(torch.tensor(prompt), [m16.new_kv(32)], [0, 16])).float()
    assert torch.allclose(l16, l32, atol=3e-2, rtol=5e-2), (l16 - l32).abs().max()


def test_chunked_prefill_equals_whole():
    """Chunking a prompt across steps must not change values (fp32, exact to float tol)."""
    c = om.CONFIGS["tiny-llama"]
    m = om.OracleModel(c, om.make_weights(c, 0, torch.float32), torch.float32)
    p = om.make_prompts(c, 1, 20, 1)[0]
    kv = m.new_kv(32)
    whole = m.forward(torch.tensor(p), [kv], [0, 20])
    kv2 = m.new_kv(32)
    a = m.forward(torch.tensor(p[:13]), [kv2], [0, 13])
    b = m.forward(torch.tensor(p[13:]), [kv2], [0, 7])
    assert torch.allclose(torch.cat([a, b]), whole, atol=1e-5, rtol=1e-5)
