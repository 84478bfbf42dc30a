"""Generate tests/golden/*.npz.  Runs ONLY in the build container (needs the installed
vLLM 0.22.0 and HF transformers 5.5; neither is read at test time).

  python tests/golden/make_golden.py

What it pins (the reference itself has no fixtures):
  ops_vllm.npz   outputs of vLLM's own importable pure-PyTorch functions on seeded inputs:
                 rms_norm / fused_add_rms_norm (vllm/ir/ops/layernorm.py), SiluAndMul.forward_native,
                 RotaryEmbedding.forward_static with Llama3RotaryEmbedding's cos/sin cache,
                 apply_top_k_top_p_pytorch.
  hf_models.npz  fp32 logits + greedy ids of HF OPTForCausalLM / LlamaForCausalLM built from the
                 oracle's config.json + seeded weights (tiny shapes) - an independent second
                 opinion on graph structure, weight naming and RoPE scaling.
bf16 tensors are stored as their int16 bit patterns.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import model as om  # noqa: E402


def bits(t):
    return t.contiguous().view(torch.int16).numpy()


def ops_vllm():
    from vllm.ir.ops.layernorm import fused_add_rms_norm, rms_norm
    from vllm.model_executor.layers.activation import SiluAndMul
    from vllm.model_executor.layers.rotary_embedding.base import RotaryEmbedding
    from vllm.model_executor.layers.rotary_embedding.llama3_rope import Llama3RotaryEmbedding
    from vllm.v1.sample.ops.topk_topp_sampler import apply_top_k_top_p_pytorch

    out = {}
    g = torch.Generator().manual_seed(1234)
    bf = torch.bfloat16
    for d in (768, 2048):
        x = torch.randn(7, d, generator=g).to(bf)
        r = torch.randn(7, d, generator=g).to(bf)
        w = (1 + 0.1 * torch.randn(d, generator=g)).to(bf)
        out[f"rms_x_{d}"], out[f"rms_r_{d}"], out[f"rms_w_{d}"] = bits(x), bits(r), bits(w)
        out[f"rms_y_{d}"] = bits(rms_norm(x, w, 1e-5))
        y, ro = fused_add_rms_norm(x.clone(), r.clone(), w, 1e-5)
        out[f"frms_y_{d}"], out[f"frms_res_{d}"] = bits(y), bits(ro)
    x = torch.randn(5, 2 * 512, generator=g).to(bf) * 2
    out["silu_x"] = bits(x)
    out["silu_y"] = bits(SiluAndMul.forward_native(x))

    for name in ("llama-3.2-1b", "llama-3.1-8b"):
        c = om.CONFIGS[name]
        rs = c.rope_scaling
        obj = object.__new__(Llama3RotaryEmbedding)
        for k, v in dict(head_size=c.head_dim, rotary_dim=c.head_dim, base=c.rope_theta,
                         max_position_embeddings=4096, is_neox_style=True,
                         scaling_factor=rs["factor"], low_freq_factor=rs["low_freq_factor"],
                         high_freq_factor=rs["high_freq_factor"],
                         orig_max_position=rs["original_max_position_embeddings"]).items():
            object.__setattr__(obj, k, v)
        inv = Llama3RotaryEmbedding._compute_inv_freq(obj, c.rope_theta)
        cache = Llama3RotaryEmbedding._compute_cos_sin_cache(obj).to(bf)
        out[f"rope_inv_{name}"] = inv.numpy()
        pos = torch.tensor([0, 1, 2, 17, 255, 1024, 4095])
        out[f"rope_pos_{name}"] = pos.numpy()
        out[f"rope_cache_rows_{name}"] = bits(cache.index_select(0, pos))
        q = torch.randn(7, 4 * c.head_dim, generator=g).to(bf)
        k = torch.randn(7, 2 * c.head_dim, generator=g).to(bf)
        qo, ko = RotaryEmbedding.forward_static(pos, q.clone(), k.clone(), c.head_dim,
                                                c.head_dim, cache, True)
        out[f"rope_q_{name}"], out[f"rope_k_{name}"] = bits(q), bits(k)
        out[f"rope_qo_{name}"], out[f"rope_ko_{name}"] = bits(qo), bits(ko)

    lg = torch.randn(6, 1000, generator=g) * 3
    k = torch.tensor([1000, 5, 50, 1000, 1, 20])
    p = torch.tensor([1.0, 1.0, 0.9, 0.5, 0.3, 0.95])
    out["tkp_logits"], out["tkp_k"], out["tkp_p"] = lg.numpy(), k.numpy(), p.numpy()
    out["tkp_masked"] = apply_top_k_top_p_pytorch(lg.clone(), k.clone(), p.clone()).numpy()
    np.savez_compressed(os.path.join(HERE, "ops_vllm.npz"), **out)
    print("ops_vllm.npz", len(out), "arrays")


def hf_models():
    from transformers import (LlamaConfig, LlamaForCausalLM, OPTConfig, OPTForCausalLM, Qwen2Config, Qwen2ForCausalLM,
                              Qwen3Config, Qwen3ForCausalLM)
    out = {}
    for name in ("tiny-opt", "tiny-llama", "tiny-llama-hd128", "tiny-qwen2", "tiny-qwen3"):
        c = om.CONFIGS[name]
        w = om.make_weights(c, seed=0, dtype=torch.float32)
        hf = c.to_hf()
        hf.pop("torch_dtype")
        if c.model_type == "opt":
            m = OPTForCausalLM(OPTConfig(**hf))
        elif c.model_type == "qwen2":
            m = Qwen2ForCausalLM(Qwen2Config(**hf))
        elif c.model_type == "qwen3":
            m = Qwen3ForCausalLM(Qwen3Config(**hf))
        else:
            m = LlamaForCausalLM(LlamaConfig(**hf))
        sd = dict(w)
        if c.tied:
            sd["lm_head.weight"] = (w["model.decoder.embed_tokens.weight"] if c.model_type == "opt"
                                    else w["model.embed_tokens.weight"])
        missing, unexpected = m.load_state_dict(sd, strict=False)
        assert not unexpected, unexpected
        assert all("rotary" in k or "inv_freq" in k for k in missing), missing
        m = m.float().eval()
        prompt = om.make_prompts(c, 1, 24, seed=3)[0]
        ids = torch.tensor([prompt])
        with torch.no_grad():
            lg = m(ids).logits[0]
            gen = m.generate(ids, max_new_tokens=8, do_sample=False, pad_token_id=0,
                             eos_token_id=None)[0, len(prompt):]
        out[f"{name}_prompt"] = np.array(prompt, dtype=np.int64)
        out[f"{name}_logits"] = lg.numpy().astype(np.float32)
        out[f"{name}_greedy"] = gen.numpy().astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "hf_models.npz"), **out)
    print("hf_models.npz", len(out), "arrays")


if __name__ == "__main__":
    torch.manual_seed(0)
    ops_vllm()
    hf_models()
