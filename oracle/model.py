"""Model-side oracle: configs, seeded weights, and the decoder forward pass.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Restates, op by op and cast by cast:
  - OPT      vllm:model_executor/models/opt.py:61-70,120-124,170-197,268-295,384-385
  - Llama    vllm:model_executor/models/llama.py:94-121,223-233,316-333,395-434
  - Qwen2    vllm:model_executor/models/qwen2.py:160-175 (Llama graph + q/k/v bias)
  - Qwen3    vllm:model_executor/models/qwen3.py:142-160 (Llama graph + per-head RMSNorm of q and k before RoPE)
  - RMSNorm  vllm:ir/ops/layernorm.py:10-21,40-58
  - SiLU*mul vllm:model_executor/layers/activation.py:137-141
  - RoPE     vllm:model_executor/layers/rotary_embedding/base.py:69-93,140-184
             vllm:.../rotary_embedding/common.py:143-183, llama3_rope.py:33-54
  - logits   vllm:model_executor/layers/logits_processor.py:54-99 (model dtype, last rows)
Attention semantics (causal, GQA, scale hd^-0.5, fp32 softmax) follow
vllm:v1/attention/backends/cpu_attn.py:292-398.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------- configs


@dataclass
class ModelConfig:
    model_type: str  # "opt" | "llama"
    vocab_size: int
    hidden_size: int
    num_layers: int
    num_heads: int
    num_kv_heads: int
    head_dim: int
    ffn: int
    max_position: int
    eps: float = 1e-5
    tied: bool = True
    rope_theta: float = 500000.0
    rope_scaling: dict | None = None
    bos: int = 0
    eos: int = 0
    name: str = ""
    qkv_bias: bool = False      # qwen2
    qk_norm: bool = False       # qwen3

    def to_hf(self) -> dict:
        """HF-layout config.json (the same file the product's loader reads)."""
        if self.model_type == "opt":
            return {
                "model_type": "opt", "architectures": ["OPTForCausalLM"],
                "vocab_size": self.vocab_size, "hidden_size": self.hidden_size,
                "num_hidden_layers": self.num_layers, "ffn_dim": self.ffn,
                "num_attention_heads": self.num_heads,
                "max_position_embeddings": self.max_position,
                "do_layer_norm_before": True, "word_embed_proj_dim": self.hidden_size,
                "enable_bias": True, "layer_norm_elementwise_affine": True,
                "activation_function": "relu", "tie_word_embeddings": True,
                "bos_token_id": self.bos, "eos_token_id": self.eos, "pad_token_id": 1,
                "torch_dtype": "bfloat16",
            }
        arch = {"llama": "LlamaForCausalLM", "qwen2": "Qwen2ForCausalLM", "qwen3": "Qwen3ForCausalLM"}[self.model_type]
        d = {
            "model_type": self.model_type, "architectures": [arch],
            "vocab_size": self.vocab_size, "hidden_size": self.hidden_size,
            "num_hidden_layers": self.num_layers, "intermediate_size": self.ffn,
            "num_attention_heads": self.num_heads, "num_key_value_heads": self.num_kv_heads,
            "head_dim": self.head_dim, "rms_norm_eps": self.eps,
            "rope_theta": self.rope_theta, "max_position_embeddings": self.max_position,
            "tie_word_embeddings": self.tied, "hidden_act": "silu",
            "bos_token_id": self.bos, "eos_token_id": self.eos,
            "torch_dtype": "bfloat16", "attention_bias": False, "mlp_bias": False,
        }
        if self.model_type != "llama":
            d.pop("attention_bias"); d.pop("mlp_bias")
        if self.rope_scaling:
            d["rope_scaling"] = dict(self.rope_scaling)
        return d


def _llama3_scaling(factor: float) -> dict:
    return {"factor": factor, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
            "original_max_position_embeddings": 8192, "rope_type": "llama3"}


# Published hyper-parameters; the reference holds none of these.
CONFIGS: dict[str, ModelConfig] = {
    "opt-125m": ModelConfig("opt", 50272, 768, 12, 12, 12, 64, 3072, 2048, 1e-5, True,
                            bos=2, eos=2, name="facebook/opt-125m"),
    "llama-3.2-1b": ModelConfig("llama", 128256, 2048, 16, 32, 8, 64, 8192, 131072, 1e-5, True,
                                500000.0, _llama3_scaling(32.0), 128000, 128001,
                                "meta-llama/Llama-3.2-1B"),
    "llama-3.1-8b": ModelConfig("llama", 128256, 4096, 32, 32, 8, 128, 14336, 131072, 1e-5, False,
                                500000.0, _llama3_scaling(8.0), 128000, 128001,
                                "meta-llama/Llama-3.1-8B"),
    "llama-3.1-70b": ModelConfig("llama", 128256, 8192, 80, 64, 8, 128, 28672, 131072, 1e-5, False,
                                 500000.0, _llama3_scaling(8.0), 128000, 128001,
                                 "meta-llama/Llama-3.1-70B"),
    # Small shapes of the same two architectures, sized so the oracle finishes in seconds.
    "tiny-opt": ModelConfig("opt", 1000, 256, 2, 4, 4, 64, 512, 256, 1e-5, True,
                            bos=2, eos=2, name="tiny-opt"),
    "tiny-llama": ModelConfig("llama", 1024, 256, 2, 4, 2, 64, 512, 16384, 1e-5, False,
                              500000.0, _llama3_scaling(8.0), 1, 2, "tiny-llama"),
    "tiny-llama-hd128": ModelConfig("llama", 2048, 512, 3, 4, 1, 128, 1024, 16384, 1e-5, True,
                                    500000.0, _llama3_scaling(32.0), 1, 2, "tiny-llama-hd128"),
    # the model families the reference's values files deploy (Qwen2.5 / Qwen3), tiny shapes
    "tiny-qwen2": ModelConfig("qwen2", 1024, 256, 2, 4, 2, 64, 512, 4096, 1e-6, True, 1000000.0, None, 1, 2,
                              "tiny-qwen2", qkv_bias=True),
    "tiny-qwen3": ModelConfig("qwen3", 1024, 256, 2, 2, 1, 128, 512, 4096, 1e-6, True, 1000000.0, None, 1, 2,
                              "tiny-qwen3", qk_norm=True),
    "tiny-llama-tp": ModelConfig("llama", 1024, 512, 2, 8, 4, 64, 1024, 16384, 1e-5, False,
                                 500000.0, _llama3_scaling(8.0), 1, 2, "tiny-llama-tp (8 heads / 4 kv heads for TP 2 and 4)"),
    "small-opt": ModelConfig("opt", 50272, 768, 2, 12, 12, 64, 3072, 2048, 1e-5, True,
                             bos=2, eos=2, name="small-opt (opt-125m shape, 2 layers)"),
    "small-llama": ModelConfig("llama", 128256, 2048, 2, 32, 8, 64, 8192, 131072, 1e-5, True,
                               500000.0, _llama3_scaling(32.0), 128000, 128001,
                               "small-llama (llama-3.2-1b shape, 2 layers)"),
}


def config_from_hf(d: dict) -> ModelConfig:
    mt = d["model_type"]
    if mt == "opt":
        return ModelConfig("opt", d["vocab_size"], d["hidden_size"], d["num_hidden_layers"],
                           d["num_attention_heads"], d["num_attention_heads"],
                           d["hidden_size"] // d["num_attention_heads"], d["ffn_dim"],
                           d["max_position_embeddings"], 1e-5, True,
                           bos=d.get("bos_token_id", 2), eos=d.get("eos_token_id", 2))
    if mt in ("llama", "qwen2", "qwen3"):
        hd = d.get("head_dim") or d["hidden_size"] // d["num_attention_heads"]
        return ModelConfig(mt, d["vocab_size"], d["hidden_size"], d["num_hidden_layers"],
                           d["num_attention_heads"], d.get("num_key_value_heads", d["num_attention_heads"]),
                           hd, d["intermediate_size"], d["max_position_embeddings"],
                           d.get("rms_norm_eps", 1e-5), d.get("tie_word_embeddings", False),
                           d.get("rope_theta", 10000.0 if mt == "llama" else 1000000.0), d.get("rope_scaling"),
                           d.get("bos_token_id", 1), d.get("eos_token_id", 2),
                           qkv_bias=(mt == "qwen2"), qk_norm=(mt == "qwen3"))
    raise ValueError(f"unsupported model_type {mt}")


# --------------------------------------------------------------------------- weights


def make_weights(cfg: ModelConfig, seed: int = 0, dtype=torch.bfloat16,
                 std: float = 0.02) -> dict[str, torch.Tensor]:
    """Seeded HF-named weights.  Matrices N(0, std); norm weights 1 + N(0, 0.1);
    biases N(0, std).  (vLLM's dummy init U(+-1e-3) gives flat logits and meaningless
    argmax parity - SURVEY.md section 7 'hard parts' - so it is not used.)
    One CPU generator, fixed order, so the same (cfg, seed) gives the same bytes on
    every machine with the same torch."""
    g = torch.Generator().manual_seed(seed)
    d, f, L = cfg.hidden_size, cfg.ffn, cfg.num_layers

    def mat(*shape):
        return (torch.randn(*shape, generator=g) * std).to(dtype)

    def norm_w(n):
        return (1.0 + 0.1 * torch.randn(n, generator=g)).to(dtype)

    w: dict[str, torch.Tensor] = {}
    if cfg.model_type == "opt":
        p = "model.decoder."
        w[p + "embed_tokens.weight"] = mat(cfg.vocab_size, d)
        w[p + "embed_positions.weight"] = mat(cfg.max_position + 2, d)
        for i in range(L):
            q = f"{p}layers.{i}."
            for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
                w[q + f"self_attn.{nm}.weight"] = mat(d, d)
                w[q + f"self_attn.{nm}.bias"] = mat(d)
            w[q + "self_attn_layer_norm.weight"] = norm_w(d)
            w[q + "self_attn_layer_norm.bias"] = mat(d)
            w[q + "fc1.weight"] = mat(f, d)
            w[q + "fc1.bias"] = mat(f)
            w[q + "fc2.weight"] = mat(d, f)
            w[q + "fc2.bias"] = mat(d)
            w[q + "final_layer_norm.weight"] = norm_w(d)
            w[q + "final_layer_norm.bias"] = mat(d)
        w[p + "final_layer_norm.weight"] = norm_w(d)
        w[p + "final_layer_norm.bias"] = mat(d)
    else:
        H, KV, hd = cfg.num_heads, cfg.num_kv_heads, cfg.head_dim
        w["model.embed_tokens.weight"] = mat(cfg.vocab_size, d)
        for i in range(L):
            q = f"model.layers.{i}."
            w[q + "self_attn.q_proj.weight"] = mat(H * hd, d)
            w[q + "self_attn.k_proj.weight"] = mat(KV * hd, d)
            w[q + "self_attn.v_proj.weight"] = mat(KV * hd, d)
            w[q + "self_attn.o_proj.weight"] = mat(d, H * hd)
            if cfg.qkv_bias:
                w[q + "self_attn.q_proj.bias"] = mat(H * hd)
                w[q + "self_attn.k_proj.bias"] = mat(KV * hd)
                w[q + "self_attn.v_proj.bias"] = mat(KV * hd)
            if cfg.qk_norm:
                w[q + "self_attn.q_norm.weight"] = norm_w(hd)
                w[q + "self_attn.k_norm.weight"] = norm_w(hd)
            w[q + "mlp.gate_proj.weight"] = mat(f, d)
            w[q + "mlp.up_proj.weight"] = mat(f, d)
            w[q + "mlp.down_proj.weight"] = mat(d, f)
            w[q + "input_layernorm.weight"] = norm_w(d)
            w[q + "post_attention_layernorm.weight"] = norm_w(d)
        w["model.norm.weight"] = norm_w(d)
        if not cfg.tied:
            w["lm_head.weight"] = mat(cfg.vocab_size, d)
    return w


def make_timing_weights(cfg: ModelConfig, dtype=torch.bfloat16) -> dict[str, torch.Tensor]:
    """Weights for TIMING the oracle at full model size without holding the full model: one layer's
    tensors are generated (cheap uniform init, std 0.02) and every layer aliases them, so the
    arithmetic and the bytes streamed per layer are those of the real shape while host memory stays
    at one layer + embeddings.  Values are meaningless; never use for parity."""
    one = ModelConfig(**{**cfg.__dict__, "num_layers": 1})
    g = torch.Generator().manual_seed(0)
    a = 0.02 * 3 ** 0.5
    w1 = {}
    for k, v in make_shapes(one).items():
        w1[k] = torch.empty(v, dtype=dtype).uniform_(-a, a, generator=g) if "norm" not in k else torch.ones(v, dtype=dtype)
    w = {}
    for k, v in w1.items():
        if ".layers.0." in k:
            for i in range(cfg.num_layers):
                w[k.replace(".layers.0.", f".layers.{i}.")] = v
        else:
            w[k] = v
    return w


def make_shapes(cfg: ModelConfig) -> dict[str, tuple]:
    """Tensor name -> shape of the HF checkpoint for cfg (same names as make_weights)."""
    d, f, L = cfg.hidden_size, cfg.ffn, cfg.num_layers
    s: dict[str, tuple] = {}
    if cfg.model_type == "opt":
        p = "model.decoder."
        s[p + "embed_tokens.weight"] = (cfg.vocab_size, d)
        s[p + "embed_positions.weight"] = (cfg.max_position + 2, d)
        for i in range(L):
            q = f"{p}layers.{i}."
            for nm in ("q_proj", "k_proj", "v_proj", "out_proj"):
                s[q + f"self_attn.{nm}.weight"] = (d, d)
                s[q + f"self_attn.{nm}.bias"] = (d,)
            s[q + "self_attn_layer_norm.weight"] = (d,)
            s[q + "self_attn_layer_norm.bias"] = (d,)
            s[q + "fc1.weight"] = (f, d)
            s[q + "fc1.bias"] = (f,)
            s[q + "fc2.weight"] = (d, f)
            s[q + "fc2.bias"] = (d,)
            s[q + "final_layer_norm.weight"] = (d,)
            s[q + "final_layer_norm.bias"] = (d,)
        s[p + "final_layer_norm.weight"] = (d,)
        s[p + "final_layer_norm.bias"] = (d,)
    else:
        H, KV, hd = cfg.num_heads, cfg.num_kv_heads, cfg.head_dim
        s["model.embed_tokens.weight"] = (cfg.vocab_size, d)
        for i in range(L):
            q = f"model.layers.{i}."
            s[q + "self_attn.q_proj.weight"] = (H * hd, d)
            s[q + "self_attn.k_proj.weight"] = (KV * hd, d)
            s[q + "self_attn.v_proj.weight"] = (KV * hd, d)
            s[q + "self_attn.o_proj.weight"] = (d, H * hd)
            if cfg.qkv_bias:
                s[q + "self_attn.q_proj.bias"] = (H * hd,)
                s[q + "self_attn.k_proj.bias"] = (KV * hd,)
                s[q + "self_attn.v_proj.bias"] = (KV * hd,)
            if cfg.qk_norm:
                s[q + "self_attn.q_norm.weight"] = (hd,)
                s[q + "self_attn.k_norm.weight"] = (hd,)
            s[q + "mlp.gate_proj.weight"] = (f, d)
            s[q + "mlp.up_proj.weight"] = (f, d)
            s[q + "mlp.down_proj.weight"] = (d, f)
            s[q + "input_layernorm.weight"] = (d,)
            s[q + "post_attention_layernorm.weight"] = (d,)
        s["model.norm.weight"] = (d,)
        if not cfg.tied:
            s["lm_head.weight"] = (cfg.vocab_size, d)
    return s


def save_model_dir(path: str, cfg: ModelConfig, weights: dict[str, torch.Tensor], torch_dtype: str = "bfloat16") -> None:
    """Write an HF-layout directory: config.json + model.safetensors (the layout the
    reference mounts at /models/<name>, ref:values-01-minimal-example3.yaml:8,22-30).  torch_dtype is what config.json
    declares (what `--dtype auto` follows); the tensors are stored in whatever dtype `weights` holds."""
    from safetensors.torch import save_file
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as fh:
        json.dump(dict(cfg.to_hf(), torch_dtype=torch_dtype), fh, indent=1)
    save_file({k: v.contiguous() for k, v in weights.items()},
              os.path.join(path, "model.safetensors"))


def make_prompts(cfg: ModelConfig, n: int, length: int, seed: int = 0) -> list[list[int]]:
    """Token-id prompts ~ U[lo, hi) (no tokenizer offline)."""
    g = torch.Generator().manual_seed(seed)
    lo = 1000 if cfg.vocab_size > 4000 else 3
    hi = cfg.vocab_size - 1000 if cfg.vocab_size > 4000 else cfg.vocab_size
    return torch.randint(lo, hi, (n, length), generator=g).tolist()


# --------------------------------------------------------------------------- ops


def rms_norm(x, w, eps):
    """vllm:ir/ops/layernorm.py:10-21 - weight multiply happens in the weight dtype."""
    od = x.dtype
    x = x.to(torch.float32)
    var = x.pow(2).mean(dim=-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    x = x.to(w.dtype) * w
    return x.to(od)


def fused_add_rms_norm(x, res, w, eps):
    """vllm:ir/ops/layernorm.py:40-58 - residual is the rounded sum, norm uses the
    un-rounded fp32 sum."""
    od = x.dtype
    x = x.to(torch.float32) + res.to(torch.float32)
    res_out = x.to(od)
    var = x.pow(2).mean(dim=-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    x = x.to(w.dtype) * w
    return x.to(od), res_out


def layer_norm(x, w, b, eps=1e-5):
    """nn.LayerNorm as used by vllm:model_executor/models/opt.py:148-150,166-168:
    statistics and affine in fp32, one rounding to the model dtype."""
    od = x.dtype
    xf = x.to(torch.float32)
    mean = xf.mean(dim=-1, keepdim=True)
    var = (xf - mean).pow(2).mean(dim=-1, keepdim=True)
    y = (xf - mean) * torch.rsqrt(var + eps) * w.to(torch.float32) + b.to(torch.float32)
    return y.to(od)


def silu_and_mul(x):
    """vllm:model_executor/layers/activation.py:137-141."""
    d = x.shape[-1] // 2
    return F.silu(x[..., :d]) * x[..., d:]


def rope_inv_freq(cfg: ModelConfig) -> torch.Tensor:
    """base.py:69-81 then llama3_rope.py:33-54."""
    rd = cfg.head_dim
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, rd, 2, dtype=torch.float) / rd))
    rs = cfg.rope_scaling
    if not rs:
        return inv
    low, high = rs["low_freq_factor"], rs["high_freq_factor"]
    orig, factor = rs["original_max_position_embeddings"], rs["factor"]
    low_wl, high_wl = orig / low, orig / high
    wl = 2 * math.pi / inv
    smooth = (orig / wl - low) / (high - low) if low != high else 0
    return torch.where(wl < high_wl, inv,
                       torch.where(wl > low_wl, inv / factor,
                                   (1 - smooth) * inv / factor + smooth * inv))


def rope_cos_sin_cache(cfg: ModelConfig, max_pos: int, dtype) -> torch.Tensor:
    """base.py:83-93 and :58-61 - fp32 cos|sin, cast to the model dtype."""
    inv = rope_inv_freq(cfg)
    t = torch.arange(max_pos, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def apply_rope_neox(x, cos, sin):
    """common.py:160-183, neox (half-split) pairing; every op rounds in x.dtype.
    x [T, heads, hd]; cos/sin [T, hd/2]."""
    cos = cos.unsqueeze(-2).to(x.dtype)
    sin = sin.unsqueeze(-2).to(x.dtype)
    x1, x2 = torch.chunk(x, 2, dim=-1)
    o1 = x1 * cos - x2 * sin
    o2 = x2 * cos + x1 * sin
    return torch.cat((o1, o2), dim=-1)


def causal_attention(q, k, v, scale, q_pos0: int):
    """One request.  q [Tq,H,hd] for positions q_pos0..q_pos0+Tq-1; k,v [ctx,KV,hd] with
    ctx = q_pos0+Tq.  fp32 scores/softmax/PV, output rounded to the model dtype."""
    Tq, H, hd = q.shape
    ctx, KV, _ = k.shape
    g = H // KV
    qf = q.to(torch.float32).transpose(0, 1)                     # [H,Tq,hd]
    kf = k.to(torch.float32).transpose(0, 1).repeat_interleave(g, dim=0)  # [H,ctx,hd]
    vf = v.to(torch.float32).transpose(0, 1).repeat_interleave(g, dim=0)
    s = torch.matmul(qf, kf.transpose(1, 2)) * scale             # [H,Tq,ctx]
    qpos = torch.arange(q_pos0, q_pos0 + Tq).view(1, Tq, 1)
    kpos = torch.arange(ctx).view(1, 1, ctx)
    s = s.masked_fill(kpos > qpos, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = torch.matmul(p, vf)                                      # [H,Tq,hd]
    return o.transpose(0, 1).reshape(Tq, H * hd).to(q.dtype)


# --------------------------------------------------------------------------- model


@dataclass
class RequestKV:
    """Per-request contiguous KV (the oracle needs no paging: paging only changes where
    rows live, not their values)."""
    k: list = field(default_factory=list)   # per layer [cap,KV,hd]
    v: list = field(default_factory=list)
    length: int = 0


class OracleModel:
    def __init__(self, cfg: ModelConfig, weights: dict[str, torch.Tensor], dtype=torch.bfloat16,
                 max_pos: int | None = None):
        self.cfg, self.dtype = cfg, dtype
        self.w = {k: v.to(dtype) for k, v in weights.items()}
        c = cfg
        self.scale = c.head_dim ** -0.5
        L = c.num_layers
        _memo: dict = {}

        def cat(names):
            # layers may alias one storage (make_timing_weights): fuse each distinct set once
            key = tuple(self.w[n].data_ptr() for n in names)
            if key not in _memo:
                _memo[key] = torch.cat([self.w[n] for n in names], 0)
            return _memo[key]
        if c.model_type != "opt":
            self.cos_sin = rope_cos_sin_cache(c, max_pos or min(c.max_position, 8192), dtype)
            # fused row orders as in QKVParallelLinear / MergedColumnParallelLinear
            self.qkv = [cat([f"model.layers.{i}.self_attn.{n}_proj.weight" for n in "qkv"]) for i in range(L)]
            self.gate_up = [cat([f"model.layers.{i}.mlp.gate_proj.weight", f"model.layers.{i}.mlp.up_proj.weight"])
                            for i in range(L)]
            self.qkv_b = ([cat([f"model.layers.{i}.self_attn.{n}_proj.bias" for n in "qkv"]) for i in range(L)]
                          if c.qkv_bias else [None] * L)
            self.lm_head = self.w["model.embed_tokens.weight"] if c.tied else self.w["lm_head.weight"]
        else:
            p = "model.decoder.layers."
            self.qkv = [cat([f"{p}{i}.self_attn.{n}_proj.weight" for n in "qkv"]) for i in range(L)]
            self.qkv_b = [cat([f"{p}{i}.self_attn.{n}_proj.bias" for n in "qkv"]) for i in range(L)]
            self.lm_head = self.w["model.decoder.embed_tokens.weight"]

    # -- kv
    def new_kv(self, cap: int) -> RequestKV:
        c = self.cfg
        kv = RequestKV()
        for _ in range(c.num_layers):
            kv.k.append(torch.zeros(cap, c.num_kv_heads, c.head_dim, dtype=self.dtype))
            kv.v.append(torch.zeros(cap, c.num_kv_heads, c.head_dim, dtype=self.dtype))
        return kv

    # -- attention over the flattened batch
    def _attend(self, layer, q, k, v, kvs, qsl):
        c = self.cfg
        outs = []
        for b, kv in enumerate(kvs):
            s, e = qsl[b], qsl[b + 1]
            n = e - s
            p0 = kv.length
            kv.k[layer][p0:p0 + n] = k[s:e].view(n, c.num_kv_heads, c.head_dim)
            kv.v[layer][p0:p0 + n] = v[s:e].view(n, c.num_kv_heads, c.head_dim)
            outs.append(causal_attention(q[s:e].view(n, c.num_heads, c.head_dim),
                                         kv.k[layer][:p0 + n], kv.v[layer][:p0 + n],
                                         self.scale, p0))
        return torch.cat(outs, 0)

    def forward(self, ids: torch.Tensor, kvs: list[RequestKV], qsl: list[int],
                trace: dict | None = None) -> torch.Tensor:
        """ids [T] flattened over requests; qsl = query_start_loc [B+1]; each request's new
        tokens continue at position kvs[b].length.  Returns hidden [T,d] after the final norm
        and advances every kv.length."""
        c, w = self.cfg, self.w
        pos = torch.cat([torch.arange(kv.length, kv.length + qsl[b + 1] - qsl[b])
                         for b, kv in enumerate(kvs)])
        H, KV, hd = c.num_heads, c.num_kv_heads, c.head_dim
        if c.model_type != "opt":
            h = F.embedding(ids, w["model.embed_tokens.weight"])
            res = None
            cs = self.cos_sin.index_select(0, pos)
            cos, sin = cs.chunk(2, dim=-1)
            for i in range(c.num_layers):
                pre = f"model.layers.{i}."
                if res is None:
                    res = h
                    h = rms_norm(h, w[pre + "input_layernorm.weight"], c.eps)
                else:
                    h, res = fused_add_rms_norm(h, res, w[pre + "input_layernorm.weight"], c.eps)
                qkv = F.linear(h, self.qkv[i], self.qkv_b[i])
                q, k, v = qkv.split([H * hd, KV * hd, KV * hd], dim=-1)
                T = q.shape[0]
                q, k = q.reshape(T, H, hd), k.reshape(T, KV, hd)
                if c.qk_norm:   # qwen3.py:150-158: RMSNorm over head_dim, per head, before RoPE
                    q = rms_norm(q, w[pre + "self_attn.q_norm.weight"], c.eps)
                    k = rms_norm(k, w[pre + "self_attn.k_norm.weight"], c.eps)
                q = apply_rope_neox(q, cos, sin).reshape(T, H * hd)
                k = apply_rope_neox(k, cos, sin).reshape(T, KV * hd)
                a = self._attend(i, q, k, v, kvs, qsl)
                h = F.linear(a, w[pre + "self_attn.o_proj.weight"])
                h, res = fused_add_rms_norm(h, res, w[pre + "post_attention_layernorm.weight"], c.eps)
                gu = F.linear(h, self.gate_up[i])
                h = F.linear(silu_and_mul(gu), w[pre + "mlp.down_proj.weight"])
                if trace is not None:
                    trace[f"layer{i}.mlp_out"] = h.clone()
                    trace[f"layer{i}.residual"] = res.clone()
            h, _ = fused_add_rms_norm(h, res, w["model.norm.weight"], c.eps)
        else:
            p = "model.decoder."
            h = F.embedding(ids, w[p + "embed_tokens.weight"]) + \
                F.embedding(pos + 2, w[p + "embed_positions.weight"])
            for i in range(c.num_layers):
                pre = f"{p}layers.{i}."
                res = h
                x = layer_norm(h, w[pre + "self_attn_layer_norm.weight"],
                               w[pre + "self_attn_layer_norm.bias"])
                qkv = F.linear(x, self.qkv[i], self.qkv_b[i])
                q, k, v = qkv.chunk(3, dim=-1)
                a = self._attend(i, q, k, v, kvs, qsl)
                h = res + F.linear(a, w[pre + "self_attn.out_proj.weight"],
                                   w[pre + "self_attn.out_proj.bias"])
                res = h
                x = layer_norm(h, w[pre + "final_layer_norm.weight"], w[pre + "final_layer_norm.bias"])
                x = torch.relu(F.linear(x, w[pre + "fc1.weight"], w[pre + "fc1.bias"]))
                h = res + F.linear(x, w[pre + "fc2.weight"], w[pre + "fc2.bias"])
                if trace is not None:
                    trace[f"layer{i}.out"] = h.clone()
            h = layer_norm(h, w[p + "final_layer_norm.weight"], w[p + "final_layer_norm.bias"])
        for b, kv in enumerate(kvs):
            kv.length += qsl[b + 1] - qsl[b]
        return h

    def logits(self, hidden_rows: torch.Tensor) -> torch.Tensor:
        """logits_processor.py:54-99 - GEMM in the model dtype against the (tied) head;
        vocab padding columns are never produced here."""
        return F.linear(hidden_rows, self.lm_head)

    # -- convenience: whole-batch greedy generation (prefill per request, batched decode)
    @torch.no_grad()
    def generate_greedy(self, prompts: list[list[int]], max_tokens: int,
                        return_logits: bool = False):
        B = len(prompts)
        kvs = [self.new_kv(len(p) + max_tokens) for p in prompts]
        out = [[] for _ in range(B)]
        all_logits = []
        # prefill one request at a time (chunking does not change values)
        last = []
        for b, p in enumerate(prompts):
            h = self.forward(torch.tensor(p), [kvs[b]], [0, len(p)])
            last.append(h[-1:])
        lg = self.logits(torch.cat(last, 0)).to(torch.float32)
        for step in range(max_tokens):
            if return_logits:
                all_logits.append(lg.clone())
            nxt = lg.argmax(dim=-1)                 # first max wins (sampler.py:235-236)
            for b in range(B):
                out[b].append(int(nxt[b]))
            if step == max_tokens - 1:
                break
            h = self.forward(nxt, kvs, list(range(B + 1)))
            lg = self.logits(h).to(torch.float32)
        return (out, all_logits) if return_logits else out
