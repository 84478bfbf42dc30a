// Engine construction: weight loading (TP-sliced), dummy weights, activation arena, KV pool.
#include "engine.h"

#include <stdlib.h>
#include <string.h>

#include <algorithm>

using namespace kge;

#include <cuda_fp16.h>

namespace kge {

const Ops kOpsBf16 = {kge_embed, kge_rmsnorm, kge_add_rmsnorm, kge_head_rmsnorm, kge_layernorm, kge_silu_mul, kge_gemm_bf16,
                      kge_gemm_silu_gate_ws_bf16, kge_rope_kvwrite, kge_rope_cache_host, kge_paged_attn, kge_argmax, kge_sample,
                      kge_logprobs, kge_tp_reduce_add_norm};
const Ops kOpsF16 = {kge_embed_f16, kge_rmsnorm_f16, kge_add_rmsnorm_f16, kge_head_rmsnorm_f16, kge_layernorm_f16, kge_silu_mul_f16,
                     kge_gemm_f16, kge_gemm_silu_gate_ws_f16, kge_rope_kvwrite_f16, kge_rope_cache_host_f16, kge_paged_attn_f16,
                     kge_argmax_f16, kge_sample_f16, kge_logprobs_f16, kge_tp_reduce_add_norm_f16};

// 16-bit pattern of v in the engine's element type
__device__ __forceinline__ uint16_t to_elt_bits(float v, bool f16) {
  return f16 ? __half_as_ushort(__float2half_rn(v)) : __bfloat16_as_ushort(__float2bfloat16_rn(v));
}

__global__ void fill_uniform_kernel(bf16* p, int64_t n, uint64_t seed, float scale, bool f16) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t z = (uint64_t)i * 0x9E3779B97F4A7C15ull + seed;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const float u = (float)(z >> 40) * (1.0f / 16777216.0f);   // [0,1)
    reinterpret_cast<uint16_t*>(p)[i] = to_elt_bits((2.f * u - 1.f) * scale, f16);
  }
}
__global__ void fill_const_kernel(bf16* p, int64_t n, float v, bool f16) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    reinterpret_cast<uint16_t*>(p)[i] = to_elt_bits(v, f16);
}

struct Builder {
  kge_engine* e;
  const ModelDir* md;
  int64_t dummy_seed;
  uint64_t counter = 1;
  std::string err;

  bf16* dalloc(int64_t elems) {
    void* p = nullptr;
    if (cudaMalloc(&p, (size_t)std::max<int64_t>(elems, 8) * 2) != cudaSuccess) {
      err = std::string("cudaMalloc failed: ") + cudaGetErrorString(cudaGetLastError());
      return nullptr;
    }
    e->allocs.push_back(p);
    e->weight_bytes += elems * 2;
    return (bf16*)p;
  }
  bool dummy() const { return !md->has_weights(); }
  // The device buffers are sized from config.json; a checkpoint tensor of another shape would be copied past the end of
  // its buffer (or leave part of it unset), so every tensor's shape is checked against the config before anything loads.
  bool expect(const std::string& name, int64_t rows, int64_t cols /* 0: 1-D */) {
    const HostTensor* t = md->find(name);
    if (!t) { err = "missing tensor " + name; return false; }
    const bool ok = cols == 0 ? (t->shape.size() == 1 && t->shape[0] == rows)
                              : (t->shape.size() == 2 && t->shape[0] == rows && t->shape[1] == cols);
    if (!ok) {
      std::string got;
      for (auto s : t->shape) got += (got.empty() ? "" : "x") + std::to_string(s);
      err = name + ": checkpoint shape " + got + " does not match config.json (expected " + std::to_string(rows) +
            (cols ? "x" + std::to_string(cols) : "") + ")";
    }
    return ok;
  }
  // dst rows [dst_r0, ...) of a [.., ncols] device matrix <- tensor slice
  bool put(bf16* dst, int64_t dst_off, const std::string& name, int64_t r0, int64_t r1, int64_t c0, int64_t c1) {
    std::vector<uint16_t> buf;
    if (!md->read_16(name, r0, r1, c0, c1, e->f16, &buf, &err)) return false;
    if (cudaMemcpy(dst + dst_off, buf.data(), buf.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess) {
      err = "cudaMemcpy failed for " + name;
      return false;
    }
    return true;
  }
  // `shard`: this rank's slice of a sharded tensor draws its own values (replicated tensors must agree on every rank,
  // which the common counter order guarantees)
  void rnd(bf16* p, int64_t n, float scale, bool shard = false) {
    const uint64_t salt = shard ? (uint64_t)e->tp_rank * 0xA24BAED4963EE407ull : 0;
    fill_uniform_kernel<<<(int)std::min<int64_t>((n + 255) / 256, 148 * 8), 256, 0, e->stream>>>(
        p, n, ((uint64_t)dummy_seed * 0x100000001B3ull + (counter++) * 0x9E3779B9ull) ^ salt, scale, e->f16);
  }
  void cst(bf16* p, int64_t n, float v) {
    fill_const_kernel<<<(int)std::min<int64_t>((n + 255) / 256, 148 * 8), 256, 0, e->stream>>>(p, n, v, e->f16);
  }
};

static const float kDummyScale = 0.02f * 1.7320508f;   // uniform with std 0.02

static bool build_llama(Builder& b) {
  kge_engine* e = b.e;
  const ModelSpec& s = e->spec;
  const int tp = e->tp_size, rk = e->tp_rank, hd = s.hd, d = s.d;
  const int Hl = e->H_loc, KVl = e->KV_loc, fl = e->ffn_loc;
  const int kv_rep = tp > s.KV ? tp / s.KV : 1;
  const int kv0 = (rk / kv_rep) * KVl;                 // first kv head of this rank
  if (!b.dummy()) {
    if (!b.expect("model.embed_tokens.weight", s.vocab, d) || !b.expect("model.norm.weight", d, 0)) return false;
    if (!s.tied && !b.expect("lm_head.weight", s.vocab, d)) return false;
    for (int i = 0; i < s.L; ++i) {
      const std::string p = "model.layers." + std::to_string(i) + ".";
      if (!b.expect(p + "self_attn.q_proj.weight", (int64_t)s.H * hd, d) || !b.expect(p + "self_attn.k_proj.weight", (int64_t)s.KV * hd, d) ||
          !b.expect(p + "self_attn.v_proj.weight", (int64_t)s.KV * hd, d) || !b.expect(p + "self_attn.o_proj.weight", d, (int64_t)s.H * hd) ||
          !b.expect(p + "mlp.gate_proj.weight", s.ffn, d) || !b.expect(p + "mlp.up_proj.weight", s.ffn, d) ||
          !b.expect(p + "mlp.down_proj.weight", d, s.ffn) || !b.expect(p + "input_layernorm.weight", d, 0) ||
          !b.expect(p + "post_attention_layernorm.weight", d, 0))
        return false;
      if (s.qkv_bias && (!b.expect(p + "self_attn.q_proj.bias", (int64_t)s.H * hd, 0) || !b.expect(p + "self_attn.k_proj.bias", (int64_t)s.KV * hd, 0) ||
                         !b.expect(p + "self_attn.v_proj.bias", (int64_t)s.KV * hd, 0)))
        return false;
      if (s.qk_norm && (!b.expect(p + "self_attn.q_norm.weight", hd, 0) || !b.expect(p + "self_attn.k_norm.weight", hd, 0))) return false;
    }
  }
  if (!(e->embed = b.dalloc((int64_t)s.vocab * d))) return false;
  if (b.dummy()) b.rnd(e->embed, (int64_t)s.vocab * d, kDummyScale);
  else if (!b.put(e->embed, 0, "model.embed_tokens.weight", 0, -1, 0, -1)) return false;
  // LM head: under tensor parallelism each rank holds the rows (vocabulary entries) [vocab_col0, vocab_col0 + vocab_loc)
  // (vllm:model_executor/layers/vocab_parallel_embedding.py:81-83; the embedding itself stays replicated here)
  if (s.tied) e->lm_head = e->embed + (int64_t)e->vocab_col0 * d;
  else {
    if (!(e->lm_head = b.dalloc((int64_t)e->vocab_loc * d))) return false;
    if (b.dummy()) b.rnd(e->lm_head, (int64_t)e->vocab_loc * d, kDummyScale, tp > 1);
    else if (!b.put(e->lm_head, 0, "lm_head.weight", e->vocab_col0, e->vocab_col0 + e->vocab_loc, 0, -1)) return false;
  }
  if (!(e->final_w = b.dalloc(d))) return false;
  if (b.dummy()) b.cst(e->final_w, d, 1.f);
  else if (!b.put(e->final_w, 0, "model.norm.weight", 0, -1, 0, -1)) return false;
  e->layers.resize(s.L);
  for (int i = 0; i < s.L; ++i) {
    LayerW& L = e->layers[i];
    const std::string p = "model.layers." + std::to_string(i) + ".";
    const int nq = Hl * hd, nkv = KVl * hd;
    L.qkv.N = nq + 2 * nkv; L.qkv.K = d;
    L.o.N = d; L.o.K = nq;
    L.up.N = 2 * fl; L.up.K = d;
    L.down.N = d; L.down.K = fl;
    if (!(L.qkv.w = b.dalloc((int64_t)L.qkv.N * d)) || !(L.o.w = b.dalloc((int64_t)d * nq)) ||
        !(L.up.w = b.dalloc((int64_t)2 * fl * d)) || !(L.down.w = b.dalloc((int64_t)d * fl)) ||
        !(L.ln1_w = b.dalloc(d)) || !(L.ln2_w = b.dalloc(d)))
      return false;
    if (s.qkv_bias && !(L.qkv.b = b.dalloc(L.qkv.N))) return false;
    if (s.qk_norm && (!(L.q_norm = b.dalloc(hd)) || !(L.k_norm = b.dalloc(hd)))) return false;
    if (b.dummy()) {
      if (s.qkv_bias) b.cst(L.qkv.b, L.qkv.N, 0.f);
      if (s.qk_norm) { b.cst(L.q_norm, hd, 1.f); b.cst(L.k_norm, hd, 1.f); }
      b.rnd(L.qkv.w, (int64_t)L.qkv.N * d, kDummyScale, tp > 1);
      b.rnd(L.o.w, (int64_t)d * nq, kDummyScale, tp > 1);
      b.rnd(L.up.w, (int64_t)2 * fl * d, kDummyScale, tp > 1);
      b.rnd(L.down.w, (int64_t)d * fl, kDummyScale, tp > 1);
      b.cst(L.ln1_w, d, 1.f);
      b.cst(L.ln2_w, d, 1.f);
      continue;
    }
    // fused rows [q_local | k_local | v_local] (vllm:model_executor/layers/linear.py:1072-1080,1363-1367)
    if (!b.put(L.qkv.w, 0, p + "self_attn.q_proj.weight", (int64_t)rk * nq, (int64_t)(rk + 1) * nq, 0, -1)) return false;
    if (!b.put(L.qkv.w, (int64_t)nq * d, p + "self_attn.k_proj.weight", (int64_t)kv0 * hd, (int64_t)(kv0 + KVl) * hd, 0, -1)) return false;
    if (!b.put(L.qkv.w, (int64_t)(nq + nkv) * d, p + "self_attn.v_proj.weight", (int64_t)kv0 * hd, (int64_t)(kv0 + KVl) * hd, 0, -1)) return false;
    if (s.qkv_bias) {
      if (!b.put(L.qkv.b, 0, p + "self_attn.q_proj.bias", (int64_t)rk * nq, (int64_t)(rk + 1) * nq, 0, -1)) return false;
      if (!b.put(L.qkv.b, nq, p + "self_attn.k_proj.bias", (int64_t)kv0 * hd, (int64_t)(kv0 + KVl) * hd, 0, -1)) return false;
      if (!b.put(L.qkv.b, nq + nkv, p + "self_attn.v_proj.bias", (int64_t)kv0 * hd, (int64_t)(kv0 + KVl) * hd, 0, -1)) return false;
    }
    if (s.qk_norm) {
      if (!b.put(L.q_norm, 0, p + "self_attn.q_norm.weight", 0, -1, 0, -1)) return false;
      if (!b.put(L.k_norm, 0, p + "self_attn.k_norm.weight", 0, -1, 0, -1)) return false;
    }
    if (!b.put(L.o.w, 0, p + "self_attn.o_proj.weight", 0, -1, (int64_t)rk * nq, (int64_t)(rk + 1) * nq)) return false;
    if (!b.put(L.up.w, 0, p + "mlp.gate_proj.weight", (int64_t)rk * fl, (int64_t)(rk + 1) * fl, 0, -1)) return false;
    if (!b.put(L.up.w, (int64_t)fl * d, p + "mlp.up_proj.weight", (int64_t)rk * fl, (int64_t)(rk + 1) * fl, 0, -1)) return false;
    if (!b.put(L.down.w, 0, p + "mlp.down_proj.weight", 0, -1, (int64_t)rk * fl, (int64_t)(rk + 1) * fl)) return false;
    if (!b.put(L.ln1_w, 0, p + "input_layernorm.weight", 0, -1, 0, -1)) return false;
    if (!b.put(L.ln2_w, 0, p + "post_attention_layernorm.weight", 0, -1, 0, -1)) return false;
  }
  // RoPE cache (host, vLLM's arithmetic) up to max_model_len positions
  const int max_pos = std::min(e->cfg.max_model_len, s.max_pos);
  std::vector<uint16_t> cs((size_t)max_pos * hd);
  if (e->op->rope_cache_host(cs.data(), max_pos, hd, s.rope_theta, s.rope_factor, s.rope_low, s.rope_high, s.rope_orig_max)) {
    b.err = kge_last_error();
    return false;
  }
  if (!(e->cos_sin = b.dalloc((int64_t)max_pos * hd))) return false;
  cudaMemcpy(e->cos_sin, cs.data(), cs.size() * 2, cudaMemcpyHostToDevice);
  return true;
}

static bool build_opt(Builder& b) {
  kge_engine* e = b.e;
  const ModelSpec& s = e->spec;
  const int tp = e->tp_size, rk = e->tp_rank, hd = s.hd, d = s.d;
  const int Hl = e->H_loc, fl = e->ffn_loc;
  (void)tp;
  const std::string dp = "model.decoder.";
  if (!b.dummy()) {
    if (!b.expect(dp + "embed_tokens.weight", s.vocab, d) || !b.expect(dp + "embed_positions.weight", s.max_pos + 2, d) ||
        !b.expect(dp + "final_layer_norm.weight", d, 0) || !b.expect(dp + "final_layer_norm.bias", d, 0))
      return false;
    for (int i = 0; i < s.L; ++i) {
      const std::string p = dp + "layers." + std::to_string(i) + ".";
      for (const char* nm : {"q_proj", "k_proj", "v_proj", "out_proj"})
        if (!b.expect(p + "self_attn." + nm + ".weight", d, d) || !b.expect(p + "self_attn." + nm + ".bias", d, 0)) return false;
      if (!b.expect(p + "fc1.weight", s.ffn, d) || !b.expect(p + "fc1.bias", s.ffn, 0) || !b.expect(p + "fc2.weight", d, s.ffn) ||
          !b.expect(p + "fc2.bias", d, 0) || !b.expect(p + "self_attn_layer_norm.weight", d, 0) ||
          !b.expect(p + "self_attn_layer_norm.bias", d, 0) || !b.expect(p + "final_layer_norm.weight", d, 0) ||
          !b.expect(p + "final_layer_norm.bias", d, 0))
        return false;
    }
  }
  if (!(e->embed = b.dalloc((int64_t)s.vocab * d)) || !(e->pos_embed = b.dalloc((int64_t)(s.max_pos + 2) * d)) ||
      !(e->final_w = b.dalloc(d)) || !(e->final_b = b.dalloc(d)))
    return false;
  e->lm_head = e->embed + (int64_t)e->vocab_col0 * d;
  if (b.dummy()) {
    b.rnd(e->embed, (int64_t)s.vocab * d, kDummyScale);
    b.rnd(e->pos_embed, (int64_t)(s.max_pos + 2) * d, kDummyScale);
    b.cst(e->final_w, d, 1.f);
    b.cst(e->final_b, d, 0.f);
  } else {
    if (!b.put(e->embed, 0, dp + "embed_tokens.weight", 0, -1, 0, -1)) return false;
    if (!b.put(e->pos_embed, 0, dp + "embed_positions.weight", 0, -1, 0, -1)) return false;
    if (!b.put(e->final_w, 0, dp + "final_layer_norm.weight", 0, -1, 0, -1)) return false;
    if (!b.put(e->final_b, 0, dp + "final_layer_norm.bias", 0, -1, 0, -1)) return false;
  }
  e->layers.resize(s.L);
  for (int i = 0; i < s.L; ++i) {
    LayerW& L = e->layers[i];
    const std::string p = dp + "layers." + std::to_string(i) + ".";
    const int nq = Hl * hd;
    L.qkv.N = 3 * nq; L.qkv.K = d;
    L.o.N = d; L.o.K = nq;
    L.up.N = fl; L.up.K = d;
    L.down.N = d; L.down.K = fl;
    if (!(L.qkv.w = b.dalloc((int64_t)3 * nq * d)) || !(L.qkv.b = b.dalloc(3 * nq)) ||
        !(L.o.w = b.dalloc((int64_t)d * nq)) || !(L.o.b = b.dalloc(d)) ||
        !(L.up.w = b.dalloc((int64_t)fl * d)) || !(L.up.b = b.dalloc(fl)) ||
        !(L.down.w = b.dalloc((int64_t)d * fl)) || !(L.down.b = b.dalloc(d)) ||
        !(L.ln1_w = b.dalloc(d)) || !(L.ln1_b = b.dalloc(d)) || !(L.ln2_w = b.dalloc(d)) || !(L.ln2_b = b.dalloc(d)))
      return false;
    if (b.dummy()) {
      b.rnd(L.qkv.w, (int64_t)3 * nq * d, kDummyScale); b.cst(L.qkv.b, 3 * nq, 0.f);
      b.rnd(L.o.w, (int64_t)d * nq, kDummyScale); b.cst(L.o.b, d, 0.f);
      b.rnd(L.up.w, (int64_t)fl * d, kDummyScale); b.cst(L.up.b, fl, 0.f);
      b.rnd(L.down.w, (int64_t)d * fl, kDummyScale); b.cst(L.down.b, d, 0.f);
      b.cst(L.ln1_w, d, 1.f); b.cst(L.ln1_b, d, 0.f); b.cst(L.ln2_w, d, 1.f); b.cst(L.ln2_b, d, 0.f);
      continue;
    }
    const char* nm[3] = {"q_proj", "k_proj", "v_proj"};
    for (int j = 0; j < 3; ++j) {
      if (!b.put(L.qkv.w, (int64_t)j * nq * d, p + "self_attn." + nm[j] + ".weight", (int64_t)rk * nq, (int64_t)(rk + 1) * nq, 0, -1)) return false;
      if (!b.put(L.qkv.b, (int64_t)j * nq, p + "self_attn." + nm[j] + ".bias", (int64_t)rk * nq, (int64_t)(rk + 1) * nq, 0, -1)) return false;
    }
    if (!b.put(L.o.w, 0, p + "self_attn.out_proj.weight", 0, -1, (int64_t)rk * nq, (int64_t)(rk + 1) * nq)) return false;
    if (!b.put(L.up.w, 0, p + "fc1.weight", (int64_t)rk * fl, (int64_t)(rk + 1) * fl, 0, -1)) return false;
    if (!b.put(L.up.b, 0, p + "fc1.bias", (int64_t)rk * fl, (int64_t)(rk + 1) * fl, 0, -1)) return false;
    if (!b.put(L.down.w, 0, p + "fc2.weight", 0, -1, (int64_t)rk * fl, (int64_t)(rk + 1) * fl)) return false;
    // row-parallel biases are added on rank 0 only (vllm:model_executor/layers/linear.py:1553-1559)
    if (rk == 0) {
      if (!b.put(L.o.b, 0, p + "self_attn.out_proj.bias", 0, -1, 0, -1)) return false;
      if (!b.put(L.down.b, 0, p + "fc2.bias", 0, -1, 0, -1)) return false;
    } else {
      b.cst(L.o.b, d, 0.f);
      b.cst(L.down.b, d, 0.f);
    }
    if (!b.put(L.ln1_w, 0, p + "self_attn_layer_norm.weight", 0, -1, 0, -1)) return false;
    if (!b.put(L.ln1_b, 0, p + "self_attn_layer_norm.bias", 0, -1, 0, -1)) return false;
    if (!b.put(L.ln2_w, 0, p + "final_layer_norm.weight", 0, -1, 0, -1)) return false;
    if (!b.put(L.ln2_b, 0, p + "final_layer_norm.bias", 0, -1, 0, -1)) return false;
  }
  return true;
}

static int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

StepMeta carve_meta(uint8_t* base, int T, int B, int R, int max_blocks, int64_t* total) {
  StepMeta m;
  int64_t off = 0;
  auto take = [&](int64_t bytes) { uint8_t* p = base ? base + off : nullptr; off = align_up(off + bytes, 16); return p; };
  m.slots = (int64_t*)take((int64_t)T * 8);
  m.seeds = (uint64_t*)take((int64_t)R * 8);
  m.ids = (int32_t*)take((int64_t)T * 4);
  m.pos = (int32_t*)take((int64_t)T * 4);
  m.qsl = (int32_t*)take((int64_t)(B + 1) * 4);
  m.seq_lens = (int32_t*)take((int64_t)B * 4);
  m.rows = (int32_t*)take((int64_t)R * 4);
  m.top_k = (int32_t*)take((int64_t)R * 4);
  m.temp = (float*)take((int64_t)R * 4);
  m.top_p = (float*)take((int64_t)R * 4);
  m.btab = (int32_t*)take((int64_t)B * max_blocks * 4);
  if (total) *total = off;
  return m;
}

}  // namespace kge

#define ENGINE_FAIL(...)            \
  do {                              \
    kge::set_error(__VA_ARGS__);    \
    kge_engine_destroy(e);          \
    return -1;                      \
  } while (0)

extern "C" int kge_engine_create(kge_engine** out, const char* model_dir, const kge_engine_config* cfg,
                                 int64_t dummy_seed) {
  KGE_REQUIRE(out && model_dir && cfg, "kge_engine_create: null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    (void)cudaGetLastError();
    kge::set_error("kge_engine_create: no CUDA device (this engine has no CPU path)");
    return -2;
  }
  KGE_REQUIRE(cfg->device >= 0 && cfg->device < ndev, "kge_engine_create: device %d of %d", cfg->device, ndev);
  KGE_CUDA_OK(cudaSetDevice(cfg->device));
  kge_engine* e = new kge_engine();
  e->cfg = *cfg;
  ModelDir md;
  std::string err;
  if (!md.open(model_dir, &err)) ENGINE_FAIL("kge_engine_create: %s", err.c_str());
  e->spec = md.spec();
  const ModelSpec& s = e->spec;
  if (!md.has_weights() && dummy_seed < 0) ENGINE_FAIL("kge_engine_create: no *.safetensors in %s (use --load-format dummy)", model_dir);
  // --dtype: auto follows the checkpoint (float16 stays float16; everything else runs in bfloat16), or as told
  if (cfg->dtype < 0 || cfg->dtype > 2) ENGINE_FAIL("kge_engine_create: dtype %d (0 auto, 1 bfloat16, 2 float16)", cfg->dtype);
  e->f16 = cfg->dtype == 2 || (cfg->dtype == 0 && (s.torch_dtype == "float16" || s.torch_dtype == "half"));
  e->op = e->f16 ? &kOpsF16 : &kOpsBf16;
  e->tp_size = cfg->tp_size > 0 ? cfg->tp_size : 1;
  e->tp_rank = cfg->tp_rank;
  if (e->tp_rank < 0 || e->tp_rank >= e->tp_size) ENGINE_FAIL("kge_engine_create: tp_rank %d of %d", e->tp_rank, e->tp_size);
  if (s.H % e->tp_size || s.ffn % e->tp_size) ENGINE_FAIL("kge_engine_create: heads %d / ffn %d not divisible by tensor-parallel-size %d", s.H, s.ffn, e->tp_size);
  if (e->tp_size > s.KV && e->tp_size % s.KV) ENGINE_FAIL("kge_engine_create: tensor-parallel-size %d incompatible with %d kv heads", e->tp_size, s.KV);
  e->H_loc = s.H / e->tp_size;
  e->KV_loc = std::max(1, s.KV / e->tp_size);
  e->ffn_loc = s.ffn / e->tp_size;
  // vocabulary shard of the LM head: equal widths, a multiple of 8 so gathered rows move as 16-byte vectors
  e->vocab_shard = e->tp_size == 1 ? s.vocab : ((s.vocab + e->tp_size - 1) / e->tp_size + 7) / 8 * 8;
  e->vocab_col0 = e->tp_rank * e->vocab_shard;
  e->vocab_loc = std::min(e->vocab_shard, s.vocab - e->vocab_col0);
  if (e->vocab_loc <= 0) ENGINE_FAIL("kge_engine_create: vocabulary of %d is too small to shard over %d ranks", s.vocab, e->tp_size);
  if (getenv("KGE_DISABLE_GRAPHS")) e->use_graphs = false;               // diagnostics only
  if (e->tp_size > 1) {
    // Until the ranks are connected over peer memory (kge_engine_tp_connect, which switches both back on) the exchange
    // is NCCL's: plain stream order and eager decode steps there (with programmatic dependent
    // launch around the NCCL kernels, Llama-70B at TP=4 deadlocked between ranks in round 1).
    // KGE_TP_PDL=1 keeps programmatic dependent launch between the engine's own kernels on that path as well.
    if (!getenv("KGE_TP_PDL")) kge::pdl_set_enabled(false);
    e->use_graphs = false;
  }
  // default KV block: 64 tokens, so that a (block, kv head) run of the cache is one 8 KB TMA box per 64-column half
  // (16 and 32, vLLM's other sizes, are accepted; profiles/r02_attn_block_size.jsonl has what they cost)
  if (e->cfg.block_size <= 0) e->cfg.block_size = 64;
  if (e->cfg.block_size & (e->cfg.block_size - 1)) ENGINE_FAIL("kge_engine_create: block_size must be a power of two");
  if (e->cfg.max_model_len <= 0) e->cfg.max_model_len = std::min(s.max_pos, 8192);
  if (e->cfg.max_model_len > s.max_pos) ENGINE_FAIL("kge_engine_create: max_model_len %d > model max positions %d", e->cfg.max_model_len, s.max_pos);
  if (e->cfg.max_num_seqs <= 0) e->cfg.max_num_seqs = 1024;                     // vllm:engine/arg_utils.py:2311-2331
  if (e->cfg.max_num_batched_tokens <= 0) e->cfg.max_num_batched_tokens = 8192;
  if (e->cfg.gpu_memory_utilization <= 0.f) e->cfg.gpu_memory_utilization = 0.9f;
  if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) ENGINE_FAIL("cudaStreamCreate failed");

  Builder b{e, &md, dummy_seed};
  const bool ok = s.is_opt() ? build_opt(b) : build_llama(b);
  if (!ok) ENGINE_FAIL("kge_engine_create: %s", b.err.c_str());

  // activation arena
  const int Tm = e->cfg.max_num_batched_tokens, Rm = e->cfg.max_num_seqs, d = s.d;
  const int qkv_n = e->layers[0].qkv.N, up_n = e->layers[0].up.N;
  e->ld_logits = (int)align_up(s.vocab, 64);          // vocab padded to 64 (vocab_parallel_embedding.py:32)
  int64_t act_bytes = 0;
  auto aalloc = [&](int64_t elems, int esz) -> void* {
    void* p = nullptr;
    if (cudaMalloc(&p, (size_t)elems * esz) != cudaSuccess) return nullptr;
    e->allocs.push_back(p);
    act_bytes += elems * esz;
    return p;
  };
  if (e->tp_size > 1) {
    // the exchange window: [h | xn | gathered logits], same offsets on every rank
    e->off_h = 0;
    e->off_xn = align_up((int64_t)Tm * d * 2, 256);
    e->off_logits = e->off_xn + align_up((int64_t)Tm * d * 2, 256);
    const int64_t win = e->off_logits + align_up((int64_t)Rm * e->ld_logits * 2, 256);
    if (kge_tp_comm_create(&e->tpc, e->tp_rank, e->tp_size, e->cfg.device, win)) { kge_engine_destroy(e); return -2; }
    uint8_t* base = (uint8_t*)kge_tp_comm_data(e->tpc);
    e->h = (bf16*)(base + e->off_h);
    e->xn = (bf16*)(base + e->off_xn);
    e->logits = (bf16*)(base + e->off_logits);
    act_bytes += win;
    e->logits_loc = (bf16*)aalloc((int64_t)Rm * e->vocab_shard, 2);
    if (!e->logits_loc) ENGINE_FAIL("kge_engine_create: activation arena allocation failed");
  } else {
    e->h = (bf16*)aalloc((int64_t)Tm * d, 2);
    e->xn = (bf16*)aalloc((int64_t)Tm * d, 2);
    e->logits = (bf16*)aalloc((int64_t)Rm * e->ld_logits, 2);
  }
  e->res = (bf16*)aalloc((int64_t)Tm * d, 2);
  e->qkv = (bf16*)aalloc((int64_t)Tm * qkv_n, 2);
  e->attn = (bf16*)aalloc((int64_t)Tm * e->H_loc * s.hd, 2);
  e->gu = (bf16*)aalloc((int64_t)Tm * up_n, 2);
  e->act = (bf16*)aalloc((int64_t)Tm * e->ffn_loc, 2);
  e->xsel = (bf16*)aalloc((int64_t)Rm * d, 2);
  e->logits_f32 = nullptr;
  e->d_out = (int32_t*)aalloc(Rm, 4);
  e->lp_vals = (float*)aalloc((int64_t)Rm * 21, 4);
  e->lp_ids = (int32_t*)aalloc((int64_t)Rm * 21, 4);
  e->lp_rank = (int32_t*)aalloc(Rm, 4);
  e->mask_pairs = (int32_t*)aalloc(2 * kge_engine::kMaxMaskPairs, 4);
  e->ws_gemm_bytes = 1024 + (int64_t)148 * 256 * 128 * 4;
  e->ws_gemm = aalloc(e->ws_gemm_bytes, 1);
  e->ws_attn_bytes = kge_attn_workspace_bytes(Rm, e->H_loc, s.hd, e->cfg.max_model_len);
  e->ws_attn = aalloc(e->ws_attn_bytes, 1);
  e->max_blocks_per_seq = (e->cfg.max_model_len + e->cfg.block_size - 1) / e->cfg.block_size;
  carve_meta(nullptr, Tm, Rm, Rm, e->max_blocks_per_seq, &e->meta_bytes);
  e->meta_dev = (uint8_t*)aalloc(e->meta_bytes, 1);
  if (!e->h || !e->res || !e->xn || !e->qkv || !e->attn || !e->gu || !e->act || !e->xsel || !e->logits ||
      !e->d_out || !e->lp_vals || !e->lp_ids || !e->lp_rank || !e->mask_pairs || !e->ws_gemm || !e->ws_attn || !e->meta_dev)
    ENGINE_FAIL("kge_engine_create: activation arena allocation failed (%s)", cudaGetErrorString(cudaGetLastError()));
  if (cudaMallocHost((void**)&e->meta_host, e->meta_bytes) != cudaSuccess ||
      cudaMallocHost((void**)&e->out_host, (size_t)Rm * 4) != cudaSuccess ||
      cudaMallocHost((void**)&e->lp_vals_host, (size_t)Rm * 21 * 4) != cudaSuccess ||
      cudaMallocHost((void**)&e->lp_ids_host, (size_t)Rm * 21 * 4) != cudaSuccess ||
      cudaMallocHost((void**)&e->lp_rank_host, (size_t)Rm * 4) != cudaSuccess ||
      cudaMallocHost((void**)&e->mask_pairs_host, (size_t)2 * kge_engine::kMaxMaskPairs * 4) != cudaSuccess)
    ENGINE_FAIL("kge_engine_create: pinned staging allocation failed");
  cudaMemsetAsync(e->ws_gemm, 0, 1024, e->stream);     // split-K tile counters start at zero

  // paged KV pool: [num_blocks, KV_loc, block_size, hd] per layer for K and for V
  const int64_t blk_bytes = (int64_t)2 * s.L * e->KV_loc * e->cfg.block_size * s.hd * 2;
  int nb = e->cfg.num_blocks;
  if (nb <= 0) {
    size_t free_b = 0, total_b = 0;
    cudaMemGetInfo(&free_b, &total_b);
    const int64_t budget = (int64_t)((double)total_b * e->cfg.gpu_memory_utilization) - (int64_t)(total_b - free_b);
    nb = (int)std::max<int64_t>(0, budget / blk_bytes);
    const int64_t cap = (int64_t)e->cfg.max_num_seqs * e->max_blocks_per_seq + 1;
    if (nb > cap) nb = (int)cap;
  }
  if (nb < 2) ENGINE_FAIL("kge_engine_create: no memory left for the KV cache (gpu_memory_utilization=%.2f)", e->cfg.gpu_memory_utilization);
  // One request of max_model_len tokens must fit the pool (block 0 is the null block), or a request that outgrows it
  // would preempt itself for ever; vLLM refuses to start in the same situation (vllm:v1/core/kv_cache_utils.py
  // check_enough_kv_cache_memory).
  if ((int64_t)(nb - 1) * e->cfg.block_size < e->cfg.max_model_len)
    ENGINE_FAIL("kge_engine_create: the model's max seq len (%d) is larger than the maximum number of tokens that can be stored in "
                "KV cache (%lld). Try increasing gpu_memory_utilization or decreasing max_model_len when initializing the engine.",
                e->cfg.max_model_len, (long long)(nb - 1) * e->cfg.block_size);
  e->num_blocks = nb;
  e->kv_bytes = blk_bytes * nb;
  e->kc.resize(s.L); e->vc.resize(s.L);
  const int64_t per = (int64_t)nb * e->KV_loc * e->cfg.block_size * s.hd;
  for (int i = 0; i < s.L; ++i) {
    void *k = nullptr, *v = nullptr;
    if (cudaMalloc(&k, per * 2) != cudaSuccess || cudaMalloc(&v, per * 2) != cudaSuccess)
      ENGINE_FAIL("kge_engine_create: KV pool allocation failed at layer %d (%d blocks)", i, nb);
    e->allocs.push_back(k); e->allocs.push_back(v);
    e->kc[i] = (bf16*)k; e->vc[i] = (bf16*)v;
    cudaMemsetAsync(k, 0, per * 2, e->stream);
    cudaMemsetAsync(v, 0, per * 2, e->stream);
  }
  SchedConfig sc;
  sc.max_num_seqs = e->cfg.max_num_seqs;
  sc.max_num_batched_tokens = e->cfg.max_num_batched_tokens;
  sc.max_model_len = e->cfg.max_model_len;
  sc.block_size = e->cfg.block_size;
  sc.num_blocks = nb;
  sc.enable_chunked_prefill = e->cfg.enable_chunked_prefill != 0;
  sc.eos_token_id = s.eos;
  e->sched.reset(new Scheduler(sc));
  if (cudaStreamSynchronize(e->stream) != cudaSuccess)
    ENGINE_FAIL("kge_engine_create: device error during initialisation: %s", cudaGetErrorString(cudaGetLastError()));
  *out = e;
  return 0;
}

extern "C" void kge_engine_destroy(kge_engine* e) {
  if (!e) return;
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (auto& g : e->graphs) cudaGraphExecDestroy(g.exec);
  for (void* p : e->allocs) cudaFree(p);
  if (e->tpc) kge_tp_comm_destroy(e->tpc);
  if (e->meta_host) cudaFreeHost(e->meta_host);
  if (e->out_host) cudaFreeHost(e->out_host);
  if (e->lp_vals_host) cudaFreeHost(e->lp_vals_host);
  if (e->lp_ids_host) cudaFreeHost(e->lp_ids_host);
  if (e->lp_rank_host) cudaFreeHost(e->lp_rank_host);
  if (e->mask_pairs_host) cudaFreeHost(e->mask_pairs_host);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

extern "C" int kge_engine_num_blocks(const kge_engine* e) { return e ? e->num_blocks : 0; }
extern "C" int kge_engine_dtype(const kge_engine* e) { return e ? (e->f16 ? 2 : 1) : 0; }
extern "C" int kge_engine_vocab(const kge_engine* e) { return e ? e->spec.vocab : 0; }
extern "C" int kge_engine_eos_token_id(const kge_engine* e) { return e ? e->spec.eos : -1; }
extern "C" int kge_engine_max_model_len(const kge_engine* e) { return e ? e->cfg.max_model_len : 0; }
extern "C" int64_t kge_engine_weight_bytes(const kge_engine* e) { return e ? e->weight_bytes : 0; }
extern "C" int64_t kge_engine_launches(const kge_engine* e) { return e ? e->launches : 0; }
