#include "loader.h"

#include "half_bits.h"

#include <dirent.h>
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <fstream>
#include <sstream>

#include "json.h"

namespace kge {

ModelDir::~ModelDir() {
  for (auto& m : maps_) munmap(m.base, m.len);
}

static bool read_file(const std::string& path, std::string* out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::stringstream ss;
  ss << f.rdbuf();
  *out = ss.str();
  return true;
}

static bool parse_config(const Json& j, ModelSpec* s, std::string* err) {
  s->model_type = j.str_or("model_type", "");
  s->vocab = (int)j.int_or("vocab_size", 0);
  s->d = (int)j.int_or("hidden_size", 0);
  s->L = (int)j.int_or("num_hidden_layers", 0);
  s->H = (int)j.int_or("num_attention_heads", 0);
  s->max_pos = (int)j.int_or("max_position_embeddings", 2048);
  s->bos = (int)j.int_or("bos_token_id", -1);
  s->torch_dtype = j.str_or("torch_dtype", j.str_or("dtype", ""));
  if (const Json* e = j.get("eos_token_id")) {
    if (e->is_num()) s->eos = (int)e->num;
    else if (e->is_arr() && !e->arr.empty() && e->arr[0].is_num()) s->eos = (int)e->arr[0].num;
  }
  if (s->model_type == "opt") {
    s->KV = s->H;
    s->ffn = (int)j.int_or("ffn_dim", 0);
    s->tied = true;
    s->eps = 1e-5f;
    if (!j.bool_or("do_layer_norm_before", true)) { *err = "OPT with do_layer_norm_before=false (opt-350m) is not supported"; return false; }
    if (j.int_or("word_embed_proj_dim", s->d) != s->d) { *err = "OPT with word_embed_proj_dim != hidden_size is not supported"; return false; }
    if (j.str_or("activation_function", "relu") != "relu") { *err = "OPT activation_function must be relu"; return false; }
  } else if (s->model_type == "llama" || s->model_type == "qwen2" || s->model_type == "qwen3") {
    // Qwen2.5 / Qwen3 (the models the reference's values files deploy, ref:values-01-minimal-example2.yaml:8,
    // ...example8.yaml:8) share Llama's graph: vllm:model_executor/models/qwen2.py:160-175, qwen3.py:142-160
    const bool qwen = s->model_type != "llama";
    s->qkv_bias = s->model_type == "qwen2";
    s->qk_norm = s->model_type == "qwen3";
    s->KV = (int)j.int_or("num_key_value_heads", s->H);
    s->ffn = (int)j.int_or("intermediate_size", 0);
    s->tied = j.bool_or("tie_word_embeddings", false);
    s->eps = (float)j.num_or("rms_norm_eps", qwen ? 1e-6 : 1e-5);
    s->rope_theta = j.num_or("rope_theta", qwen ? 1000000.0 : 10000.0);
    if (const Json* rp = j.get("rope_parameters"))
      if (rp->is_obj()) s->rope_theta = rp->num_or("rope_theta", s->rope_theta);
    if (const Json* rs = j.get("rope_scaling")) {
      if (rs->is_obj()) {
        std::string t = rs->str_or("rope_type", rs->str_or("type", ""));
        if (t == "llama3") {
          s->rope_factor = rs->num_or("factor", 8.0);
          s->rope_low = rs->num_or("low_freq_factor", 1.0);
          s->rope_high = rs->num_or("high_freq_factor", 4.0);
          s->rope_orig_max = (int)rs->int_or("original_max_position_embeddings", 8192);
        } else if (!t.empty() && t != "default") {
          *err = "rope_scaling type '" + t + "' is not supported (llama3 only)";
          return false;
        }
      }
    }
    if (!qwen && (j.bool_or("attention_bias", false) || j.bool_or("mlp_bias", false))) { *err = "llama with attention_bias/mlp_bias is not supported"; return false; }
    if (s->model_type == "qwen3" && j.bool_or("attention_bias", false)) { *err = "qwen3 with attention_bias is not supported"; return false; }
    if (j.bool_or("use_sliding_window", false)) { *err = "sliding-window attention is not supported"; return false; }
  } else {
    *err = "unsupported model_type '" + s->model_type + "' (opt, llama, qwen2, qwen3)";
    return false;
  }
  if (s->H <= 0 || s->d <= 0 || s->L <= 0 || s->vocab <= 0 || s->ffn <= 0) { *err = "config.json misses a required size"; return false; }
  s->hd = (int)j.int_or("head_dim", s->d / s->H);
  if (s->hd != 64 && s->hd != 128) { *err = "head_dim must be 64 or 128"; return false; }
  if (s->H % s->KV != 0) { *err = "num_attention_heads must be a multiple of num_key_value_heads"; return false; }
  return true;
}

bool ModelDir::open(const std::string& dir, std::string* err) {
  std::string txt;
  if (!read_file(dir + "/config.json", &txt)) { *err = "cannot read " + dir + "/config.json"; return false; }
  Json cfg;
  std::string perr;
  if (!json_parse(txt, &cfg, &perr)) { *err = "config.json: " + perr; return false; }
  if (!parse_config(cfg, &spec_, err)) return false;

  std::vector<std::string> files;
  if (DIR* d = opendir(dir.c_str())) {
    while (dirent* e = readdir(d)) {
      std::string n = e->d_name;
      if (n.size() > 12 && n.substr(n.size() - 12) == ".safetensors") files.push_back(dir + "/" + n);
    }
    closedir(d);
  }
  std::sort(files.begin(), files.end());
  for (auto& path : files) {
    int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) { *err = "cannot open " + path; return false; }
    struct stat st;
    fstat(fd, &st);
    if (st.st_size < 8) { close(fd); *err = path + ": too small"; return false; }
    void* base = mmap(nullptr, st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (base == MAP_FAILED) { *err = "mmap failed for " + path; return false; }
    maps_.push_back({base, (size_t)st.st_size});
    const uint8_t* p = (const uint8_t*)base;
    uint64_t hlen;
    memcpy(&hlen, p, 8);
    if (hlen > (uint64_t)st.st_size - 8) { /* compared this way round: 8 + hlen could wrap */ *err = path + ": bad header length"; return false; }
    Json hdr;
    JsonParser jp((const char*)p + 8, hlen);
    if (!jp.parse(&hdr, &perr) || !hdr.is_obj()) { *err = path + ": header " + perr; return false; }
    const uint8_t* data0 = p + 8 + hlen;
    const uint64_t data_len = st.st_size - 8 - hlen;
    for (auto& kv : hdr.obj) {
      if (kv.first == "__metadata__") continue;
      const Json& t = kv.second;
      const Json* off = t.get("data_offsets");
      const Json* shp = t.get("shape");
      if (!off || !shp || off->arr.size() != 2) { *err = path + ": bad entry " + kv.first; return false; }
      HostTensor ht;
      ht.dtype = t.str_or("dtype", "");
      for (auto& s : shp->arr) ht.shape.push_back((int64_t)s.num);
      const uint64_t b = (uint64_t)off->arr[0].num, e = (uint64_t)off->arr[1].num;
      const int esz = ht.dtype == "F32" ? 4 : 2;
      if (e > data_len || b > e || (e - b) != (uint64_t)ht.numel() * esz) { *err = path + ": bad offsets for " + kv.first; return false; }
      if (ht.dtype != "BF16" && ht.dtype != "F16" && ht.dtype != "F32") { *err = kv.first + ": dtype " + ht.dtype + " unsupported"; return false; }
      ht.data = data0 + b;
      tensors_[kv.first] = ht;
    }
  }
  return true;
}

const HostTensor* ModelDir::find(const std::string& name) const {
  auto it = tensors_.find(name);
  return it == tensors_.end() ? nullptr : &it->second;
}

static inline uint16_t f32_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static inline float f16_f32(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000) << 16;
  uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ff, u;
  if (exp == 0) {
    if (man == 0) u = sign;
    else {
      exp = 127 - 15 + 1;
      while (!(man & 0x400)) { man <<= 1; --exp; }
      u = sign | (exp << 23) | ((man & 0x3ff) << 13);
    }
  } else if (exp == 31) u = sign | 0x7f800000u | (man << 13);
  else u = sign | ((exp + 127 - 15) << 23) | (man << 13);
  float f;
  memcpy(&f, &u, 4);
  return f;
}

static inline float bf16_f32(uint16_t b) {
  const uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

bool ModelDir::read_16(const std::string& name, int64_t r0, int64_t r1, int64_t c0, int64_t c1, bool f16,
                       std::vector<uint16_t>* out, std::string* err) const {
  const HostTensor* t = find(name);
  if (!t) { *err = "missing tensor " + name; return false; }
  int64_t rows, cols;
  if (t->shape.size() == 1) { rows = t->shape[0]; cols = 1; c0 = 0; c1 = 1; }
  else if (t->shape.size() == 2) { rows = t->shape[0]; cols = t->shape[1]; }
  else { *err = name + ": rank must be 1 or 2"; return false; }
  if (r1 < 0) r1 = rows;
  if (c1 < 0) c1 = cols;
  if (r0 < 0 || r1 > rows || c0 < 0 || c1 > cols || r0 > r1 || c0 > c1) {
    *err = name + ": slice out of range (shape " + std::to_string(rows) + "x" + std::to_string(cols) + ")";
    return false;
  }
  const int64_t nr = r1 - r0, nc = c1 - c0;
  out->resize((size_t)(nr * nc));
  for (int64_t r = 0; r < nr; ++r) {
    const int64_t src = (r0 + r) * cols + c0;
    uint16_t* dst = out->data() + r * nc;
    if (t->dtype == (f16 ? "F16" : "BF16")) {
      memcpy(dst, t->data + src * 2, (size_t)nc * 2);
    } else if (t->dtype == "F16") {
      const uint16_t* s = (const uint16_t*)t->data + src;
      for (int64_t c = 0; c < nc; ++c) dst[c] = f32_bf16(f16_f32(s[c]));
    } else if (t->dtype == "BF16") {
      const uint16_t* s = (const uint16_t*)t->data + src;
      for (int64_t c = 0; c < nc; ++c) dst[c] = f32_to_f16_bits(bf16_f32(s[c]));
    } else {
      const float* s = (const float*)t->data + src;
      for (int64_t c = 0; c < nc; ++c) dst[c] = f16 ? f32_to_f16_bits(s[c]) : f32_bf16(s[c]);
    }
  }
  return true;
}

}  // namespace kge

// ---------------------------------------------------------------- host-only C view (tests, tooling)
#include "common.cuh"
#include "kge.h"

struct kge_model_dir { kge::ModelDir md; };

extern "C" kge_model_dir* kge_model_open(const char* dir) {
  if (!dir) { kge::set_error("kge_model_open: null directory"); return nullptr; }
  auto* m = new kge_model_dir();
  std::string err;
  if (!m->md.open(dir, &err)) { kge::set_error("kge_model_open: %s", err.c_str()); delete m; return nullptr; }
  return m;
}
extern "C" void kge_model_close(kge_model_dir* m) { delete m; }

extern "C" int kge_model_spec_json(const kge_model_dir* m, char* out, int cap) {
  KGE_REQUIRE(m && out && cap > 0, "kge_model_spec_json: bad arguments");
  const kge::ModelSpec& s = m->md.spec();
  const int n = snprintf(out, (size_t)cap,
                         "{\"model_type\":\"%s\",\"qkv_bias\":%s,\"qk_norm\":%s,\"vocab\":%d,\"d\":%d,\"L\":%d,\"H\":%d,\"KV\":%d,\"hd\":%d,"
                         "\"ffn\":%d,\"max_pos\":%d,\"eps\":%.9g,\"tied\":%s,\"rope_theta\":%.17g,\"rope_factor\":%.17g,\"rope_low\":%.17g,"
                         "\"rope_high\":%.17g,\"rope_orig_max\":%d,\"bos\":%d,\"eos\":%d,\"has_weights\":%s}",
                         s.model_type.c_str(), s.qkv_bias ? "true" : "false", s.qk_norm ? "true" : "false", s.vocab, s.d, s.L, s.H, s.KV,
                         s.hd, s.ffn, s.max_pos, (double)s.eps, s.tied ? "true" : "false", s.rope_theta, s.rope_factor, s.rope_low,
                         s.rope_high, s.rope_orig_max, s.bos, s.eos, m->md.has_weights() ? "true" : "false");
  KGE_REQUIRE(n > 0 && n < cap, "kge_model_spec_json: buffer of %d bytes too small", cap);
  return n;
}

static int64_t model_read_16(const kge_model_dir* m, const char* name, int64_t r0, int64_t r1, int64_t c0, int64_t c1, bool f16,
                             uint16_t* out, int64_t cap);
extern "C" int64_t kge_model_read_bf16(const kge_model_dir* m, const char* name, int64_t r0, int64_t r1, int64_t c0, int64_t c1,
                                       uint16_t* out, int64_t cap) {
  return model_read_16(m, name, r0, r1, c0, c1, false, out, cap);
}
extern "C" int64_t kge_model_read_f16(const kge_model_dir* m, const char* name, int64_t r0, int64_t r1, int64_t c0, int64_t c1,
                                      uint16_t* out, int64_t cap) {
  return model_read_16(m, name, r0, r1, c0, c1, true, out, cap);
}
static int64_t model_read_16(const kge_model_dir* m, const char* name, int64_t r0, int64_t r1, int64_t c0, int64_t c1, bool f16,
                             uint16_t* out, int64_t cap) {
  if (!m || !name || !out) { kge::set_error("kge_model_read_bf16: null argument"); return -1; }
  std::vector<uint16_t> v;
  std::string err;
  if (!m->md.read_16(name, r0, r1, c0, c1, f16, &v, &err)) { kge::set_error("kge_model_read_bf16: %s", err.c_str()); return -1; }
  if ((int64_t)v.size() > cap) { kge::set_error("kge_model_read_bf16: %lld elements > capacity %lld", (long long)v.size(), (long long)cap); return -1; }
  memcpy(out, v.data(), v.size() * 2);
  return (int64_t)v.size();
}
