// HF-layout model directory reader: config.json -> ModelSpec, *.safetensors -> host tensors
// (mmap, zero-copy), converted to bf16 on demand.  Host only.
#pragma once
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

namespace kge {

struct ModelSpec {
  std::string model_type;       // "opt" | "llama" | "qwen2" | "qwen3"
  bool qkv_bias = false;        // qwen2: q/k/v projections carry a bias
  bool qk_norm = false;         // qwen3: RMSNorm over head_dim on q and k, per head, before RoPE
  int vocab = 0, d = 0, L = 0, H = 0, KV = 0, hd = 0, ffn = 0, max_pos = 0;
  float eps = 1e-5f;
  bool tied = false;
  double rope_theta = 10000.0;
  double rope_factor = 0.0, rope_low = 1.0, rope_high = 4.0;   // factor 0 => no llama3 scaling
  int rope_orig_max = 8192;
  int bos = -1, eos = -1;
  std::string torch_dtype;      // config.json torch_dtype / dtype ("float16", "bfloat16", "float32", or empty)
  bool is_opt() const { return model_type == "opt"; }
};

struct HostTensor {
  const uint8_t* data = nullptr;   // inside an mmap
  std::string dtype;               // "BF16" | "F16" | "F32"
  std::vector<int64_t> shape;
  int64_t numel() const { int64_t n = 1; for (auto s : shape) n *= s; return n; }
};

class ModelDir {
 public:
  ~ModelDir();
  // Reads config.json (required) and maps every *.safetensors file (may be none).
  bool open(const std::string& dir, std::string* err);
  const ModelSpec& spec() const { return spec_; }
  bool has_weights() const { return !tensors_.empty(); }
  const HostTensor* find(const std::string& name) const;
  // Rows [r0, r1) x cols [c0, c1) of a 2-D (or all of a 1-D, r-range only) tensor as bf16 bits.
  bool read_bf16(const std::string& name, int64_t r0, int64_t r1, int64_t c0, int64_t c1,
                 std::vector<uint16_t>* out, std::string* err) const { return read_16(name, r0, r1, c0, c1, false, out, err); }
  // The same slice as bfloat16 bits (f16 = false) or IEEE half bits (f16 = true), converted from whatever the file holds.
  bool read_16(const std::string& name, int64_t r0, int64_t r1, int64_t c0, int64_t c1, bool f16,
               std::vector<uint16_t>* out, std::string* err) const;

 private:
  struct Map { void* base; size_t len; };
  ModelSpec spec_;
  std::map<std::string, HostTensor> tensors_;
  std::vector<Map> maps_;
};

}  // namespace kge
