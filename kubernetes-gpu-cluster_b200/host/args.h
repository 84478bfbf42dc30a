// `vllm serve MODEL [flags]` command line of the image the reference's values files start
// (ref:values-01-minimal-example4.yaml:16-23, ref:values-01-minimal-example8.yaml:23-38; flag list
// vllm:engine/arg_utils.py:415-598, vllm:entrypoints/openai/cli_args.py:228-230).  Both `--flag value`
// and `--flag=value` spellings; unknown flags are an error (a typo must not silently change behaviour).
#pragma once
#include <stdint.h>
#include <stdlib.h>

#include <string>
#include <vector>

namespace kge {

struct ServeArgs {
  std::string model;                 // HF id or directory
  std::string served_model_name;     // defaults to `model`
  std::string host = "0.0.0.0";
  int port = 8000;
  int max_model_len = 0;
  std::string dtype = "auto";
  int tensor_parallel_size = 1;
  int pipeline_parallel_size = 1;
  double gpu_memory_utilization = 0.9;
  int max_num_seqs = 0;
  int max_num_batched_tokens = 0;
  int block_size = 64;      // tokens per KV block: one TMA box per (block, kv head, 64-column half); 16 and 32 are accepted
  int num_gpu_blocks_override = 0;
  bool enable_chunked_prefill = true;
  bool scheduler_reserve_full_isl = true;
  std::string api_key;                          // --api-key or $VLLM_API_KEY: Bearer token required on /v1 routes
  std::string generation_config = "auto";     // auto: defaults from the model's generation_config.json; vllm: none
  bool enable_prefix_caching = true;       // vLLM V1 default
  bool enforce_eager = false;
  bool disable_custom_all_reduce = false;
  bool trust_remote_code = false;
  std::string load_format = "auto";
  int64_t seed = 0;
  bool dry_run = false;              // print the parsed configuration as JSON and exit (no GPU needed)
};

// Returns "" on success, else the error text.
inline std::string parse_serve_args(int argc, char** argv, ServeArgs* a) {
  std::vector<std::string> v(argv + 1, argv + argc);
  size_t i = 0;
  if (i < v.size() && v[i] == "serve") ++i;
  auto is_flag = [](const std::string& s) { return s.size() > 2 && s[0] == '-' && s[1] == '-'; };
  bool have_model = false;
  while (i < v.size()) {
    std::string arg = v[i++];
    if (!is_flag(arg)) {
      if (have_model) return "unexpected positional argument '" + arg + "'";
      a->model = arg;
      have_model = true;
      continue;
    }
    std::string key = arg, val;
    bool has_val = false;
    size_t eq = arg.find('=');
    if (eq != std::string::npos) { key = arg.substr(0, eq); val = arg.substr(eq + 1); has_val = true; }
    for (auto& c : key) if (c == '_') c = '-';
    auto need = [&](std::string* out) -> bool {
      if (has_val) { *out = val; return true; }
      if (i >= v.size()) return false;
      *out = v[i++];
      return true;
    };
    auto need_int = [&](int* out) -> std::string {
      std::string s;
      if (!need(&s)) return key + " expects a value";
      char* e = nullptr;
      long x = strtol(s.c_str(), &e, 10);
      if (!e || *e || s.empty()) return key + ": '" + s + "' is not an integer";
      *out = (int)x;
      return "";
    };
    std::string err, s;
    if (key == "--model") { if (!need(&a->model)) return key + " expects a value"; have_model = true; }
    else if (key == "--served-model-name") { if (!need(&a->served_model_name)) return key + " expects a value"; }
    else if (key == "--host") { if (!need(&a->host)) return key + " expects a value"; }
    else if (key == "--port") err = need_int(&a->port);
    else if (key == "--max-model-len") err = need_int(&a->max_model_len);
    else if (key == "--dtype") {
      if (!need(&a->dtype)) return key + " expects a value";
      if (a->dtype != "auto" && a->dtype != "float16" && a->dtype != "half" && a->dtype != "bfloat16" && a->dtype != "float32")
        return "--dtype must be one of auto, float16, bfloat16, float32";
    }
    else if (key == "--tensor-parallel-size" || key == "-tp") err = need_int(&a->tensor_parallel_size);
    else if (key == "--pipeline-parallel-size") err = need_int(&a->pipeline_parallel_size);
    else if (key == "--gpu-memory-utilization") {
      if (!need(&s)) return key + " expects a value";
      char* e = nullptr;
      a->gpu_memory_utilization = strtod(s.c_str(), &e);
      if (!e || *e || a->gpu_memory_utilization <= 0 || a->gpu_memory_utilization > 1) return "--gpu-memory-utilization must be in (0, 1]";
    }
    else if (key == "--max-num-seqs") err = need_int(&a->max_num_seqs);
    else if (key == "--max-num-batched-tokens") err = need_int(&a->max_num_batched_tokens);
    else if (key == "--block-size") err = need_int(&a->block_size);
    else if (key == "--num-gpu-blocks-override") err = need_int(&a->num_gpu_blocks_override);
    else if (key == "--enable-chunked-prefill") a->enable_chunked_prefill = true;
    else if (key == "--no-enable-chunked-prefill") a->enable_chunked_prefill = false;
    else if (key == "--api-key") { if (!need(&a->api_key)) return key + " expects a value"; }
    else if (key == "--generation-config") { if (!need(&a->generation_config)) return key + " expects a value"; }
    else if (key == "--scheduler-reserve-full-isl") a->scheduler_reserve_full_isl = true;
    else if (key == "--no-scheduler-reserve-full-isl") a->scheduler_reserve_full_isl = false;
    else if (key == "--enable-prefix-caching") a->enable_prefix_caching = true;
    else if (key == "--no-enable-prefix-caching") a->enable_prefix_caching = false;
    else if (key == "--enforce-eager") a->enforce_eager = true;
    else if (key == "--disable-custom-all-reduce") a->disable_custom_all_reduce = true;
    else if (key == "--trust-remote-code") a->trust_remote_code = true;
    else if (key == "--load-format") {
      if (!need(&a->load_format)) return key + " expects a value";
      if (a->load_format != "auto" && a->load_format != "safetensors" && a->load_format != "dummy")
        return "--load-format must be one of auto, safetensors, dummy";
    }
    else if (key == "--seed") { int x = 0; err = need_int(&x); a->seed = x; }
    else if (key == "--dry-run") a->dry_run = true;
    else return "unrecognized argument " + arg;
    if (!err.empty()) return err;
  }
  if (!have_model || a->model.empty()) return "the model (HF id or directory) is required: vllm serve MODEL [flags]";
  if (a->pipeline_parallel_size != 1)
    return "--pipeline-parallel-size " + std::to_string(a->pipeline_parallel_size) +
           " is not supported (this engine has tensor parallelism and replicas only; use 1)";
  if (a->tensor_parallel_size < 1) return "--tensor-parallel-size must be >= 1";
  if (a->port <= 0 || a->port > 65535) return "--port out of range";
  if (a->served_model_name.empty()) a->served_model_name = a->model;
  if (a->api_key.empty()) if (const char* k = getenv("VLLM_API_KEY")) a->api_key = k;
  return "";
}

}  // namespace kge
