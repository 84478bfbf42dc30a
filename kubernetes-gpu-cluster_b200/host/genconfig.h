// Model-supplied sampling defaults and extra end-of-sequence ids, as vLLM's OpenAI server applies them with its
// default `--generation-config auto`: generation_config.json of the model directory gives the defaults for
// temperature / top_p / top_k when a request leaves them out (vllm:config/model.py get_diff_sampling_param), and
// every eos_token_id listed there (or in config.json, when that is a list) other than the primary one ends a request
// like a stop token unless ignore_eos (vllm:sampling_params.py:556-584 update_from_generation_config).
// `--generation-config vllm` switches the defaults off.
#pragma once
#include <stdint.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "json.h"

namespace kge {

struct GenDefaults {
  bool has_temperature = false, has_top_p = false, has_top_k = false;
  double temperature = 1.0, top_p = 1.0;
  int top_k = 0;
  std::vector<int32_t> extra_eos;          // end-of-sequence ids beyond config.json's primary eos_token_id
  std::vector<std::string> ignored;        // generation_config keys this engine does not implement (logged once)

  std::string to_json() const {
    std::string e = "[";
    for (size_t i = 0; i < extra_eos.size(); ++i) e += (i ? "," : "") + std::to_string(extra_eos[i]);
    e += "]";
    char b[256];
    snprintf(b, sizeof b, "{\"temperature\":%s,\"top_p\":%s,\"top_k\":%s,\"extra_eos_token_ids\":", has_temperature ? std::to_string(temperature).c_str() : "null",
             has_top_p ? std::to_string(top_p).c_str() : "null", has_top_k ? std::to_string(top_k).c_str() : "null");
    std::string ig = "[";
    for (size_t i = 0; i < ignored.size(); ++i) ig += (i ? ",\"" : "\"") + ignored[i] + "\"";
    return std::string(b) + e + ",\"ignored\":" + ig + "]}";
  }
};

inline bool read_json_file(const std::string& path, Json* j) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::stringstream ss;
  ss << f.rdbuf();
  std::string err;
  return json_parse(ss.str(), j, &err) && j->is_obj();
}

inline void collect_ids(const Json* e, std::vector<int32_t>* out) {
  if (!e) return;
  auto add = [&](const Json& x) {
    if (!x.is_num()) return;
    const int64_t v = x.as_i64();
    if (v >= 0 && v <= INT32_MAX && std::find(out->begin(), out->end(), (int32_t)v) == out->end()) out->push_back((int32_t)v);
  };
  if (e->is_num()) add(*e);
  else if (e->is_arr()) for (auto& x : e->arr) add(x);
}

// `use_generation_config` false = `--generation-config vllm`: no sampling defaults from the model (eos ids still apply)
inline GenDefaults load_gen_defaults(const std::string& dir, bool use_generation_config) {
  GenDefaults g;
  Json cfg, gen;
  int64_t primary = -1;
  if (read_json_file(dir + "/config.json", &cfg)) {
    const Json* e = cfg.get("eos_token_id");
    std::vector<int32_t> ids;
    collect_ids(e, &ids);
    if (!ids.empty()) primary = ids[0];
    for (size_t i = 1; i < ids.size(); ++i) g.extra_eos.push_back(ids[i]);
  }
  if (read_json_file(dir + "/generation_config.json", &gen)) {
    std::vector<int32_t> ids;
    collect_ids(gen.get("eos_token_id"), &ids);
    for (int32_t id : ids)
      if (id != primary && std::find(g.extra_eos.begin(), g.extra_eos.end(), id) == g.extra_eos.end()) g.extra_eos.push_back(id);
    if (use_generation_config) {
      if (const Json* t = gen.get("temperature")) if (t->is_num()) { g.has_temperature = true; g.temperature = t->num; }
      if (const Json* t = gen.get("top_p")) if (t->is_num()) { g.has_top_p = true; g.top_p = t->num; }
      if (const Json* t = gen.get("top_k")) if (t->is_num()) { g.has_top_k = true; g.top_k = (int)std::max<int64_t>(0, std::min<int64_t>(t->as_i64(), INT32_MAX)); }
      for (const char* k : {"repetition_penalty", "min_p"})
        if (const Json* t = gen.get(k)) if (t->is_num() && ((std::string(k) == "min_p" && t->num != 0.0) || (std::string(k) != "min_p" && t->num != 1.0))) g.ignored.push_back(k);
    }
  }
  return g;
}

}  // namespace kge
