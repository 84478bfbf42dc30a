// TEST INFRASTRUCTURE ONLY - never shipped, never linked into the product.
// A stand-in libkge.so for exercising the HTTP server (host/server.cc) on a machine without a GPU: it exports the part of
// the C ABI (include/kge.h) the server calls and drives the REAL scheduler (csrc/scheduler.cc), but "samples" tokens with
// a fixed formula instead of running a model.  tests/test_server_http_cpu.py builds it into a temporary directory and
// puts that directory first on LD_LIBRARY_PATH of the unmodified kge-serve binary.  The product's own libkge.so has no
// such path: without CUDA every compute entry point fails (tests/test_abi_cpu.py::test_no_cpu_fallback).
//
// Token rule: next = (7 * last + 3 * position + 11) mod vocab, so outputs depend on the prompt and are reproducible.
// Every request's sampling parameters are appended to $MOCK_KGE_LOG (one JSON line) so tests can see what the server
// passed down.  Log-probabilities are fabricated: sampled -0.5, then -1.5, -2.5, ... for tokens next+1, next+2, ...
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "json.h"
#include "kge.h"
#include "scheduler.h"

using namespace kge;

static thread_local char g_err[512] = "";
static void set_err(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

struct kge_engine {
  std::unique_ptr<Scheduler> sched;
  int vocab = 1024, eos = -1, num_blocks = 0, max_model_len = 0;
  std::map<int64_t, int> want_lp;                       // request id -> top-N
  std::map<int64_t, std::pair<int32_t, int>> last_lp;    // request id -> (token sampled in the last step, top-N)
};

extern "C" const char* kge_last_error(void) { return g_err; }
extern "C" int kge_abi_version(void) { return KGE_ABI_VERSION; }
extern "C" int kge_device_sm_count(void) { return 148; }

extern "C" int kge_engine_create(kge_engine** out, const char* model_dir, const kge_engine_config* cfg, int64_t) {
  if (!out || !model_dir || !cfg) { set_err("kge_engine_create: null argument"); return -1; }
  std::ifstream f(std::string(model_dir) + "/config.json");
  if (!f) { set_err("kge_engine_create: cannot read %s/config.json", model_dir); return -1; }
  std::stringstream ss;
  ss << f.rdbuf();
  Json j;
  std::string err;
  if (!json_parse(ss.str(), &j, &err) || !j.is_obj()) { set_err("kge_engine_create: config.json: %s", err.c_str()); return -1; }
  auto* e = new kge_engine();
  e->vocab = (int)j.int_or("vocab_size", 1024);
  if (const Json* eo = j.get("eos_token_id")) {
    if (eo->is_num()) e->eos = (int)eo->as_i64();
    else if (eo->is_arr() && !eo->arr.empty()) e->eos = (int)eo->arr[0].as_i64();
  }
  SchedConfig sc;
  sc.max_num_seqs = cfg->max_num_seqs > 0 ? cfg->max_num_seqs : 1024;
  sc.max_num_batched_tokens = cfg->max_num_batched_tokens > 0 ? cfg->max_num_batched_tokens : 8192;
  sc.max_model_len = e->max_model_len = cfg->max_model_len > 0 ? cfg->max_model_len : (int)std::min<int64_t>(j.int_or("max_position_embeddings", 2048), 8192);
  sc.block_size = cfg->block_size > 0 ? cfg->block_size : 16;
  sc.num_blocks = e->num_blocks = cfg->num_blocks > 0 ? cfg->num_blocks : 4096;
  sc.enable_chunked_prefill = cfg->enable_chunked_prefill != 0;
  sc.eos_token_id = e->eos;
  e->sched.reset(new Scheduler(sc));
  *out = e;
  return 0;
}
extern "C" void kge_engine_destroy(kge_engine* e) { delete e; }
extern "C" int kge_engine_num_blocks(const kge_engine* e) { return e ? e->num_blocks : 0; }
extern "C" int kge_engine_vocab(const kge_engine* e) { return e ? e->vocab : 0; }
extern "C" int kge_engine_eos_token_id(const kge_engine* e) { return e ? e->eos : -1; }
extern "C" int kge_engine_max_model_len(const kge_engine* e) { return e ? e->max_model_len : 0; }
extern "C" int64_t kge_engine_weight_bytes(const kge_engine*) { return 0; }
extern "C" void kge_engine_set_reserve_full_isl(kge_engine* e, int32_t on) { if (e) e->sched->set_reserve_full_isl(on != 0); }
extern "C" void kge_engine_set_prefix_caching(kge_engine* e, int32_t on) { if (e) e->sched->set_prefix_caching(on != 0); }
extern "C" int kge_nccl_unique_id(char*) { set_err("mock engine: no tensor parallelism"); return -1; }
extern "C" int kge_engine_init_tp(kge_engine*, const char*) { set_err("mock engine: no tensor parallelism"); return -1; }
extern "C" int kge_engine_tp_info(kge_engine*, kge_tp_peer_info*) { set_err("mock engine: no tensor parallelism"); return -1; }
extern "C" int kge_engine_tp_connect(kge_engine*, const kge_tp_peer_info*) { set_err("mock engine: no tensor parallelism"); return -1; }

extern "C" int kge_add_request(kge_engine* e, int64_t id, const int32_t* prompt, int32_t n, const kge_sampling_params* p) {
  if (!e || !prompt || !p) { set_err("kge_add_request: null argument"); return -1; }
  for (int i = 0; i < n; ++i)
    if (prompt[i] < 0 || prompt[i] >= e->vocab) { set_err("kge_add_request: token id %d out of vocabulary", prompt[i]); return -1; }
  SamplingParams sp;
  sp.max_tokens = p->max_tokens; sp.min_tokens = p->min_tokens; sp.ignore_eos = p->ignore_eos != 0;
  sp.temperature = p->temperature; sp.top_k = p->top_k; sp.top_p = p->top_p; sp.seed = p->seed; sp.has_seed = p->has_seed != 0;
  if (p->stop_token_ids && p->num_stop_token_ids > 0) sp.stop_token_ids.assign(p->stop_token_ids, p->stop_token_ids + p->num_stop_token_ids);
  const char* why = nullptr;
  if (!e->sched->add(id, prompt, n, sp, &why)) { set_err("kge_add_request: %s", why ? why : "rejected"); return -1; }
  if (const char* path = getenv("MOCK_KGE_LOG")) {
    if (FILE* f = fopen(path, "a")) {
      fprintf(f, "{\"id\":%lld,\"prompt_len\":%d,\"max_tokens\":%d,\"min_tokens\":%d,\"ignore_eos\":%d,\"temperature\":%.6g,\"top_k\":%d,"
                 "\"top_p\":%.6g,\"seed\":%llu,\"stop_token_ids\":[", (long long)id, n, p->max_tokens, p->min_tokens, p->ignore_eos,
              (double)p->temperature, p->top_k, (double)p->top_p, (unsigned long long)p->seed);
      for (size_t i = 0; i < sp.stop_token_ids.size(); ++i) fprintf(f, "%s%d", i ? "," : "", sp.stop_token_ids[i]);
      fprintf(f, "]}\n");
      fclose(f);
    }
  }
  return 0;
}

extern "C" int kge_request_set_logprobs(kge_engine* e, int64_t id, int32_t top_n) {
  if (!e || !e->sched->find(id)) { set_err("kge_request_set_logprobs: unknown request"); return -1; }
  if (top_n < -1 || top_n > 20) { set_err("kge_request_set_logprobs: top_n=%d must be in [-1, 20]", top_n); return -1; }
  if (top_n >= 0) e->want_lp[id] = top_n; else e->want_lp.erase(id);
  return 0;
}

extern "C" int kge_abort_request(kge_engine* e, int64_t id) {
  if (!e) return -1;
  e->want_lp.erase(id);
  return e->sched->abort(id) ? 0 : -1;
}

extern "C" int kge_has_unfinished(const kge_engine* e) { return e && e->sched->has_unfinished() ? 1 : 0; }

extern "C" int kge_engine_step(kge_engine* e, kge_step_output* outs, int32_t cap) {
  if (!e || !outs) { set_err("kge_engine_step: null argument"); return -1; }
  e->last_lp.clear();
  const std::vector<Scheduled>& sc = e->sched->schedule();
  if (sc.empty()) return 0;
  if ((int)sc.size() > cap) { set_err("kge_engine_step: output capacity %d < %zu", cap, sc.size()); return -1; }
  std::vector<int32_t> sampled;
  for (auto& s : sc) {
    if (!Scheduler::needs_sample(s)) continue;
    Request* r = s.req;
    const int pos = r->num_tokens();
    const int last = r->token_at(pos - 1);
    const int32_t tok = (int32_t)((7ll * last + 3ll * pos + 11) % e->vocab);
    sampled.push_back(tok);
    auto it = e->want_lp.find(r->id);
    if (it != e->want_lp.end()) e->last_lp[r->id] = {tok, it->second};
  }
  std::vector<StepResult> res;
  e->sched->update(sampled, &res);
  for (size_t i = 0; i < res.size(); ++i) {
    outs[i].request_id = res[i].request_id; outs[i].token = res[i].token;
    outs[i].finished = res[i].finished; outs[i].num_scheduled = res[i].num_scheduled;
    if (res[i].finished) e->want_lp.erase(res[i].request_id);
  }
  if (const char* us = getenv("MOCK_KGE_STEP_US")) usleep((useconds_t)atoi(us));    // lets a test abort a request mid-flight
  return (int)res.size();
}

extern "C" int kge_engine_step_logprobs(const kge_engine* e, int64_t id, float* vals, int32_t* ids, int32_t* rank, int32_t cap) {
  if (!e || !vals || !ids) { set_err("kge_engine_step_logprobs: null argument"); return -1; }
  auto it = e->last_lp.find(id);
  if (it == e->last_lp.end()) return 0;
  const int n = it->second.second + 1;
  if (cap < n) { set_err("kge_engine_step_logprobs: cap"); return -1; }
  vals[0] = -0.5f; ids[0] = it->second.first;
  for (int k = 1; k < n; ++k) {                 // top-1 is the sampled token itself (greedy), then its successors
    vals[k] = -0.5f - (float)(k - 1);
    ids[k] = (it->second.first + (k - 1)) % e->vocab;
  }
  if (rank) *rank = 1;
  return n;
}

extern "C" int kge_engine_stats(const kge_engine* e, int32_t* running, int32_t* waiting, float* kv_usage, int64_t* preemptions) {
  if (!e) return -1;
  if (running) *running = e->sched->num_running();
  if (waiting) *waiting = e->sched->num_waiting();
  if (kv_usage) *kv_usage = 1.f - (float)e->sched->free_blocks() / (float)(e->sched->usable_blocks() > 0 ? e->sched->usable_blocks() : 1);
  if (preemptions) *preemptions = e->sched->num_preemptions();
  return 0;
}

extern "C" int kge_engine_prefix_cache_stats(const kge_engine* e, int64_t* q, int64_t* h) {
  if (!e) return -1;
  if (q) *q = e->sched->prefix_cache_queries();
  if (h) *h = e->sched->prefix_cache_hits();
  return 0;
}
