// kge-serve: the container entrypoint the reference's values files expect (`vllm serve MODEL ...`,
// installed as bin/vllm -> kge-serve), serving the OpenAI-compatible surface of the engine it
// replaces: POST /v1/completions, POST /v1/chat/completions, GET /v1/models, GET /health,
// GET /metrics (vllm:entrypoints/openai/completion/api_router.py:34-66, .../models/api_router.py:20,
// vllm:entrypoints/serve/instrumentator/health.py:22-35, metrics.py:30-44).
//
// One engine thread owns the kge_engine handles (one per tensor-parallel rank) and is the only
// caller of kge_add_request / kge_engine_step; HTTP connections are served by their own threads and
// never touch the GPU, so kubelet's /health probes are answered while a step is running.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "args.h"
#include "json.h"
#include "kge.h"
#include "tokenizer.h"
#include "stop_strings.h"
#include "nfc.h"
#include "genconfig.h"

using namespace kge;

namespace {

// Prometheus histogram with vLLM's metric names and bucket edges (vllm:v1/metrics/loggers.py).
struct Histogram {
  std::vector<double> edges;
  std::vector<uint64_t> counts;      // per bucket, not cumulative; last = +Inf
  double sum = 0;
  uint64_t n = 0;
  std::mutex mu;
  explicit Histogram(std::vector<double> e) : edges(std::move(e)), counts(edges.size() + 1, 0) {}
  void observe(double v) {
    std::lock_guard<std::mutex> lk(mu);
    size_t i = 0;
    while (i < edges.size() && v > edges[i]) ++i;
    ++counts[i]; sum += v; ++n;
  }
  std::string text(const char* name, const char* help, const std::string& label) {
    std::lock_guard<std::mutex> lk(mu);
    std::string o = std::string("# HELP ") + name + " " + help + "\n# TYPE " + name + " histogram\n";
    uint64_t cum = 0;
    char b[64];
    for (size_t i = 0; i <= edges.size(); ++i) {
      cum += counts[i];
      if (i < edges.size()) snprintf(b, sizeof b, "%g", edges[i]); else snprintf(b, sizeof b, "+Inf");
      o += std::string(name) + "_bucket{" + label + ",le=\"" + b + "\"} " + std::to_string(cum) + "\n";
    }
    snprintf(b, sizeof b, "%.9g", sum);
    o += std::string(name) + "_sum{" + label + "} " + b + "\n" + name + "_count{" + label + "} " + std::to_string(n) + "\n";
    return o;
  }
};

struct ReqState {
  std::mutex mu;
  std::condition_variable cv;
  std::vector<int32_t> tokens;
  // logprobs requested: per generated token, (token id, logprob) with the sampled token first, then the top-N
  std::vector<std::vector<std::pair<int32_t, float>>> lps;
  int finished = 0;            // 0 running, 1 stop, 2 length, 3 error/aborted
  std::string error;
  int prompt_len = 0;
  std::chrono::steady_clock::time_point t_arrival = std::chrono::steady_clock::now(), t_last;   // latency metrics
};

struct Submit {
  int64_t id;
  std::vector<int32_t> prompt;
  kge_sampling_params sp;
  std::vector<int32_t> stop_ids;
  int logprobs = -1;           // SamplingParams.logprobs: top-N per token, -1 = none
  std::shared_ptr<ReqState> st;
};

struct Server {
  ServeArgs args;
  std::vector<kge_engine*> engines;       // one per TP rank
  int vocab = 0;
  int eos = -1;
  GenDefaults gen;                        // generation_config.json defaults and extra end-of-sequence ids
  Tokenizer tok;                          // loaded when the model directory has a tokenizer.json
  bool have_tok = false;
  std::mutex mu;
  std::condition_variable cv;
  std::deque<Submit> submits;
  std::deque<int64_t> aborts;
  std::map<int64_t, std::shared_ptr<ReqState>> live;   // engine thread only
  std::atomic<bool> stop{false}, healthy{true};
  std::atomic<int64_t> next_id{1};
  std::atomic<int64_t> prompt_tokens{0}, generation_tokens{0}, success_stop{0}, success_length{0}, num_steps{0};
  std::atomic<int> running{0}, waiting{0}, active_conns{0};
  std::atomic<int64_t> preemptions{0}, pc_queries{0}, pc_hits{0};
  Histogram h_ttft{{0.001, 0.005, 0.01, 0.02, 0.04, 0.06, 0.08, 0.1, 0.25, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0, 20.0, 40.0, 80.0, 160.0, 640.0, 2560.0}};
  Histogram h_itl{{0.01, 0.025, 0.05, 0.075, 0.1, 0.15, 0.2, 0.3, 0.4, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0, 20.0, 40.0, 80.0}};
  Histogram h_e2e{{0.3, 0.5, 0.8, 1.0, 1.5, 2.0, 2.5, 5.0, 10.0, 15.0, 20.0, 30.0, 40.0, 50.0, 60.0, 120.0, 240.0, 480.0, 960.0, 1920.0, 7680.0}};
  std::atomic<float> kv_usage{0.f};
  time_t created = time(nullptr);
};

Server* g_srv = nullptr;

// ---------------------------------------------------------------------------------- engine thread
// Tensor parallel: every rank runs the identical step so that the exchanges inside it can meet.  Ranks > 0 live in
// worker threads that stay for the life of the server (one per rank, started on the first step) and are handed each
// step through a generation counter; rank 0 steps in the engine-loop thread.
struct RankWorkers {
  std::mutex mu;
  std::condition_variable cv_go, cv_done;
  uint64_t gen = 0;
  int pending = 0, cap = 0;
  bool quit = false;
  std::vector<int> rcs;
  std::vector<std::vector<kge_step_output>> tmp;
  std::vector<std::thread> th;
  ~RankWorkers() {
    { std::lock_guard<std::mutex> lk(mu); quit = true; }
    cv_go.notify_all();
    for (auto& t : th) t.join();
  }
};
static RankWorkers g_workers;

int step_all(Server& s, kge_step_output* outs, int cap) {
  if (s.engines.size() == 1) return kge_engine_step(s.engines[0], outs, cap);
  RankWorkers& w = g_workers;
  if (w.th.empty()) {
    w.rcs.assign(s.engines.size(), 0);
    w.tmp.resize(s.engines.size());
    for (size_t r = 1; r < s.engines.size(); ++r) {
      w.tmp[r].resize(cap);
      w.th.emplace_back([&w, &s, r] {
        uint64_t seen = 0;
        for (;;) {
          int c;
          {
            std::unique_lock<std::mutex> lk(w.mu);
            w.cv_go.wait(lk, [&] { return w.quit || w.gen != seen; });
            if (w.quit) return;
            seen = w.gen;
            c = w.cap;
          }
          const int rc = kge_engine_step(s.engines[r], w.tmp[r].data(), c);
          std::lock_guard<std::mutex> lk(w.mu);
          w.rcs[r] = rc;
          if (--w.pending == 0) w.cv_done.notify_one();
        }
      });
    }
  }
  {
    std::lock_guard<std::mutex> lk(w.mu);
    w.cap = std::min<int>(cap, (int)w.tmp[1].size());
    w.pending = (int)s.engines.size() - 1;
    ++w.gen;
  }
  w.cv_go.notify_all();
  const int n = kge_engine_step(s.engines[0], outs, cap);
  std::unique_lock<std::mutex> lk(w.mu);
  w.cv_done.wait(lk, [&] { return w.pending == 0; });
  for (size_t r = 1; r < s.engines.size(); ++r)
    if (w.rcs[r] < 0 || w.rcs[r] != n) return -3;
  return n;
}

void engine_loop(Server* sp) {
  Server& s = *sp;
  std::vector<kge_step_output> outs(8192);
  int idle_steps = 0;
  while (!s.stop) {
    std::deque<Submit> subs;
    std::deque<int64_t> abs_;
    {
      std::unique_lock<std::mutex> lk(s.mu);
      if (s.submits.empty() && s.aborts.empty() && !kge_has_unfinished(s.engines[0]))
        s.cv.wait_for(lk, std::chrono::milliseconds(50));
      subs.swap(s.submits);
      abs_.swap(s.aborts);
    }
    for (auto& sub : subs) {
      sub.sp.stop_token_ids = sub.stop_ids.empty() ? nullptr : sub.stop_ids.data();
      sub.sp.num_stop_token_ids = (int)sub.stop_ids.size();
      int rc = 0;
      for (auto* e : s.engines) {
        rc = kge_add_request(e, sub.id, sub.prompt.data(), (int)sub.prompt.size(), &sub.sp);
        if (!rc && sub.logprobs >= 0) rc = kge_request_set_logprobs(e, sub.id, sub.logprobs);
        if (rc) break;
      }
      if (rc) {
        std::lock_guard<std::mutex> lk(sub.st->mu);
        sub.st->error = kge_last_error();
        sub.st->finished = 3;
        sub.st->cv.notify_all();
      } else {
        s.live[sub.id] = sub.st;
        s.prompt_tokens += (int64_t)sub.prompt.size();
      }
    }
    for (int64_t id : abs_) {
      for (auto* e : s.engines) kge_abort_request(e, id);
      s.live.erase(id);
    }
    if (!kge_has_unfinished(s.engines[0])) continue;
    const int n = step_all(s, outs.data(), (int)outs.size());
    if (n < 0) {
      fprintf(stderr, "engine step failed: %s\n", kge_last_error());
      s.healthy = false;
      for (auto& kv : s.live) {
        std::lock_guard<std::mutex> lk(kv.second->mu);
        kv.second->error = std::string("engine step failed: ") + kge_last_error();
        kv.second->finished = 3;
        kv.second->cv.notify_all();
      }
      s.live.clear();
      s.stop = true;
      break;
    }
    if (n == 0) {
      // Unfinished requests but nothing could be scheduled.  kge_engine_create guarantees that one request always fits the
      // KV pool, so this should not happen; if it does, spinning here would hang every client while /health stays 200.
      // Fail the requests that are stuck and carry on serving.
      if (++idle_steps >= 3) {
        fprintf(stderr, "engine made no progress with %zu live requests: failing them\n", s.live.size());
        for (auto& kv : s.live) {
          for (auto* e : s.engines) kge_abort_request(e, kv.first);
          std::lock_guard<std::mutex> lk(kv.second->mu);
          kv.second->error = "the engine could not schedule this request (KV cache too small for it)";
          kv.second->finished = 3;
          kv.second->cv.notify_all();
        }
        s.live.clear();
        idle_steps = 0;
      } else {
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
      }
      continue;
    }
    idle_steps = 0;
    ++s.num_steps;
    for (int i = 0; i < n; ++i) {
      auto it = s.live.find(outs[i].request_id);
      if (it == s.live.end()) continue;
      auto st = it->second;
      {
        std::lock_guard<std::mutex> lk(st->mu);
        if (outs[i].token >= 0) {
          float lv[21]; int32_t li[21], rk = 0;
          const int m = kge_engine_step_logprobs(s.engines[0], outs[i].request_id, lv, li, &rk, 21);
          if (m > 0) {
            st->lps.emplace_back();
            for (int k = 0; k < m; ++k) st->lps.back().push_back({li[k], lv[k]});
          }
          const auto now = std::chrono::steady_clock::now();
          if (st->tokens.empty()) s.h_ttft.observe(std::chrono::duration<double>(now - st->t_arrival).count());
          else s.h_itl.observe(std::chrono::duration<double>(now - st->t_last).count());
          st->t_last = now;
          st->tokens.push_back(outs[i].token);
          ++s.generation_tokens;
        }
        if (outs[i].finished) {
          st->finished = outs[i].finished;
          s.h_e2e.observe(std::chrono::duration<double>(std::chrono::steady_clock::now() - st->t_arrival).count());
        }
      }
      st->cv.notify_all();
      if (outs[i].finished) {
        (outs[i].finished == 1 ? s.success_stop : s.success_length)++;
        s.live.erase(it);
      }
    }
    int32_t r = 0, w = 0;
    float u = 0;
    int64_t p = 0;
    kge_engine_stats(s.engines[0], &r, &w, &u, &p);
    s.running = r; s.waiting = w; s.kv_usage = u; s.preemptions = p;
    int64_t pq = 0, ph = 0;
    kge_engine_prefix_cache_stats(s.engines[0], &pq, &ph);
    s.pc_queries = pq; s.pc_hits = ph;
  }
  // shutting down: wake every connection that is still waiting for tokens, and the ones that were never admitted
  {
    std::lock_guard<std::mutex> lk(s.mu);
    for (auto& sub : s.submits) s.live[sub.id] = sub.st;
    s.submits.clear();
  }
  for (auto& kv : s.live) {
    std::lock_guard<std::mutex> lk(kv.second->mu);
    if (!kv.second->finished) { kv.second->error = "the server is shutting down"; kv.second->finished = 3; }
    kv.second->cv.notify_all();
  }
  s.live.clear();
}

// ---------------------------------------------------------------------------------- HTTP helpers
struct HttpReq {
  std::string method, path, body, authorization;
  bool keep_alive = true;
};

bool send_all(int fd, const char* p, size_t n) {
  while (n) {
    ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
    if (w <= 0) return false;
    p += w; n -= (size_t)w;
  }
  return true;
}
bool send_str(int fd, const std::string& s) { return send_all(fd, s.data(), s.size()); }

// returns 1 ok, 0 clean EOF, -1 error
int read_request(int fd, std::string& buf, HttpReq* r) {
  size_t hend;
  while ((hend = buf.find("\r\n\r\n")) == std::string::npos) {
    if (buf.size() > (1 << 20)) return -1;
    char tmp[8192];
    ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
    if (n == 0) return buf.empty() ? 0 : -1;
    if (n < 0) return -1;
    buf.append(tmp, (size_t)n);
  }
  std::string head = buf.substr(0, hend);
  size_t le = head.find("\r\n");
  std::string line = head.substr(0, le);
  size_t s1 = line.find(' '), s2 = line.rfind(' ');
  if (s1 == std::string::npos || s2 == s1) return -1;
  r->method = line.substr(0, s1);
  r->path = line.substr(s1 + 1, s2 - s1 - 1);
  size_t q = r->path.find('?');
  if (q != std::string::npos) r->path.resize(q);
  r->keep_alive = line.find("HTTP/1.1") != std::string::npos;
  size_t clen = 0;
  size_t pos = le == std::string::npos ? head.size() : le + 2;
  while (pos < head.size()) {
    size_t e = head.find("\r\n", pos);
    if (e == std::string::npos) e = head.size();
    std::string h = head.substr(pos, e - pos);
    pos = e + 2;
    size_t c = h.find(':');
    if (c == std::string::npos) continue;
    std::string k = h.substr(0, c), v = h.substr(c + 1);
    for (auto& ch : k) ch = (char)tolower(ch);
    while (!v.empty() && v[0] == ' ') v.erase(0, 1);
    if (k == "content-length") clen = (size_t)strtoull(v.c_str(), nullptr, 10);
    if (k == "authorization") r->authorization = v;
    if (k == "connection") {
      for (auto& ch : v) ch = (char)tolower(ch);
      if (v.find("close") != std::string::npos) r->keep_alive = false;
      if (v.find("keep-alive") != std::string::npos) r->keep_alive = true;
    }
  }
  if (clen > (64u << 20)) return -1;
  buf.erase(0, hend + 4);
  while (buf.size() < clen) {
    char tmp[65536];
    ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
    if (n <= 0) return -1;
    buf.append(tmp, (size_t)n);
  }
  r->body = buf.substr(0, clen);
  buf.erase(0, clen);
  return 1;
}

const char* status_text(int code) {
  switch (code) {
    case 200: return "OK";
    case 400: return "Bad Request";
    case 401: return "Unauthorized";
    case 404: return "Not Found";
    case 405: return "Method Not Allowed";
    case 500: return "Internal Server Error";
    case 503: return "Service Unavailable";
  }
  return "OK";
}

bool respond(int fd, int code, const std::string& ctype, const std::string& body, bool keep) {
  std::string h = "HTTP/1.1 " + std::to_string(code) + " " + status_text(code) + "\r\n";
  if (!ctype.empty()) h += "Content-Type: " + ctype + "\r\n";
  h += "Content-Length: " + std::to_string(body.size()) + "\r\n";
  h += keep ? "Connection: keep-alive\r\n\r\n" : "Connection: close\r\n\r\n";
  return send_str(fd, h + body);
}

// {"error":{"message","type","param","code"}} with HTTP status = code (vllm:entrypoints/openai/engine/protocol.py:59-67)
bool respond_error(int fd, int code, const std::string& msg, const std::string& type, bool keep) {
  std::string b = "{\"error\":{\"message\":\"" + json_escape(msg) + "\",\"type\":\"" + type + "\",\"param\":null,\"code\":" +
                  std::to_string(code) + "}}";
  return respond(fd, code, "application/json", b, keep);
}

bool send_chunk(int fd, const std::string& data) {
  char hdr[32];
  snprintf(hdr, sizeof hdr, "%zx\r\n", data.size());
  return send_str(fd, std::string(hdr) + data + "\r\n");
}

std::string ids_json(const std::vector<int32_t>& v, size_t from = 0) {
  std::string s = "[";
  for (size_t i = from; i < v.size(); ++i) { if (i > from) s += ","; s += std::to_string(v[i]); }
  return s + "]";
}

// ---------------------------------------------------------------------------------- completions
struct ParsedReq {
  std::vector<std::vector<int32_t>> prompts;
  kge_sampling_params sp{};
  std::vector<int32_t> stop_ids;
  std::vector<std::string> stop;            // stop strings (need the tokenizer)
  bool include_stop = false;                // include_stop_str_in_output
  int logprobs = -1;                        // top-N log-probabilities per token, -1 = none
  bool max_tokens_given = true;
  bool given_temperature = true, given_top_p = true, given_top_k = true;
  bool stream = false, include_usage = false, echo = false;
  int n = 1;
  std::string model;
};

// prompt: str | [str] | [int] | [[int]] (vllm:entrypoints/openai/completion/protocol.py:46-52); strings need
// the model directory to carry a tokenizer.json
const char* kNoTokenizer = "string prompts need a tokenizer.json in the model directory (none was found): send token ids "
                           "(prompt: [int] or [[int]])";
std::string parse_prompt(const Server& s, const Json* p, std::vector<std::vector<int32_t>>* out) {
  const int vocab = s.vocab;
  if (!p) return "prompt is required";
  if (p->is_str()) {
    if (!s.have_tok) return kNoTokenizer;
    out->push_back(s.tok.encode(p->str, true));
    if (out->back().empty()) return "prompt is empty";
    return "";
  }
  if (!p->is_arr() || p->arr.empty()) return "prompt must be a non-empty string, array of strings or array of token ids";
  auto one = [&](const Json& a, std::vector<int32_t>* v) -> std::string {
    for (auto& t : a.arr) {
      if (!t.is_num()) return "prompt arrays must hold integers";
      const int64_t id = t.as_i64();
      if (id < 0 || id >= vocab) return "token id " + std::to_string(id) + " is out of vocabulary";
      v->push_back((int32_t)id);
    }
    if (v->empty()) return "prompt is empty";
    return "";
  };
  if (p->arr[0].is_num()) {
    out->emplace_back();
    return one(*p, &out->back());
  }
  if (p->arr[0].is_str()) {
    if (!s.have_tok) return kNoTokenizer;
    for (auto& a : p->arr) {
      if (!a.is_str()) return "prompt array mixes strings and other types";
      out->push_back(s.tok.encode(a.str, true));
      if (out->back().empty()) return "prompt is empty";
    }
    return "";
  }
  for (auto& a : p->arr) {
    if (!a.is_arr()) return "prompt must be [int] or [[int]]";
    out->emplace_back();
    std::string e = one(a, &out->back());
    if (!e.empty()) return e;
  }
  return "";
}

std::string parse_sampling(const Json& j, ParsedReq* r, bool chat) {
  // completions default to 16 tokens (completion/protocol.py:55); chat requests without a limit may use the rest of the
  // context (chat_completion/protocol.py: max_tokens = None -> max_model_len - prompt length), resolved once the prompt is known
  auto I = [&](const char* k, int64_t d) { return (int)std::max<int64_t>(INT32_MIN, std::min<int64_t>(j.int_or(k, d), INT32_MAX)); };
  r->max_tokens_given = j.get("max_tokens") || (chat && j.get("max_completion_tokens"));
  r->sp.max_tokens = I(chat ? "max_completion_tokens" : "max_tokens", I("max_tokens", 16));
  r->sp.min_tokens = I("min_tokens", 0);
  r->sp.ignore_eos = j.bool_or("ignore_eos", false) ? 1 : 0;
  r->sp.temperature = (float)j.num_or("temperature", 1.0);     // omitted => 1.0 (vllm:sampling_params.py:205) unless the model says otherwise
  r->sp.top_p = (float)j.num_or("top_p", 1.0);
  r->sp.top_k = I("top_k", 0);
  r->given_temperature = j.get("temperature") && j.get("temperature")->is_num();
  r->given_top_p = j.get("top_p") && j.get("top_p")->is_num();
  r->given_top_k = j.get("top_k") && j.get("top_k")->is_num();
  if (r->sp.top_k < 0) r->sp.top_k = 0;
  r->sp.seed = (uint64_t)j.int_or("seed", 0);
  r->sp.has_seed = (j.get("seed") && j.get("seed")->is_num()) ? 1 : 0;     // "seed": 0 is a seed
  r->n = I("n", 1);
  r->stream = j.bool_or("stream", false);
  r->echo = j.bool_or("echo", false);
  r->model = j.str_or("model", "");
  if (const Json* so = j.get("stream_options")) r->include_usage = so->bool_or("include_usage", false);
  if (const Json* st = j.get("stop_token_ids"))
    if (st->is_arr()) for (auto& t : st->arr) if (t.is_num()) r->stop_ids.push_back((int32_t)std::max<int64_t>(-1, std::min<int64_t>(t.as_i64(), INT32_MAX)));
  if (const Json* st = j.get("stop")) {       // str | [str] (vllm:entrypoints/openai/completion/protocol.py:62)
    if (st->is_str()) { if (!st->str.empty()) r->stop.push_back(st->str); }
    else if (st->is_arr()) { for (auto& x : st->arr) { if (!x.is_str()) return "stop must be a string or a list of strings"; if (!x.str.empty()) r->stop.push_back(x.str); } }
    else if (!st->is_null()) return "stop must be a string or a list of strings";
  }
  r->include_stop = j.bool_or("include_stop_str_in_output", false);
  // completions: logprobs = N (vllm:entrypoints/openai/completion/protocol.py:57); chat: logprobs = true + top_logprobs = N
  if (const Json* lp = j.get("logprobs")) {
    if (chat) { if (lp->type == Json::Bool && lp->b) r->logprobs = I("top_logprobs", 0); }
    else if (lp->is_num()) r->logprobs = (int)std::max<int64_t>(-2, std::min<int64_t>(lp->as_i64(), 1000));
    else if (!lp->is_null() && !(lp->type == Json::Bool && !lp->b)) return "logprobs must be an integer";
  }
  if (r->logprobs < -1 || r->logprobs > 20) return "logprobs / top_logprobs must be in [0, 20]";
  if (r->sp.max_tokens < 1) return "max_tokens must be at least 1";
  if (r->sp.temperature < 0) return "temperature must be non-negative";
  if (r->sp.top_p <= 0 || r->sp.top_p > 1) return "top_p must be in (0, 1]";
  if (r->n < 1 || r->n > 64) return "n must be in [1, 64]";
  return "";
}

void push_abort(Server& s, int64_t id) {
  std::lock_guard<std::mutex> lk(s.mu);
  s.aborts.push_back(id);
  s.cv.notify_all();
}

// --- logprobs in the OpenAI response shapes (vllm:entrypoints/openai/completion/serving.py _create_completion_logprobs,
// chat_completion/serving.py _create_chat_logprobs).  Token strings come from the tokenizer, or are "token_id:N".
using LpEntry = std::vector<std::pair<int32_t, float>>;
std::string lp_token(const Server& s, int32_t id) {
  return s.have_tok ? s.tok.decode(std::vector<int32_t>{id}) : "token_id:" + std::to_string(id);
}
std::string lp_num(float v) {
  char b[32];
  snprintf(b, sizeof b, "%.9g", (double)(v < -9999.f ? -9999.f : v));
  return b;
}
std::string lp_bytes(const std::string& t) {
  std::string o = "[";
  for (size_t i = 0; i < t.size(); ++i) o += (i ? "," : "") + std::to_string((int)(unsigned char)t[i]);
  return o + "]";
}
// completions: {"text_offset","token_logprobs","tokens","top_logprobs"} for entries [from, lps.size())
std::string lp_completion_json(const Server& s, const std::vector<LpEntry>& lps, size_t from, size_t* offset) {
  std::string off = "[", tl = "[", tk = "[", top = "[";
  for (size_t i = from; i < lps.size(); ++i) {
    const std::string sep = i > from ? "," : "";
    const std::string t = lp_token(s, lps[i][0].first);
    off += sep + std::to_string(*offset);
    *offset += utf8_chars(t);
    tl += sep + lp_num(lps[i][0].second);
    tk += sep + "\"" + json_escape(t) + "\"";
    top += sep + "{";
    for (size_t k = 0; k < lps[i].size(); ++k)
      top += (k ? ",\"" : "\"") + json_escape(lp_token(s, lps[i][k].first)) + "\":" + lp_num(lps[i][k].second);
    top += "}";
  }
  return "{\"text_offset\":" + off + "],\"token_logprobs\":" + tl + "],\"tokens\":" + tk + "],\"top_logprobs\":" + top + "]}";
}
// chat: {"content":[{"token","logprob","bytes","top_logprobs":[{"token","logprob","bytes"}...]}]}
std::string lp_chat_json(const Server& s, const std::vector<LpEntry>& lps, size_t from, int top_n) {
  std::string c = "{\"content\":[";
  for (size_t i = from; i < lps.size(); ++i) {
    auto one = [&](const std::pair<int32_t, float>& e) {
      const std::string t = lp_token(s, e.first);
      return "\"token\":\"" + json_escape(t) + "\",\"logprob\":" + lp_num(e.second) + ",\"bytes\":" + lp_bytes(t);
    };
    c += (i > from ? ",{" : "{") + one(lps[i][0]) + ",\"top_logprobs\":[";
    for (size_t k = 0; k < lps[i].size() && (int)k < top_n; ++k) c += (k ? ",{" : "{") + one(lps[i][k]) + "}";
    c += "]}";
  }
  return c + "]}";
}

bool handle_generate(Server& s, int fd, const HttpReq& hr, bool chat) {
  Json j;
  std::string perr;
  if (!json_parse(hr.body, &j, &perr) || !j.is_obj()) return respond_error(fd, 400, "invalid JSON body: " + perr, "BadRequestError", hr.keep_alive);
  ParsedReq pr;
  std::string err = parse_sampling(j, &pr, chat);
  if (err.empty()) {
    // what the model's generation_config.json supplies when the request is silent (vLLM: --generation-config auto)
    if (!pr.given_temperature && s.gen.has_temperature) pr.sp.temperature = (float)s.gen.temperature;
    if (!pr.given_top_p && s.gen.has_top_p) pr.sp.top_p = (float)s.gen.top_p;
    if (!pr.given_top_k && s.gen.has_top_k) pr.sp.top_k = s.gen.top_k;
    if (!pr.sp.ignore_eos)
      for (int32_t id : s.gen.extra_eos)
        if (std::find(pr.stop_ids.begin(), pr.stop_ids.end(), id) == pr.stop_ids.end()) pr.stop_ids.push_back(id);
  }
  if (err.empty()) {
    if (chat) {
      // messages are rendered with the model family's chat format (Llama-3 headers or ChatML) and tokenised;
      // without a tokenizer / known format, token ids may be passed in `prompt_token_ids`
      const Json* msgs = j.get("messages");
      if (msgs && msgs->is_arr() && !msgs->arr.empty() && s.have_tok && s.tok.has_chat_template()) {
        std::vector<ChatMessage> cm;
        for (auto& m : msgs->arr) {
          if (!m.is_obj()) { err = "messages must be objects with role and content"; break; }
          const Json* c = m.get("content");
          std::string text;
          if (c && c->is_str()) text = c->str;
          else if (c && c->is_arr()) { for (auto& part : c->arr) if (part.is_obj() && part.str_or("type", "") == "text") text += part.str_or("text", ""); }
          cm.push_back({m.str_or("role", "user"), text});
        }
        if (err.empty()) pr.prompts.push_back(s.tok.encode(s.tok.render_chat(cm), false));
      } else {
        err = parse_prompt(s, j.get("prompt_token_ids"), &pr.prompts);
        if (!err.empty()) err = "chat completions need a tokenizer.json and a supported chat template (Llama-3 or ChatML) in the model "
                                "directory; otherwise pass prompt_token_ids: [int] (" + err + ")";
      }
    } else {
      err = parse_prompt(s, j.get("prompt"), &pr.prompts);
    }
  }
  if (err.empty() && pr.echo && (!s.have_tok || chat))
    err = chat ? "echo is a completions parameter" : "echo needs a tokenizer.json in the model directory (none was found)";
  if (err.empty() && !pr.stop.empty() && !s.have_tok)
    err = "stop strings need a tokenizer.json in the model directory (none was found); use stop_token_ids";
  if (!err.empty()) return respond_error(fd, 400, err, "BadRequestError", hr.keep_alive);
  if (!pr.model.empty() && pr.model != s.args.served_model_name)
    return respond_error(fd, 404, "The model `" + pr.model + "` does not exist.", "NotFoundError", hr.keep_alive);

  // One choice of the response.  Tokens are consumed as the engine produces them, in both modes: the text stream
  // (detokeniser + stop strings, host/stop_strings.h) decides when a stop string ends the request early.
  struct Sub {
    int64_t id; std::shared_ptr<ReqState> st;
    size_t sent = 0; bool done = false;
    TextStream ts; std::vector<int32_t> toks; int fin = 0; std::string stop_reason = "null";
    std::string echo;          // echo=true: the prompt's text, sent in front of the completion (completion/serving.py)
    std::vector<std::vector<std::pair<int32_t, float>>> lps;   // per consumed token: sampled first, then the top-N, no duplicates
    size_t lp_chars = 0;       // text_offset of the next token (characters of the token strings so far)
  };
  std::vector<Sub> subs;
  int prompt_tokens = 0;
  {
    std::lock_guard<std::mutex> lk(s.mu);
    for (auto& p : pr.prompts)
      for (int k = 0; k < pr.n; ++k) {
        Submit sb;
        sb.id = s.next_id++;
        sb.prompt = p;
        sb.sp = pr.sp;
        if (chat && !pr.max_tokens_given) sb.sp.max_tokens = std::max(1, s.args.max_model_len - (int)p.size());
        if (pr.n > 1 && sb.sp.has_seed) sb.sp.seed += (uint64_t)k;
        sb.stop_ids = pr.stop_ids;
        sb.logprobs = pr.logprobs;
        sb.st = std::make_shared<ReqState>();
        sb.st->prompt_len = (int)p.size();
        subs.push_back({sb.id, sb.st});
        subs.back().ts.stop = pr.stop;
        subs.back().ts.include_stop = pr.include_stop;
        subs.back().ts.min_tokens = pr.sp.min_tokens;
        if (pr.echo) subs.back().echo = s.tok.decode(p);
        prompt_tokens += (int)p.size();
        s.submits.push_back(std::move(sb));
      }
    s.cv.notify_all();
  }
  const std::string rid = std::string(chat ? "chatcmpl-" : "cmpl-") + std::to_string(subs[0].id);
  const std::string model = json_escape(s.args.served_model_name);
  const long created = (long)time(nullptr);
  auto reason = [](int f) { return f == 1 ? "\"stop\"" : f == 2 ? "\"length\"" : "null"; };

  // Feed one engine token into choice `sb`; returns the text that may be handed out now.  `fin` is the engine's
  // finish flag and applies to the last token of the batch it came with.
  auto feed = [&](Sub& sb, int32_t tokid, int fin, bool last_of_batch, const LpEntry* lp) -> std::string {
    const bool last = fin != 0 && last_of_batch;
    sb.toks.push_back(tokid);
    if (lp && !lp->empty()) {                       // the sampled token first; drop its duplicate among the top-N
      sb.lps.emplace_back();
      for (size_t k = 0; k < lp->size(); ++k)
        if (k == 0 || (*lp)[k].first != (*lp)[0].first) sb.lps.back().push_back((*lp)[k]);
    }
    if (last) {
      sb.fin = fin;
      // a stop token id (not EOS) is reported as stop_reason (vllm:v1/core/sched/utils.py:105-112)
      if (fin == 1 && !(tokid == s.eos && !pr.sp.ignore_eos))
        for (int32_t x : pr.stop_ids) if (x == tokid) sb.stop_reason = std::to_string(tokid);
    }
    if (!s.have_tok) return "";
    if (sb.ts.push(s.tok, tokid, last && fin == 1, last)) {
      sb.fin = 1;
      sb.stop_reason = "\"" + json_escape(sb.ts.stop[sb.ts.matched]) + "\"";
      if (!last) { push_abort(s, sb.id); ++s.success_stop; }   // the engine still runs this request: free its KV blocks
    }
    return sb.ts.next_delta(sb.fin != 0);
  };

  if (!pr.stream) {
    std::string choices;
    int completion = 0;
    for (size_t i = 0; i < subs.size(); ++i) {
      Sub& sb = subs[i];
      while (!sb.fin) {
        std::vector<int32_t> fresh;
        std::vector<LpEntry> fresh_lp;
        int fin = 0;
        {
          std::unique_lock<std::mutex> lk(sb.st->mu);
          sb.st->cv.wait(lk, [&] { return sb.st->tokens.size() > sb.sent || sb.st->finished != 0; });
          fresh.assign(sb.st->tokens.begin() + sb.sent, sb.st->tokens.end());
          if (sb.st->lps.size() >= sb.st->tokens.size()) fresh_lp.assign(sb.st->lps.begin() + sb.sent, sb.st->lps.end());
          fin = sb.st->finished;
        }
        if (fin == 3) {
          std::string e;
          { std::lock_guard<std::mutex> lk(sb.st->mu); e = sb.st->error; }
          for (auto& o : subs) push_abort(s, o.id);
          return respond_error(fd, 400, e, "BadRequestError", hr.keep_alive);
        }
        for (size_t k = 0; k < fresh.size() && !sb.fin; ++k)
          feed(sb, fresh[k], fin, k + 1 == fresh.size(), k < fresh_lp.size() ? &fresh_lp[k] : nullptr);
        sb.sent += fresh.size();
      }
      completion += (int)sb.toks.size();
      if (i) choices += ",";
      const std::string text = json_escape(sb.echo + sb.ts.text);
      std::string lpj = "null";
      if (pr.logprobs >= 0 && !sb.lps.empty()) {
        size_t off = utf8_chars(sb.echo);
        lpj = chat ? lp_chat_json(s, sb.lps, 0, pr.logprobs) : lp_completion_json(s, sb.lps, 0, &off);
      }
      if (chat)
        choices += "{\"index\":" + std::to_string(i) + ",\"message\":{\"role\":\"assistant\",\"content\":\"" + text + "\"},\"token_ids\":" +
                   ids_json(sb.toks) + ",\"logprobs\":" + lpj + ",\"finish_reason\":" + reason(sb.fin) + ",\"stop_reason\":" + sb.stop_reason + "}";
      else
        choices += "{\"index\":" + std::to_string(i) + ",\"text\":\"" + text + "\",\"token_ids\":" + ids_json(sb.toks) +
                   ",\"logprobs\":" + lpj + ",\"finish_reason\":" + reason(sb.fin) + ",\"stop_reason\":" + sb.stop_reason + "}";
    }
    std::string body = "{\"id\":\"" + rid + "\",\"object\":\"" + (chat ? "chat.completion" : "text_completion") + "\",\"created\":" +
                       std::to_string(created) + ",\"model\":\"" + model + "\",\"choices\":[" + choices + "],\"usage\":{\"prompt_tokens\":" +
                       std::to_string(prompt_tokens) + ",\"total_tokens\":" + std::to_string(prompt_tokens + completion) +
                       ",\"completion_tokens\":" + std::to_string(completion) + "}}";
    return respond(fd, 200, "application/json", body, hr.keep_alive);
  }

  // SSE: data: {chunk}\n\n ... data: [DONE]\n\n (vllm:entrypoints/openai/completion/serving.py:436,463)
  std::string h = "HTTP/1.1 200 OK\r\nContent-Type: text/event-stream; charset=utf-8\r\nCache-Control: no-cache\r\n"
                  "Transfer-Encoding: chunked\r\n";
  h += hr.keep_alive ? "Connection: keep-alive\r\n\r\n" : "Connection: close\r\n\r\n";
  bool ok = send_str(fd, h);
  const char* obj = chat ? "chat.completion.chunk" : "text_completion";
  int completion = 0;
  size_t remaining = subs.size();
  while (ok && remaining) {
    for (size_t i = 0; i < subs.size() && ok; ++i) {
      Sub& sb = subs[i];
      if (sb.done) continue;
      std::vector<int32_t> fresh;
      std::vector<LpEntry> fresh_lp;
      int fin = 0;
      std::string e;
      {
        std::unique_lock<std::mutex> lk(sb.st->mu);
        sb.st->cv.wait_for(lk, std::chrono::milliseconds(subs.size() > 1 ? 2 : 1000),
                           [&] { return sb.st->tokens.size() > sb.sent || sb.st->finished != 0; });
        fresh.assign(sb.st->tokens.begin() + sb.sent, sb.st->tokens.end());
        if (sb.st->lps.size() >= sb.st->tokens.size()) fresh_lp.assign(sb.st->lps.begin() + sb.sent, sb.st->lps.end());
        fin = sb.st->finished;
        e = sb.st->error;
      }
      if (fin == 3) {
        ok = send_chunk(fd, "data: {\"error\":{\"message\":\"" + json_escape(e) + "\",\"type\":\"BadRequestError\",\"param\":null,\"code\":400}}\n\n");
        sb.done = true; --remaining;
        continue;
      }
      for (size_t k = 0; k < fresh.size() && ok && !sb.fin; ++k) {
        std::string piece = feed(sb, fresh[k], fin, k + 1 == fresh.size(), k < fresh_lp.size() ? &fresh_lp[k] : nullptr);
        if (sb.toks.size() == 1 && !sb.echo.empty()) { piece = sb.echo + piece; sb.lp_chars = utf8_chars(sb.echo); }
        piece = json_escape(piece);
        std::string lpj = "null";
        if (pr.logprobs >= 0 && sb.lps.size() == sb.toks.size())
          lpj = chat ? lp_chat_json(s, sb.lps, sb.lps.size() - 1, pr.logprobs) : lp_completion_json(s, sb.lps, sb.lps.size() - 1, &sb.lp_chars);
        std::string c = "{\"id\":\"" + rid + "\",\"object\":\"" + obj + "\",\"created\":" + std::to_string(created) + ",\"model\":\"" + model +
                        "\",\"choices\":[{\"index\":" + std::to_string(i);
        if (chat) c += ",\"delta\":{\"role\":\"assistant\",\"content\":\"" + piece + "\"}";
        else c += ",\"text\":\"" + piece + "\"";
        c += ",\"token_ids\":[" + std::to_string(fresh[k]) + "],\"logprobs\":" + lpj + ",\"finish_reason\":" + (sb.fin ? reason(sb.fin) : "null") +
             ",\"stop_reason\":" + (sb.fin ? sb.stop_reason : std::string("null")) + "}]}";
        ok = send_chunk(fd, "data: " + c + "\n\n");
        ++completion;
      }
      sb.sent += fresh.size();
      if (sb.fin) { sb.done = true; --remaining; }
    }
  }
  if (!ok) {                                     // client went away: free the KV blocks
    for (auto& o : subs) if (!o.done) push_abort(s, o.id);
    return false;
  }
  if (pr.include_usage)
    ok = send_chunk(fd, "data: {\"id\":\"" + rid + "\",\"object\":\"" + obj + "\",\"created\":" + std::to_string(created) + ",\"model\":\"" + model +
                            "\",\"choices\":[],\"usage\":{\"prompt_tokens\":" + std::to_string(prompt_tokens) + ",\"total_tokens\":" +
                            std::to_string(prompt_tokens + completion) + ",\"completion_tokens\":" + std::to_string(completion) + "}}\n\n");
  ok = ok && send_chunk(fd, "data: [DONE]\n\n") && send_str(fd, "0\r\n\r\n");
  return ok;
}

std::string metrics_text(Server& s) {
  const std::string m = "model_name=\"" + s.args.served_model_name + "\"";
  std::string o;
  auto gauge = [&](const char* name, const char* help, double v) {
    o += std::string("# HELP ") + name + " " + help + "\n# TYPE " + name + " gauge\n" + name + "{" + m + "} " + std::to_string(v) + "\n";
  };
  auto counter = [&](const char* name, const char* help, const std::string& labels, double v, bool head) {
    if (head) o += std::string("# HELP ") + name + " " + help + "\n# TYPE " + name + " counter\n";
    o += std::string(name) + "{" + m + labels + "} " + std::to_string(v) + "\n";
  };
  gauge("vllm:num_requests_running", "Number of requests in model execution batches.", s.running);
  gauge("vllm:num_requests_waiting", "Number of requests waiting to be processed.", s.waiting);
  gauge("vllm:kv_cache_usage_perc", "KV-cache usage. 1 means 100 percent usage.", s.kv_usage);
  gauge("vllm:gpu_cache_usage_perc", "GPU KV-cache usage (v0.8 name). 1 means 100 percent usage.", s.kv_usage);
  counter("vllm:num_preemptions_total", "Cumulative number of preemption from the engine.", "", (double)s.preemptions, true);
  counter("vllm:prompt_tokens_total", "Number of prefill tokens processed.", "", (double)s.prompt_tokens, true);
  counter("vllm:generation_tokens_total", "Number of generation tokens processed.", "", (double)s.generation_tokens, true);
  counter("vllm:request_success_total", "Count of successfully processed requests.", ",finished_reason=\"stop\"", (double)s.success_stop, true);
  counter("vllm:request_success_total", "", ",finished_reason=\"length\"", (double)s.success_length, false);
  counter("vllm:prefix_cache_queries_total", "Prefix cache queries, in terms of number of queried tokens.", "", (double)s.pc_queries, true);
  counter("vllm:prefix_cache_hits_total", "Prefix cache hits, in terms of number of cached tokens.", "", (double)s.pc_hits, true);
  o += s.h_ttft.text("vllm:time_to_first_token_seconds", "Histogram of time to first token in seconds.", m);
  o += s.h_itl.text("vllm:inter_token_latency_seconds", "Histogram of inter-token latency in seconds.", m);
  o += s.h_itl.text("vllm:time_per_output_token_seconds", "Histogram of time per output token in seconds (v0.8 name of inter_token_latency).", m);
  o += s.h_e2e.text("vllm:e2e_request_latency_seconds", "Histogram of end to end request latency in seconds.", m);
  counter("kge:engine_steps_total", "Engine steps executed.", "", (double)s.num_steps, true);
  return o;
}

// POST /tokenize {"prompt": str, "add_special_tokens": true} | {"messages": [...]} -> {"count", "max_model_len", "tokens"}
// POST /detokenize {"tokens": [...]} -> {"prompt"}   (vllm:entrypoints/serve/tokenize/api_router.py, protocol.py)
bool handle_tokenize(Server& s, int fd, const HttpReq& hr, bool detok) {
  Json j;
  std::string perr;
  if (!json_parse(hr.body, &j, &perr) || !j.is_obj()) return respond_error(fd, 400, "invalid JSON body: " + perr, "BadRequestError", hr.keep_alive);
  if (!s.have_tok) return respond_error(fd, 400, kNoTokenizer, "BadRequestError", hr.keep_alive);
  if (detok) {
    const Json* t = j.get("tokens");
    if (!t || !t->is_arr()) return respond_error(fd, 400, "tokens must be an array of integers", "BadRequestError", hr.keep_alive);
    std::vector<int32_t> ids;
    for (auto& x : t->arr) {
      if (!x.is_num()) return respond_error(fd, 400, "tokens must be an array of integers", "BadRequestError", hr.keep_alive);
      ids.push_back((int32_t)std::max<int64_t>(-1, std::min<int64_t>(x.as_i64(), INT32_MAX)));
    }
    return respond(fd, 200, "application/json", "{\"prompt\":\"" + json_escape(s.tok.decode(ids)) + "\"}", hr.keep_alive);
  }
  std::vector<int32_t> ids;
  if (const Json* msgs = j.get("messages")) {
    if (!msgs->is_arr() || !s.tok.has_chat_template()) return respond_error(fd, 400, "messages need a supported chat template", "BadRequestError", hr.keep_alive);
    std::vector<ChatMessage> cm;
    for (auto& m : msgs->arr) if (m.is_obj()) cm.push_back({m.str_or("role", "user"), m.str_or("content", "")});
    ids = s.tok.encode(s.tok.render_chat(cm), false);
  } else {
    const Json* p = j.get("prompt");
    if (!p || !p->is_str()) return respond_error(fd, 400, "prompt must be a string", "BadRequestError", hr.keep_alive);
    ids = s.tok.encode(p->str, j.bool_or("add_special_tokens", true));
  }
  return respond(fd, 200, "application/json", "{\"count\":" + std::to_string(ids.size()) + ",\"max_model_len\":" + std::to_string(s.args.max_model_len) +
                 ",\"tokens\":" + ids_json(ids) + ",\"token_strs\":null}", hr.keep_alive);
}

void serve_conn(Server* sp, int fd) {
  Server& s = *sp;
  struct Live { std::atomic<int>& n; ~Live() { n.fetch_sub(1, std::memory_order_release); } } live{s.active_conns};   // counted in main before the thread starts
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
  std::string buf;
  while (!s.stop) {
    HttpReq r;
    const int rc = read_request(fd, buf, &r);
    if (rc <= 0) break;
    bool ok = true;
    // --api-key / VLLM_API_KEY guards the /v1 routes (vllm:entrypoints/openai/server_utils.py AuthenticationMiddleware)
    if (!s.args.api_key.empty() && r.path.compare(0, 3, "/v1") == 0 && r.method != "OPTIONS" && r.authorization != "Bearer " + s.args.api_key) {
      ok = respond(fd, 401, "application/json", "{\"error\":\"Unauthorized\"}", r.keep_alive);
      if (!ok || !r.keep_alive) break;
      continue;
    }
    if (r.path == "/health" || r.path == "/ping") {
      ok = respond(fd, s.healthy ? 200 : 503, "", "", r.keep_alive);
    } else if (r.path == "/v1/models" && r.method == "GET") {
      const std::string m = json_escape(s.args.served_model_name);
      ok = respond(fd, 200, "application/json",
                   "{\"object\":\"list\",\"data\":[{\"id\":\"" + m + "\",\"object\":\"model\",\"created\":" + std::to_string((long)s.created) +
                       ",\"owned_by\":\"vllm\",\"root\":\"" + json_escape(s.args.model) + "\",\"parent\":null,\"max_model_len\":" +
                       std::to_string(s.args.max_model_len) + "}]}",
                   r.keep_alive);
    } else if (r.path == "/metrics" && r.method == "GET") {
      ok = respond(fd, 200, "text/plain; version=0.0.4; charset=utf-8", metrics_text(s), r.keep_alive);
    } else if (r.path == "/version") {
      ok = respond(fd, 200, "application/json", "{\"version\":\"kge-b200-0.1\"}", r.keep_alive);
    } else if ((r.path == "/tokenize" || r.path == "/detokenize") && r.method == "POST") {
      ok = handle_tokenize(s, fd, r, r.path == "/detokenize");
    } else if (r.path == "/v1/completions" || r.path == "/v1/chat/completions") {
      if (r.method != "POST") ok = respond_error(fd, 405, "Method Not Allowed", "MethodNotAllowed", r.keep_alive);
      else if (!s.healthy) ok = respond_error(fd, 503, "engine is not healthy", "ServiceUnavailable", r.keep_alive);
      else ok = handle_generate(s, fd, r, r.path != "/v1/completions");
    } else {
      ok = respond_error(fd, 404, "Not Found", "NotFoundError", r.keep_alive);
    }
    if (!ok || !r.keep_alive) break;
  }
  ::close(fd);
}

void on_signal(int) { if (g_srv) g_srv->stop = true; }

bool is_dir(const std::string& p) {
  struct stat st;
  return stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}

// HF id -> local snapshot directory (no network in the pod's engine: the chart mounts or pre-pulls it)
std::string resolve_model(const std::string& model) {
  if (is_dir(model)) return model;
  const char* hf = getenv("HF_HOME");
  std::string base = hf ? std::string(hf) : std::string(getenv("HOME") ? getenv("HOME") : "/root") + "/.cache/huggingface";
  std::string name = model;
  for (auto& c : name) if (c == '/') c = '-';
  std::string repl = model;
  size_t p;
  while ((p = repl.find('/')) != std::string::npos) repl.replace(p, 1, "--");
  std::string snap = base + "/hub/models--" + repl + "/snapshots";
  if (is_dir(snap)) {
    // first snapshot
    std::string cmd = "ls -1 '" + snap + "' 2>/dev/null | head -1";
    if (FILE* f = popen(cmd.c_str(), "r")) {
      char line[512] = {0};
      if (fgets(line, sizeof line, f)) {
        std::string l = line;
        while (!l.empty() && (l.back() == '\n' || l.back() == '\r')) l.pop_back();
        pclose(f);
        if (!l.empty()) return snap + "/" + l;
      } else pclose(f);
    }
  }
  return "";
}

}  // namespace

// `vllm tokenize DIR`: one JSON object per stdin line, {"text": ...} -> {"ids": [...], "pieces": [...]},
// {"ids": [...]} -> {"text": ...}, {"messages": [...]} -> {"ids": [...], "text": rendered}.  No GPU.
int tokenize_main(const std::string& dir) {
  Tokenizer tok;
  std::string err;
  if (!tok.load(dir, &err)) { fprintf(stderr, "error: %s\n", err.c_str()); return 1; }
  std::string line;
  char buf[1 << 16];
  while (fgets(buf, sizeof buf, stdin)) {
    line = buf;
    while (!line.empty() && line.back() != '\n' && fgets(buf, sizeof buf, stdin)) line += buf;
    Json j;
    if (!json_parse(line, &j, &err) || !j.is_obj()) { printf("{\"error\":\"%s\"}\n", json_escape(err).c_str()); continue; }
    if (const Json* t = j.get("text")) {
      auto ids = tok.encode(t->str, j.bool_or("add_special_tokens", true));
      std::string pieces = "[";
      bool first = true;
      for (auto& p : tok.split(t->str)) { pieces += (first ? "\"" : ",\"") + json_escape(p) + "\""; first = false; }
      printf("{\"ids\":%s,\"pieces\":%s]}\n", ids_json(ids).c_str(), pieces.c_str());
    } else if (const Json* n = j.get("nfc")) {
      printf("{\"text\":\"%s\"}\n", json_escape(nfc_utf8(n->str)).c_str());
    } else if (j.get("ids") && j.get("stop")) {
      // {"ids", "stop": [str], "include_stop_str_in_output", "min_tokens", "stop_terminated"}: the request's text stream
      // -> {"text", "deltas": [text handed out after each token], "matched": stop index or -1, "tokens_used"}
      TextStream ts;
      for (auto& x : j.get("stop")->arr) ts.stop.push_back(x.str);
      ts.include_stop = j.bool_or("include_stop_str_in_output", false);
      ts.min_tokens = (int)j.int_or("min_tokens", 0);
      const bool stop_term = j.bool_or("stop_terminated", false);
      const auto& arr = j.get("ids")->arr;
      std::string deltas = "[";
      int used = 0;
      for (size_t i = 0; i < arr.size(); ++i) {
        const bool last = i + 1 == arr.size();
        const bool hit = ts.push(tok, (int32_t)arr[i].num, last && stop_term, last);
        ++used;
        deltas += (i ? ",\"" : "\"") + json_escape(ts.next_delta(hit || last)) + "\"";
        if (hit) break;
      }
      printf("{\"text\":\"%s\",\"deltas\":%s],\"matched\":%d,\"tokens_used\":%d}\n", json_escape(ts.text).c_str(), deltas.c_str(),
             ts.matched, used);
    } else if (const Json* ids = j.get("ids")) {
      std::vector<int32_t> v;
      for (auto& x : ids->arr) v.push_back((int32_t)std::max<int64_t>(-1, std::min<int64_t>(x.as_i64(), INT32_MAX)));
      std::string streamed, pending;
      for (int32_t id : v) streamed += tok.decode_stream(id, &pending);
      streamed += Tokenizer::flush_stream(&pending);
      printf("{\"text\":\"%s\",\"streamed\":\"%s\"}\n", json_escape(tok.decode(v)).c_str(), json_escape(streamed).c_str());
    } else if (const Json* ms = j.get("messages")) {
      std::vector<ChatMessage> cm;
      for (auto& m : ms->arr) cm.push_back({m.str_or("role", "user"), m.str_or("content", "")});
      const std::string r = tok.has_chat_template() ? tok.render_chat(cm) : "";
      printf("{\"text\":\"%s\",\"ids\":%s}\n", json_escape(r).c_str(), ids_json(tok.encode(r, false)).c_str());
    }
    fflush(stdout);
  }
  return 0;
}

int main(int argc, char** argv) {
  if (argc >= 3 && std::string(argv[1]) == "tokenize") return tokenize_main(argv[2]);
  Server s;
  std::string err = parse_serve_args(argc, argv, &s.args);
  if (!err.empty()) {
    fprintf(stderr, "usage: vllm serve MODEL [--host H] [--port P] [--max-model-len N] [--tensor-parallel-size N] ...\nerror: %s\n", err.c_str());
    return 2;
  }
  const ServeArgs& a = s.args;
  if (a.dry_run) {
    printf("{\"model\":\"%s\",\"served_model_name\":\"%s\",\"host\":\"%s\",\"port\":%d,\"max_model_len\":%d,\"dtype\":\"%s\","
           "\"tensor_parallel_size\":%d,\"pipeline_parallel_size\":%d,\"gpu_memory_utilization\":%.4f,\"max_num_seqs\":%d,"
           "\"max_num_batched_tokens\":%d,\"block_size\":%d,\"enable_chunked_prefill\":%s,\"enable_prefix_caching\":%s,"
           "\"enforce_eager\":%s,\"disable_custom_all_reduce\":%s,\"trust_remote_code\":%s,\"load_format\":\"%s\",\"seed\":%lld,"
           "\"generation_config\":\"%s\",\"generation_defaults\":%s}\n",
           json_escape(a.model).c_str(), json_escape(a.served_model_name).c_str(), a.host.c_str(), a.port, a.max_model_len, a.dtype.c_str(),
           a.tensor_parallel_size, a.pipeline_parallel_size, a.gpu_memory_utilization, a.max_num_seqs, a.max_num_batched_tokens,
           a.block_size, a.enable_chunked_prefill ? "true" : "false", a.enable_prefix_caching ? "true" : "false",
           a.enforce_eager ? "true" : "false", a.disable_custom_all_reduce ? "true" : "false", a.trust_remote_code ? "true" : "false",
           a.load_format.c_str(), (long long)a.seed, json_escape(a.generation_config).c_str(),
           load_gen_defaults(a.generation_config == "auto" || a.generation_config == "vllm" ? a.model : a.generation_config,
                             a.generation_config != "vllm").to_json().c_str());
    return 0;
  }
  if (!a.enable_prefix_caching)
    fprintf(stderr, "INFO: prefix caching is off (--no-enable-prefix-caching): every prompt is computed in full\n");
  if (a.dtype == "float32")
    fprintf(stderr, "WARNING: --dtype float32: this engine computes in 16-bit storage with fp32 accumulation; running in bfloat16\n");
  const int sms = kge_device_sm_count();
  if (sms <= 0) { fprintf(stderr, "error: no CUDA device visible (this engine has no CPU backend)\n"); return 1; }
  const std::string dir = resolve_model(a.model);
  if (dir.empty()) { fprintf(stderr, "error: model '%s' is neither a directory nor a cached HF snapshot under $HF_HOME\n", a.model.c_str()); return 1; }

  // engines: one per tensor-parallel rank on devices 0..tp-1 (the pod sees exactly requestGPU devices)
  s.engines.resize(a.tensor_parallel_size, nullptr);
  std::vector<int> rcs(a.tensor_parallel_size, 0);
  std::vector<std::string> errs(a.tensor_parallel_size);
  auto create = [&](int r) {
    kge_engine_config c{};
    c.max_num_seqs = a.max_num_seqs; c.max_num_batched_tokens = a.max_num_batched_tokens;
    c.max_model_len = a.max_model_len; c.block_size = a.block_size; c.num_blocks = a.num_gpu_blocks_override;
    c.dtype = (a.dtype == "float16" || a.dtype == "half") ? 2 : (a.dtype == "bfloat16" || a.dtype == "float32") ? 1 : 0;
    c.gpu_memory_utilization = (float)a.gpu_memory_utilization; c.tp_rank = r; c.tp_size = a.tensor_parallel_size;
    c.enable_chunked_prefill = a.enable_chunked_prefill; c.device = r;
    rcs[r] = kge_engine_create(&s.engines[r], dir.c_str(), &c, a.load_format == "dummy" ? a.seed : -1);
    if (rcs[r]) errs[r] = kge_last_error();
    else {
      kge_engine_set_reserve_full_isl(s.engines[r], a.scheduler_reserve_full_isl);
      kge_engine_set_prefix_caching(s.engines[r], a.enable_prefix_caching);
    }
  };
  {
    std::vector<std::thread> th;
    for (int r = 0; r < a.tensor_parallel_size; ++r) th.emplace_back(create, r);
    for (auto& t : th) t.join();
  }
  for (int r = 0; r < a.tensor_parallel_size; ++r)
    if (rcs[r]) { fprintf(stderr, "error: rank %d: %s\n", r, errs[r].c_str()); return 1; }
  if (a.tensor_parallel_size > 1 && !a.disable_custom_all_reduce) {
    // default exchange: the ranks (threads of this process) map each other's windows over NVLink peer access; the
    // forward pass then runs the fused all-reduce + residual + norm kernel, with decode graphs
    std::vector<kge_tp_peer_info> infos(a.tensor_parallel_size);
    for (int r = 0; r < a.tensor_parallel_size; ++r)
      if (kge_engine_tp_info(s.engines[r], &infos[r])) { fprintf(stderr, "error: rank %d: %s\n", r, kge_last_error()); return 1; }
    for (int r = 0; r < a.tensor_parallel_size; ++r)
      if (kge_engine_tp_connect(s.engines[r], infos.data())) { fprintf(stderr, "error: rank %d: %s\n", r, kge_last_error()); return 1; }
    fprintf(stderr, "INFO: tensor parallel exchange over peer memory (fused all-reduce + residual + norm)\n");
  } else if (a.tensor_parallel_size > 1) {
    // --disable-custom-all-reduce (ref:values-01-minimal-example8.yaml:32): NCCL all-reduce, as the reference deploys it
    char id[128];
    if (kge_nccl_unique_id(id)) { fprintf(stderr, "error: %s\n", kge_last_error()); return 1; }
    std::vector<std::thread> th;
    for (int r = 0; r < a.tensor_parallel_size; ++r)
      th.emplace_back([&, r] { rcs[r] = kge_engine_init_tp(s.engines[r], id); if (rcs[r]) errs[r] = kge_last_error(); });
    for (auto& t : th) t.join();
    for (int r = 0; r < a.tensor_parallel_size; ++r)
      if (rcs[r]) { fprintf(stderr, "error: rank %d: %s\n", r, errs[r].c_str()); return 1; }
    fprintf(stderr, "INFO: tensor parallel exchange over NCCL (--disable-custom-all-reduce)\n");
  }
  // sampling defaults and extra end-of-sequence ids the model ships (--generation-config auto | vllm | DIR)
  s.gen = load_gen_defaults(a.generation_config == "auto" || a.generation_config == "vllm" ? dir : a.generation_config, a.generation_config != "vllm");
  if (s.gen.has_temperature && s.gen.temperature < 0) s.gen.has_temperature = false;
  if (s.gen.has_top_p && !(s.gen.top_p > 0 && s.gen.top_p <= 1)) s.gen.has_top_p = false;
  if (s.gen.has_temperature || s.gen.has_top_p || s.gen.has_top_k || !s.gen.extra_eos.empty())
    fprintf(stderr, "INFO: model defaults applied to requests that leave them out: %s\n", s.gen.to_json().c_str());
  for (auto& k : s.gen.ignored) fprintf(stderr, "WARNING: generation_config.json sets %s, which this engine does not implement\n", k.c_str());
  s.vocab = kge_engine_vocab(s.engines[0]);
  s.eos = kge_engine_eos_token_id(s.engines[0]);
  {
    std::string terr;
    s.have_tok = s.tok.load(dir, &terr);
    if (s.have_tok) fprintf(stderr, "INFO: tokenizer loaded (%d entries%s)\n", s.tok.vocab_size(), s.tok.has_chat_template() ? ", chat template" : "");
    else fprintf(stderr, "INFO: %s - string prompts and chat messages will be refused; token-id prompts work\n", terr.c_str());
  }
  s.args.max_model_len = kge_engine_max_model_len(s.engines[0]);      // what the engine settled on when the flag was left out
  fprintf(stderr, "INFO: model %s loaded: %.2f GB weights/rank, %d KV blocks, tp=%d\n", dir.c_str(),
          kge_engine_weight_bytes(s.engines[0]) / 1e9, kge_engine_num_blocks(s.engines[0]), a.tensor_parallel_size);

  g_srv = &s;
  signal(SIGINT, on_signal);
  signal(SIGTERM, on_signal);
  signal(SIGPIPE, SIG_IGN);
  int lfd = socket(AF_INET, SOCK_STREAM, 0);
  int one = 1;
  setsockopt(lfd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons((uint16_t)a.port);
  if (inet_pton(AF_INET, a.host.c_str(), &addr.sin_addr) != 1) addr.sin_addr.s_addr = INADDR_ANY;
  if (bind(lfd, (sockaddr*)&addr, sizeof addr) || listen(lfd, 1024)) { perror("bind/listen"); return 1; }
  std::thread eng(engine_loop, &s);
  fprintf(stderr, "INFO: Uvicorn-compatible server running on http://%s:%d (Press CTRL+C to quit)\n", a.host.c_str(), a.port);
  while (!s.stop) {
    fd_set rf;
    FD_ZERO(&rf);
    FD_SET(lfd, &rf);
    timeval tv{0, 200000};
    if (select(lfd + 1, &rf, nullptr, nullptr, &tv) <= 0) continue;
    int fd = accept(lfd, nullptr, nullptr);
    if (fd < 0) continue;
    s.active_conns.fetch_add(1);
    std::thread(serve_conn, &s, fd).detach();
  }
  s.stop = true;
  { std::lock_guard<std::mutex> lk(s.mu); s.cv.notify_all(); }
  eng.join();
  ::close(lfd);
  // connection threads are detached: give them a moment to send their last bytes and leave before the state they use
  // goes away; whatever is still parked on an idle keep-alive socket after that ends with the process
  for (int i = 0; i < 300 && s.active_conns.load(std::memory_order_acquire) > 0; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(10));
  const int code = s.healthy ? 0 : 1;
  if (s.active_conns.load(std::memory_order_acquire) > 0) { fflush(nullptr); _exit(code); }
  for (auto* e : s.engines) kge_engine_destroy(e);
  return code;
}
