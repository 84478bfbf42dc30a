// kge-router: the request router in front of N engine replicas (BASELINE config 5: `replicaCount` engines behind
// Service vllm-router-service:80, ref:values-01-minimal-example2.yaml:10, ref:old_README.md:1175).  It mirrors the
// command line and behaviour of the production-stack router the reference's chart deploys, for the static case:
//   kge-router --host 0.0.0.0 --port 8001 --service-discovery static
//              --static-backends http://10.0.0.1:8000,http://10.0.0.2:8000 [--static-models m1,m2]
//              --routing-logic roundrobin|session [--session-key x-user-id]
// POST /v1/completions and /v1/chat/completions go to one healthy backend that serves the request's `model`
// (round-robin, or the same backend for the same session header value) and the backend's response - JSON or an SSE
// stream - is relayed byte for byte as it arrives.  GET /v1/models merges the backends' lists, GET /health is 200 while
// at least one backend is healthy, GET /metrics reports per-backend request counts and health.  No GPU, no libkge.
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../csrc/json.h"

namespace {
using kge::Json;

struct Backend {
  std::string url, host, model;      // model: from --static-models ("" = serves anything)
  int port = 80;
  std::atomic<bool> healthy{false};
  std::atomic<int64_t> requests{0}, failures{0};
};

struct Router {
  std::string host = "0.0.0.0", logic = "roundrobin", session_key = "x-user-id";
  int port = 8001;
  std::vector<std::unique_ptr<Backend>> backends;
  std::atomic<uint64_t> rr{0};
  std::atomic<bool> stop{false};
  std::atomic<int64_t> no_backend{0};
  std::atomic<int> active_conns{0};                  // detached connection threads still running
};
Router* g_router = nullptr;
void on_signal(int) { if (g_router) g_router->stop = true; }

bool send_all(int fd, const char* p, size_t n) {
  while (n) {
    const ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
    if (w <= 0) return false;
    p += w; n -= (size_t)w;
  }
  return true;
}
bool send_str(int fd, const std::string& s) { return send_all(fd, s.data(), s.size()); }

std::string json_escape(const std::string& s) {
  std::string o;
  for (unsigned char c : s) {
    if (c == '"' || c == '\\') { o += '\\'; o += (char)c; }
    else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
    else o += (char)c;
  }
  return o;
}

bool respond(int fd, int code, const char* status, const std::string& ctype, const std::string& body) {
  std::string h = "HTTP/1.1 " + std::to_string(code) + " " + status + "\r\n";
  if (!ctype.empty()) h += "Content-Type: " + ctype + "\r\n";
  h += "Content-Length: " + std::to_string(body.size()) + "\r\nConnection: close\r\n\r\n";
  return send_str(fd, h + body);
}
bool respond_error(int fd, int code, const char* status, const std::string& msg, const char* type) {
  return respond(fd, code, status, "application/json",
                 "{\"error\":{\"message\":\"" + json_escape(msg) + "\",\"type\":\"" + type + "\",\"param\":null,\"code\":" + std::to_string(code) + "}}");
}

int connect_to(const std::string& host, int port, int timeout_ms) {
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET; hints.ai_socktype = SOCK_STREAM;
  if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) || !res) return -1;
  const int fd = socket(res->ai_family, res->ai_socktype, res->ai_protocol);
  if (fd < 0) { freeaddrinfo(res); return -1; }
  timeval tv{timeout_ms / 1000, (timeout_ms % 1000) * 1000};
  setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof tv);
  const int rc = connect(fd, res->ai_addr, res->ai_addrlen);
  freeaddrinfo(res);
  if (rc) { close(fd); return -1; }
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
  return fd;
}

// Small GET against a backend; the whole response body (Connection: close) or "" on failure.  *status = HTTP code.
std::string backend_get(const Backend& b, const std::string& path, int timeout_ms, int* status) {
  *status = 0;
  const int fd = connect_to(b.host, b.port, timeout_ms);
  if (fd < 0) return "";
  timeval tv{timeout_ms / 1000, (timeout_ms % 1000) * 1000};
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
  std::string resp;
  if (send_str(fd, "GET " + path + " HTTP/1.1\r\nHost: " + b.host + "\r\nConnection: close\r\n\r\n")) {
    char buf[8192];
    ssize_t n;
    while ((n = recv(fd, buf, sizeof buf, 0)) > 0) resp.append(buf, (size_t)n);
  }
  close(fd);
  if (resp.compare(0, 5, "HTTP/") == 0) {
    const size_t sp = resp.find(' ');
    if (sp != std::string::npos) *status = atoi(resp.c_str() + sp + 1);
  }
  const size_t he = resp.find("\r\n\r\n");
  return he == std::string::npos ? "" : resp.substr(he + 4);
}

void health_loop(Router* r) {
  while (!r->stop) {
    for (auto& b : r->backends) {
      int st = 0;
      backend_get(*b, "/health", 1000, &st);
      b->healthy = st == 200;
    }
    for (int i = 0; i < 20 && !r->stop; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(100));
  }
}

struct HttpReq {
  std::string method, path, head, body;
  std::string header(const std::string& name) const {       // lower-case name
    size_t pos = 0;
    while (pos < head.size()) {
      size_t e = head.find("\r\n", pos);
      if (e == std::string::npos) e = head.size();
      const std::string line = head.substr(pos, e - pos);
      pos = e + 2;
      const size_t c = line.find(':');
      if (c == std::string::npos) continue;
      std::string k = line.substr(0, c);
      for (auto& ch : k) ch = (char)tolower(ch);
      if (k == name) { size_t v = c + 1; while (v < line.size() && line[v] == ' ') ++v; return line.substr(v); }
    }
    return "";
  }
};

bool read_request(int fd, HttpReq* r) {
  std::string buf;
  size_t hend;
  char tmp[16384];
  while ((hend = buf.find("\r\n\r\n")) == std::string::npos) {
    if (buf.size() > (1 << 20)) return false;
    const ssize_t n = recv(fd, tmp, sizeof tmp, 0);
    if (n <= 0) return false;
    buf.append(tmp, (size_t)n);
  }
  r->head = buf.substr(0, hend);
  const size_t le = r->head.find("\r\n");
  const std::string line = r->head.substr(0, le);
  const size_t s1 = line.find(' '), s2 = line.rfind(' ');
  if (s1 == std::string::npos || s2 == s1) return false;
  r->method = line.substr(0, s1);
  r->path = line.substr(s1 + 1, s2 - s1 - 1);
  const size_t q = r->path.find('?');
  if (q != std::string::npos) r->path.resize(q);
  const size_t clen = (size_t)strtoull(r->header("content-length").c_str(), nullptr, 10);
  if (clen > (64u << 20)) return false;
  buf.erase(0, hend + 4);
  while (buf.size() < clen) {
    const ssize_t n = recv(fd, tmp, sizeof tmp, 0);
    if (n <= 0) return false;
    buf.append(tmp, (size_t)n);
  }
  r->body = buf.substr(0, clen);
  return true;
}

uint64_t fnv1a(const std::string& s) {
  uint64_t h = 1469598103934665603ull;
  for (unsigned char c : s) { h ^= c; h *= 1099511628211ull; }
  return h;
}

// Backends that are healthy and serve `model` ("" = any), in configuration order.
std::vector<Backend*> candidates(Router& r, const std::string& model) {
  std::vector<Backend*> v;
  for (auto& b : r.backends)
    if (b->healthy && (model.empty() || b->model.empty() || b->model == model)) v.push_back(b.get());
  return v;
}

// Round-robin, or (routing-logic session) the backend the session key hashes to among the current candidates;
// requests without the session header fall back to round-robin.
Backend* pick(Router& r, const std::vector<Backend*>& c, const std::string& session) {
  if (c.empty()) return nullptr;
  if (r.logic == "session" && !session.empty()) return c[fnv1a(session) % c.size()];
  return c[r.rr.fetch_add(1) % c.size()];
}

// Forward the request and relay the backend's response verbatim as it arrives (JSON body or SSE stream).
// Returns false when nothing could be relayed (so the caller may try another backend).
bool proxy(Backend& b, const HttpReq& req, int client_fd, bool* client_gone) {
  const int fd = connect_to(b.host, b.port, 2000);
  if (fd < 0) return false;
  std::string h = req.method + " " + req.path + " HTTP/1.1\r\nHost: " + b.host + ":" + std::to_string(b.port) + "\r\n";
  const std::string ct = req.header("content-type"), auth = req.header("authorization");
  if (!ct.empty()) h += "Content-Type: " + ct + "\r\n";
  if (!auth.empty()) h += "Authorization: " + auth + "\r\n";
  h += "Content-Length: " + std::to_string(req.body.size()) + "\r\nConnection: close\r\n\r\n";
  if (!send_str(fd, h + req.body)) { close(fd); return false; }
  bool any = false;
  char buf[16384];
  ssize_t n;
  while ((n = recv(fd, buf, sizeof buf, 0)) > 0) {
    any = true;
    if (!send_all(client_fd, buf, (size_t)n)) { *client_gone = true; break; }   // closing `fd` makes the engine abort the request
  }
  close(fd);
  return any;
}

std::string models_json(Router& r) {
  std::string data;
  std::vector<std::string> seen;
  for (auto& b : r.backends) {
    if (!b->healthy) continue;
    int st = 0;
    const std::string body = backend_get(*b, "/v1/models", 2000, &st);
    Json j;
    std::string err;
    if (st != 200 || !kge::json_parse(body, &j, &err)) continue;
    const Json* d = j.get("data");
    if (!d || !d->is_arr()) continue;
    for (auto& m : d->arr) {
      const std::string id = m.str_or("id", "");
      bool dup = false;
      for (auto& s : seen) dup |= s == id;
      if (id.empty() || dup) continue;
      seen.push_back(id);
      data += std::string(data.empty() ? "" : ",") + "{\"id\":\"" + json_escape(id) + "\",\"object\":\"model\",\"created\":" +
              std::to_string((long long)m.int_or("created", 0)) + ",\"owned_by\":\"" + json_escape(m.str_or("owned_by", "vllm")) + "\",\"root\":\"" +
              json_escape(m.str_or("root", id)) + "\",\"parent\":null,\"max_model_len\":" + std::to_string((long long)m.int_or("max_model_len", 0)) + "}";
    }
  }
  return "{\"object\":\"list\",\"data\":[" + data + "]}";
}

std::string metrics_text(Router& r) {
  std::string o = "# HELP vllm_router:healthy_pods_total Number of healthy engine backends.\n# TYPE vllm_router:healthy_pods_total gauge\n";
  int healthy = 0;
  for (auto& b : r.backends) healthy += b->healthy ? 1 : 0;
  o += "vllm_router:healthy_pods_total " + std::to_string(healthy) + "\n";
  o += "# HELP vllm_router:requests_total Requests routed to a backend.\n# TYPE vllm_router:requests_total counter\n";
  for (auto& b : r.backends) o += "vllm_router:requests_total{server=\"" + b->url + "\"} " + std::to_string((long long)b->requests) + "\n";
  o += "# HELP vllm_router:request_failures_total Requests a backend could not be reached for.\n# TYPE vllm_router:request_failures_total counter\n";
  for (auto& b : r.backends) o += "vllm_router:request_failures_total{server=\"" + b->url + "\"} " + std::to_string((long long)b->failures) + "\n";
  o += "# HELP vllm_router:no_backend_total Requests rejected because no healthy backend serves the model.\n# TYPE vllm_router:no_backend_total counter\n";
  o += "vllm_router:no_backend_total " + std::to_string((long long)r.no_backend) + "\n";
  return o;
}

void serve_conn(Router* rp, int fd) {
  Router& r = *rp;
  struct Live { std::atomic<int>& n; ~Live() { n.fetch_sub(1, std::memory_order_release); } } live{r.active_conns};
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
  HttpReq req;
  if (read_request(fd, &req)) {
    if (req.path == "/health" || req.path == "/ping") {
      bool any = false;
      for (auto& b : r.backends) any |= b->healthy.load();
      respond(fd, any ? 200 : 503, any ? "OK" : "Service Unavailable", "", "");
    } else if (req.path == "/v1/models" && req.method == "GET") {
      respond(fd, 200, "OK", "application/json", models_json(r));
    } else if (req.path == "/metrics" && req.method == "GET") {
      respond(fd, 200, "OK", "text/plain; version=0.0.4; charset=utf-8", metrics_text(r));
    } else if (req.method == "POST" && (req.path == "/v1/completions" || req.path == "/v1/chat/completions")) {
      Json j;
      std::string err, model;
      if (kge::json_parse(req.body, &j, &err) && j.is_obj()) model = j.str_or("model", "");
      const std::string session = req.header(r.session_key);
      std::vector<Backend*> c = candidates(r, model);
      bool done = false, client_gone = false;
      // a backend that cannot be reached is marked unhealthy and the next candidate is tried
      while (!c.empty() && !done && !client_gone) {
        Backend* b = pick(r, c, session);
        ++b->requests;
        done = proxy(*b, req, fd, &client_gone);
        if (!done && !client_gone) {
          ++b->failures;
          b->healthy = false;
          std::vector<Backend*> rest;
          for (Backend* x : c) if (x != b) rest.push_back(x);
          c.swap(rest);
        }
      }
      if (!done && !client_gone) {
        ++r.no_backend;
        const bool known = model.empty() || [&] { for (auto& b : r.backends) if (b->model.empty() || b->model == model) return true; return false; }();
        if (!known) respond_error(fd, 404, "Not Found", "The model `" + model + "` does not exist.", "NotFoundError");
        else respond_error(fd, 503, "Service Unavailable", "no healthy engine backend for model `" + model + "`", "ServiceUnavailableError");
      }
    } else {
      respond_error(fd, 404, "Not Found", "unknown route " + req.path, "NotFoundError");
    }
  }
  shutdown(fd, SHUT_RDWR);
  close(fd);
}

std::vector<std::string> split_csv(const std::string& s) {
  std::vector<std::string> v;
  size_t i = 0;
  while (i <= s.size()) {
    size_t e = s.find(',', i);
    if (e == std::string::npos) e = s.size();
    if (e > i) v.push_back(s.substr(i, e - i));
    i = e + 1;
  }
  return v;
}

}  // namespace

int main(int argc, char** argv) {
  Router r;
  g_router = &r;
  std::vector<std::string> urls, models;
  std::string discovery = "static";
  for (int i = 1; i < argc; ++i) {
    std::string k = argv[i], v;
    const size_t eq = k.find('=');
    if (eq != std::string::npos) { v = k.substr(eq + 1); k.resize(eq); }
    else if (i + 1 < argc && argv[i + 1][0] != '-') v = argv[++i];
    if (k == "--host") r.host = v;
    else if (k == "--port") r.port = atoi(v.c_str());
    else if (k == "--service-discovery") discovery = v;
    else if (k == "--static-backends") urls = split_csv(v);
    else if (k == "--static-models") models = split_csv(v);
    else if (k == "--routing-logic") r.logic = v;
    else if (k == "--session-key") { r.session_key = v; for (auto& c : r.session_key) c = (char)tolower(c); }
    else if (k == "--help" || k == "-h") { printf("kge-router --port P --service-discovery static --static-backends URL[,URL] [--static-models M[,M]] [--routing-logic roundrobin|session] [--session-key HEADER]\n"); return 0; }
    // other production-stack flags (--engine-stats-interval, --log-stats, ...) are accepted and ignored
  }
  if (discovery != "static") { fprintf(stderr, "error: --service-discovery %s is not supported (static only: there is no Kubernetes API here)\n", discovery.c_str()); return 2; }
  if (urls.empty()) { fprintf(stderr, "error: --static-backends is required\n"); return 2; }
  if (!models.empty() && models.size() != urls.size()) { fprintf(stderr, "error: --static-models must name one model per backend\n"); return 2; }
  if (r.logic != "roundrobin" && r.logic != "session") { fprintf(stderr, "error: --routing-logic must be roundrobin or session\n"); return 2; }
  for (size_t i = 0; i < urls.size(); ++i) {
    auto b = std::make_unique<Backend>();
    b->url = urls[i];
    std::string u = urls[i];
    if (u.compare(0, 7, "http://") == 0) u = u.substr(7);
    while (!u.empty() && u.back() == '/') u.pop_back();
    const size_t c = u.rfind(':');
    b->host = c == std::string::npos ? u : u.substr(0, c);
    b->port = c == std::string::npos ? 80 : atoi(u.c_str() + c + 1);
    if (!models.empty()) b->model = models[i];
    r.backends.push_back(std::move(b));
  }
  signal(SIGINT, on_signal);
  signal(SIGTERM, on_signal);
  signal(SIGPIPE, SIG_IGN);
  const int lfd = socket(AF_INET, SOCK_STREAM, 0);
  int one = 1;
  setsockopt(lfd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons((uint16_t)r.port);
  addr.sin_addr.s_addr = r.host == "0.0.0.0" ? INADDR_ANY : inet_addr(r.host.c_str());
  if (bind(lfd, (sockaddr*)&addr, sizeof addr) || listen(lfd, 1024)) { perror("bind/listen"); return 1; }
  for (auto& b : r.backends) { int st = 0; backend_get(*b, "/health", 1000, &st); b->healthy = st == 200; }
  std::thread hc(health_loop, &r);
  fprintf(stderr, "INFO: kge-router on %s:%d, %zu backends, routing %s\n", r.host.c_str(), r.port, r.backends.size(), r.logic.c_str());
  timeval tv{0, 200000};
  setsockopt(lfd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
  while (!r.stop) {
    const int fd = accept(lfd, nullptr, nullptr);
    if (fd < 0) continue;
    r.active_conns.fetch_add(1);
    std::thread(serve_conn, &r, fd).detach();
  }
  close(lfd);
  hc.join();
  // let the detached connection threads finish before the state they use goes away
  for (int i = 0; i < 300 && r.active_conns.load(std::memory_order_acquire) > 0; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(10));
  if (r.active_conns.load(std::memory_order_acquire) > 0) { fflush(nullptr); _exit(0); }
  return 0;
}
