#include "scheduler.h"

#include <algorithm>

namespace kge {

Scheduler::Scheduler(const SchedConfig& c) : cfg_(c) {
  free_.reserve(c.num_blocks > 0 ? c.num_blocks : 0);
  for (int32_t b = c.num_blocks - 1; b >= 1; --b) free_.push_back(b);   // pop() hands out 1, 2, ...
}

Request* Scheduler::find(int64_t id) {
  auto it = reqs_.find(id);
  return it == reqs_.end() ? nullptr : it->second.get();
}

bool Scheduler::add(int64_t id, const int32_t* prompt, int32_t n, const SamplingParams& sp, const char** why) {
  static const char* kEmpty = "prompt is empty";
  static const char* kLong = "prompt is longer than max_model_len";
  static const char* kDup = "duplicate request id";
  static const char* kKv = "prompt needs more KV blocks than the pool holds";
  if (n <= 0) { if (why) *why = kEmpty; return false; }
  if (n >= cfg_.max_model_len) { if (why) *why = kLong; return false; }
  if (reqs_.count(id)) { if (why) *why = kDup; return false; }
  if ((n + cfg_.block_size - 1) / cfg_.block_size > usable_blocks()) { if (why) *why = kKv; return false; }
  auto r = std::make_unique<Request>();
  r->id = id;
  r->prompt.assign(prompt, prompt + n);
  r->sp = sp;
  waiting_.push_back(r.get());
  reqs_[id] = std::move(r);
  return true;
}

bool Scheduler::alloc(Request* r, int32_t n_new) {
  const int32_t bs = cfg_.block_size;
  const int32_t need = (r->num_computed + n_new + bs - 1) / bs - (int32_t)r->blocks.size();
  if (need > (int32_t)free_.size()) return false;
  for (int32_t i = 0; i < need; ++i) { r->blocks.push_back(free_.back()); free_.pop_back(); }
  return true;
}

void Scheduler::release(Request* r) {
  for (auto it = r->blocks.rbegin(); it != r->blocks.rend(); ++it) free_.push_back(*it);
  r->blocks.clear();
}

void Scheduler::finish(Request* r, ReqStatus st) {
  r->status = st;
  release(r);
  running_.erase(std::remove(running_.begin(), running_.end(), r), running_.end());
  waiting_.erase(std::remove(waiting_.begin(), waiting_.end(), r), waiting_.end());
  reqs_.erase(r->id);
}

bool Scheduler::abort(int64_t id) {
  Request* r = find(id);
  if (!r) return false;
  finish(r, ReqStatus::Aborted);
  return true;
}

const std::vector<Scheduled>& Scheduler::schedule() {
  sched_.clear();
  int32_t budget = cfg_.max_num_batched_tokens;
  bool preempted = false;
  size_t i = 0;
  while (i < running_.size() && budget > 0) {
    Request* r = running_[i];
    int32_t n = std::min({r->num_tokens() - r->num_computed, budget, cfg_.max_model_len - 1 - r->num_computed});
    if (n <= 0) { ++i; continue; }
    bool ok = true;
    while (!alloc(r, n)) {
      Request* victim = running_.back();
      running_.pop_back();
      release(victim);
      victim->num_computed = 0;
      victim->status = ReqStatus::Preempted;
      victim->num_preemptions++;
      ++preemptions_;
      waiting_.push_front(victim);
      preempted = true;
      for (size_t k = 0; k < sched_.size(); ++k)
        if (sched_[k].req == victim) { budget += sched_[k].num_tokens; sched_.erase(sched_.begin() + k); break; }
      if (victim == r) { ok = false; break; }
    }
    if (!ok) break;
    sched_.push_back({r, n});
    budget -= n;
    ++i;
  }
  if (!preempted) {
    while (!waiting_.empty() && budget > 0) {
      if ((int32_t)running_.size() == cfg_.max_num_seqs) break;
      Request* r = waiting_.front();
      int32_t n = r->num_tokens() - r->num_computed;
      if (!cfg_.enable_chunked_prefill && n > budget) break;
      n = std::min(n, budget);
      if (cfg_.reserve_full_isl) {
        const int32_t full = std::min(r->num_tokens(), cfg_.max_model_len);
        if ((full + cfg_.block_size - 1) / cfg_.block_size - (int32_t)r->blocks.size() > free_blocks()) break;
      }
      if (!alloc(r, n)) break;
      waiting_.pop_front();
      r->status = ReqStatus::Running;
      running_.push_back(r);
      sched_.push_back({r, n});
      budget -= n;
    }
  }
  for (auto& s : sched_) s.req->num_computed += s.num_tokens;
  return sched_;
}

void Scheduler::update(const std::vector<int32_t>& sampled, std::vector<StepResult>* out) {
  size_t si = 0;
  // copy: finish() edits running_ and frees requests
  std::vector<Scheduled> cur = sched_;
  sched_.clear();
  for (auto& s : cur) {
    Request* r = s.req;
    StepResult res{r->id, -1, 0, s.num_tokens};
    if (r->num_computed == r->num_tokens()) {
      const int32_t tok = si < sampled.size() ? sampled[si] : 0;
      ++si;
      r->output.push_back(tok);
      res.token = tok;
      const int32_t nout = (int32_t)r->output.size();
      if (nout >= r->sp.min_tokens) {       // utils.py:100-101: nothing stops a request earlier
        if (!r->sp.ignore_eos && tok == cfg_.eos_token_id) res.finished = 1;
        else if (std::find(r->sp.stop_token_ids.begin(), r->sp.stop_token_ids.end(), tok) != r->sp.stop_token_ids.end())
          res.finished = 1;
        else if (r->num_tokens() >= cfg_.max_model_len || nout >= r->sp.max_tokens) res.finished = 2;
      }
    }
    if (out) out->push_back(res);
    if (res.finished) finish(r, res.finished == 1 ? ReqStatus::FinishedStop : ReqStatus::FinishedLength);
  }
}

}  // namespace kge
