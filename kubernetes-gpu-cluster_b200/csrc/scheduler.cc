#include "scheduler.h"

#include <algorithm>

namespace kge {

// ------------------------------------------------------------------------------- block pool
void BlockPool::init(int32_t num_blocks) {
  const int32_t n = num_blocks > 1 ? num_blocks : 1;
  prev_.assign(n, 0); next_.assign(n, 0); ref_.assign(n, 0);
  in_free_.assign(n, 0); has_hash_.assign(n, 0);
  hash_.assign(n, BlockHash());
  index_.clear();
  num_free_ = 0;
  prev_[0] = next_[0] = 0;                                   // empty ring around the null block
  for (int32_t b = 1; b < n; ++b) push_back(b);              // the queue hands out 1, 2, ...
}

void BlockPool::push_back(int32_t b) {
  const int32_t tail = prev_[0];
  next_[tail] = b; prev_[b] = tail; next_[b] = 0; prev_[0] = b;
  in_free_[b] = 1;
  ++num_free_;
}

void BlockPool::unlink(int32_t b) {
  next_[prev_[b]] = next_[b];
  prev_[next_[b]] = prev_[b];
  in_free_[b] = 0;
  --num_free_;
}

int32_t BlockPool::cached_block(const BlockHash& h) const {
  auto it = index_.find(h);
  return it == index_.end() || it->second.empty() ? -1 : it->second.front();
}

void BlockPool::touch(const std::vector<int32_t>& blocks) {
  for (int32_t b : blocks) {
    if (ref_[b] == 0 && in_free_[b]) unlink(b);              // an eviction candidate becomes live again
    ++ref_[b];
  }
}

int32_t BlockPool::take() {
  const int32_t b = next_[0];
  if (b == 0) return -1;
  unlink(b);
  if (has_hash_[b]) {                                        // reusing the block evicts its cache entry
    auto it = index_.find(hash_[b]);
    if (it != index_.end()) {
      auto& v = it->second;
      v.erase(std::remove(v.begin(), v.end(), b), v.end());
      if (v.empty()) index_.erase(it);
    }
    has_hash_[b] = 0;
  }
  ref_[b] = 1;
  return b;
}

void BlockPool::release(const std::vector<int32_t>& order) {
  for (int32_t b : order) --ref_[b];
  for (int32_t b : order)
    if (ref_[b] == 0 && !in_free_[b]) push_back(b);          // hashes stay: the blocks remain cache hits until reused
}

void BlockPool::forget_hashes() {
  index_.clear();
  std::fill(has_hash_.begin(), has_hash_.end(), (uint8_t)0);
}

void BlockPool::set_hash(int32_t b, const BlockHash& h) {
  hash_[b] = h;
  has_hash_[b] = 1;
  index_[h].push_back(b);
}

// ------------------------------------------------------------------------------- scheduler
Scheduler::Scheduler(const SchedConfig& c) : cfg_(c) { pool_.init(c.num_blocks); }

static inline uint64_t mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// hashes of the first `nblocks` full blocks of prompt + output (chained: a block's hash covers its whole prefix)
void Scheduler::extend_hashes(Request* r, int32_t nblocks) {
  const int32_t bs = cfg_.block_size;
  while ((int32_t)r->block_hashes.size() < nblocks) {
    const int32_t i = (int32_t)r->block_hashes.size();
    BlockHash h = i ? r->block_hashes[i - 1] : BlockHash{0x6B67655F62323030ull, 0x70726566697863ull};
    for (int32_t t = i * bs; t < (i + 1) * bs; ++t) {
      const uint64_t tok = (uint64_t)(uint32_t)r->token_at(t);
      h.a = mix64(h.a ^ (tok + 0x9E3779B97F4A7C15ull));
      h.b = mix64(h.b + tok * 0xD6E8FEB86659FD93ull + 0x2545F4914F6CDD1Dull);
    }
    h.a = mix64(h.a ^ (uint64_t)i);                          // position in the chain
    r->block_hashes.push_back(h);
  }
}

// blocks that became full get their hash (vllm cache_blocks: up to min(computed + new, num_tokens) tokens)
void Scheduler::register_full_blocks(Request* r, int32_t num_tokens) {
  if (!cfg_.enable_prefix_caching) return;
  const int32_t nfull = num_tokens / cfg_.block_size;
  if (r->num_hashed_blocks >= nfull) return;
  extend_hashes(r, nfull);
  for (int32_t i = r->num_hashed_blocks; i < nfull; ++i) pool_.set_hash(r->blocks[i], r->block_hashes[i]);
  r->num_hashed_blocks = nfull;
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

void Scheduler::invalidate_prefix_cache() {
  if (!cfg_.enable_prefix_caching) return;
  pool_.forget_hashes();
  for (auto& kv : reqs_) kv.second->num_hashed_blocks = (int32_t)kv.second->blocks.size();   // nothing of theirs gets registered later
}

bool Scheduler::alloc(Request* r, int32_t n_new) {          // a RUNNING request grows
  const int32_t bs = cfg_.block_size;
  const int32_t need = (r->num_computed + n_new + bs - 1) / bs - (int32_t)r->blocks.size();
  if (need > pool_.num_free()) return false;
  for (int32_t i = 0; i < need; ++i) r->blocks.push_back(pool_.take());
  register_full_blocks(r, std::min(r->num_computed + n_new, r->num_tokens()));
  return true;
}

// WAITING -> RUNNING (vllm scheduler.py:560-760).  Looks the prompt up in the prefix cache when nothing of the request
// is computed (new or preempted), applies the full-sequence admission gate and allocates the first chunk.
int32_t Scheduler::admit(Request* r, int32_t budget) {
  const int32_t bs = cfg_.block_size;
  std::vector<int32_t> hits;
  if (cfg_.enable_prefix_caching && r->num_computed == 0) {
    const int32_t max_blocks = (r->num_tokens() - 1) / bs;   // the last token is always recomputed: its logits are needed
    extend_hashes(r, max_blocks);
    for (int32_t i = 0; i < max_blocks; ++i) {
      const int32_t b = pool_.cached_block(r->block_hashes[i]);
      if (b < 0) break;
      hits.push_back(b);
    }
  }
  const int32_t nh = (int32_t)hits.size();
  const int32_t computed = r->num_computed + nh * bs;
  int32_t n = r->num_tokens() - computed;
  if (!cfg_.enable_chunked_prefill && n > budget) return -1;
  n = std::min(n, budget);
  int32_t evictable = 0;                                     // hit blocks sitting in the free queue: touching them shrinks it
  for (int32_t b : hits) evictable += pool_.ref(b) == 0;
  const int32_t have = (int32_t)r->blocks.size() + nh;
  if (cfg_.reserve_full_isl) {
    const int32_t full = (std::min(r->num_tokens(), cfg_.max_model_len) + bs - 1) / bs;
    if (std::max(full - have, 0) + evictable > pool_.num_free()) return -1;
  }
  const int32_t need = std::max((computed + n + bs - 1) / bs - have, 0);
  if (need + evictable > pool_.num_free()) return -1;
  if (nh) {
    pool_.touch(hits);
    r->blocks = hits;
    r->num_hashed_blocks = nh;
    r->num_computed = computed;
  }
  if (cfg_.enable_prefix_caching && r->num_computed == nh * bs) {   // a lookup happened
    r->num_cached_tokens = nh * bs;
    pc_queries_ += r->num_tokens();
    pc_hits_ += nh * bs;
  }
  for (int32_t i = 0; i < need; ++i) r->blocks.push_back(pool_.take());
  register_full_blocks(r, std::min(computed + n, r->num_tokens()));
  return n;
}

void Scheduler::release(Request* r) {
  // tail first, so that a request's last blocks are evicted before the blocks a shared prefix lives in
  std::vector<int32_t> order(r->blocks.rbegin(), r->blocks.rend());
  pool_.release(order);
  r->blocks.clear();
  r->num_hashed_blocks = 0;
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
      const int32_t n = admit(r, budget);
      if (n < 0) break;
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
