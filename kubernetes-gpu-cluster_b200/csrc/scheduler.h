// Continuous-batching scheduler and paged-KV block manager (host side, no CUDA).
// Policy follows vllm:v1/core/sched/scheduler.py:329-830 (FCFS: RUNNING first under the token
// budget, tail preemption with recompute when blocks run out, WAITING admission up to
// max_num_seqs), :1283 update_from_output, stop rules vllm:v1/core/sched/utils.py:94-130 and the
// block accounting of vllm:v1/core/kv_cache_manager.py:236-430 (block 0 is the null block).
// Automatic prefix caching (--enable-prefix-caching): block pool with vLLM's free-queue order and hash map
// (vllm:v1/core/block_pool.py:184-433), longest-hit lookup and the "hits that are eviction candidates count as
// allocations" rule (vllm:v1/core/single_type_kv_cache_manager.py:88-240,483-528), at most num_tokens-1 cached
// tokens per admission (vllm:v1/core/kv_cache_manager.py:194-235).
#pragma once
#include <stdint.h>

#include <deque>
#include <memory>
#include <unordered_map>
#include <vector>

namespace kge {

struct SamplingParams {
  int32_t max_tokens = 16;
  int32_t min_tokens = 0;
  bool ignore_eos = false;
  float temperature = 1.0f;
  int32_t top_k = 0;
  float top_p = 1.0f;
  uint64_t seed = 0;
  bool has_seed = false;
  std::vector<int32_t> stop_token_ids;
  int32_t logprobs = -1;               // top-N log-probabilities per sampled token (0 = the sampled token's only), -1 = none
};

// 128-bit chained hash of a full block: (parent block's hash, the block's token ids).
struct BlockHash {
  uint64_t a = 0, b = 0;
  bool operator==(const BlockHash& o) const { return a == o.a && b == o.b; }
};
struct BlockHashHasher { size_t operator()(const BlockHash& h) const { return (size_t)(h.a ^ (h.b * 0x9E3779B97F4A7C15ull)); } };

enum class ReqStatus { Waiting, Running, Preempted, FinishedStop, FinishedLength, Aborted };

struct Request {
  int64_t id = 0;
  std::vector<int32_t> prompt;
  std::vector<int32_t> output;
  SamplingParams sp;
  int32_t num_computed = 0;
  std::vector<int32_t> blocks;
  std::vector<BlockHash> block_hashes;   // of the leading full blocks of prompt + output, extended on demand
  int32_t num_hashed_blocks = 0;         // leading entries of `blocks` that are registered in the prefix cache
  int32_t num_cached_tokens = 0;         // tokens taken from the prefix cache at the last admission
  ReqStatus status = ReqStatus::Waiting;
  int32_t num_preemptions = 0;
  int32_t num_tokens() const { return (int32_t)(prompt.size() + output.size()); }
  int32_t token_at(int32_t i) const {
    return i < (int32_t)prompt.size() ? prompt[i] : output[i - prompt.size()];
  }
};

struct SchedConfig {
  int32_t max_num_seqs = 1024;
  int32_t max_num_batched_tokens = 8192;
  int32_t max_model_len = 4096;
  int32_t block_size = 16;
  int32_t num_blocks = 0;              // including null block 0
  bool enable_chunked_prefill = true;
  int32_t eos_token_id = -1;
  // vllm --scheduler-reserve-full-isl (default on): a WAITING request is admitted only when the pool has
  // free blocks for its whole known sequence, not just its first chunk (vllm:v1/core/kv_cache_manager.py:346-360).
  bool reserve_full_isl = true;
  // vllm --enable-prefix-caching: full blocks stay addressable by the hash of their token prefix after their request
  // is gone, and a new request whose prompt starts with the same tokens reuses them instead of recomputing.
  bool enable_prefix_caching = false;
};

// Fixed pool of KV blocks: free queue (doubly linked, O(1) removal from the middle), reference counts, and the
// prefix-cache index hash -> blocks (insertion order; the first one serves lookups).  Block 0 is the null block.
class BlockPool {
 public:
  void init(int32_t num_blocks);
  int32_t num_free() const { return num_free_; }
  int32_t ref(int32_t b) const { return ref_[b]; }
  int32_t cached_block(const BlockHash& h) const;          // -1 = miss
  void touch(const std::vector<int32_t>& blocks);          // a hit: +1 reference, leaves the free queue if it was there
  int32_t take();                                          // pops the FRONT of the free queue, evicting its hash
  void release(const std::vector<int32_t>& blocks_in_eviction_order);
  void set_hash(int32_t b, const BlockHash& h);
  int64_t num_cached_hashes() const { return (int64_t)index_.size(); }
  void forget_hashes();                                    // drop every cache entry (block contents were overwritten behind the pool's back)

 private:
  void unlink(int32_t b);
  void push_back(int32_t b);
  int32_t num_free_ = 0;
  std::vector<int32_t> prev_, next_, ref_;                 // links use block 0 as the list head
  std::vector<uint8_t> in_free_, has_hash_;
  std::vector<BlockHash> hash_;
  std::unordered_map<BlockHash, std::vector<int32_t>, BlockHashHasher> index_;
};

struct Scheduled {
  Request* req;
  int32_t num_tokens;                  // tokens run this step, starting at req->num_computed - num_tokens
};

struct StepResult {
  int64_t request_id;
  int32_t token;                       // -1: prefill chunk only
  int32_t finished;                    // 0 running, 1 stop, 2 length
  int32_t num_scheduled;
};

class Scheduler {
 public:
  explicit Scheduler(const SchedConfig& c);
  // Returns false (with reason) when the request can never run.
  bool add(int64_t id, const int32_t* prompt, int32_t n, const SamplingParams& sp, const char** why);
  bool abort(int64_t id);
  // One scheduling pass; advances num_computed of every scheduled request.
  const std::vector<Scheduled>& schedule();
  // `tokens[i]` is the sampled token for the i-th scheduled request that reached its last known
  // token this step (needs_sample order).  Appends one StepResult per scheduled request.
  void update(const std::vector<int32_t>& sampled, std::vector<StepResult>* out);
  static bool needs_sample(const Scheduled& s) { return s.req->num_computed == s.req->num_tokens(); }
  bool has_unfinished() const { return !running_.empty() || !waiting_.empty(); }
  int32_t num_running() const { return (int32_t)running_.size(); }
  int32_t num_waiting() const { return (int32_t)waiting_.size(); }
  int32_t free_blocks() const { return pool_.num_free(); }
  int32_t usable_blocks() const { return cfg_.num_blocks - 1; }
  int64_t num_preemptions() const { return preemptions_; }
  const SchedConfig& config() const { return cfg_; }
  void set_reserve_full_isl(bool on) { cfg_.reserve_full_isl = on; }
  void set_prefix_caching(bool on) { cfg_.enable_prefix_caching = on; }
  // KV blocks were written without going through schedule() (kge_step with caller-built tables, the resident replay):
  // nothing cached can be trusted any more.  Requests that are running keep their blocks but register no more hashes.
  void invalidate_prefix_cache();
  int64_t prefix_cache_queries() const { return pc_queries_; }   // tokens looked up / tokens served from the cache
  int64_t prefix_cache_hits() const { return pc_hits_; }
  Request* find(int64_t id);

 private:
  bool alloc(Request* r, int32_t n_new);
  int32_t admit(Request* r, int32_t budget);            // tokens scheduled, or -1: the request has to wait
  void release(Request* r);
  void finish(Request* r, ReqStatus st);
  void extend_hashes(Request* r, int32_t nblocks);
  void register_full_blocks(Request* r, int32_t num_tokens);
  SchedConfig cfg_;
  BlockPool pool_;
  int64_t pc_queries_ = 0, pc_hits_ = 0;
  std::deque<Request*> waiting_;
  std::vector<Request*> running_;
  std::unordered_map<int64_t, std::unique_ptr<Request>> reqs_;
  std::vector<Scheduled> sched_;
  int64_t preemptions_ = 0;
};

}  // namespace kge
