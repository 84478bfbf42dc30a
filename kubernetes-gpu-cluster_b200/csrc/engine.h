// Engine object behind the kge_engine handle: device weights, activation arena, paged KV pool,
// per-step metadata staging, scheduler.  One engine = one GPU (one TP rank).
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "kge.h"
#include "loader.h"
#include "scheduler.h"

namespace kge {

// The element-type dependent ops behind one table: the engine is compiled once and runs a bfloat16 or a float16 model
// by calling the C entry points of the matching build of the kernel sources (common.cuh).
struct Ops {
  decltype(&kge_embed) embed;
  decltype(&kge_rmsnorm) rmsnorm;
  decltype(&kge_add_rmsnorm) add_rmsnorm;
  decltype(&kge_head_rmsnorm) head_rmsnorm;
  decltype(&kge_layernorm) layernorm;
  decltype(&kge_silu_mul) silu_mul;
  decltype(&kge_gemm_bf16) gemm;
  decltype(&kge_gemm_silu_gate_ws_bf16) gemm_silu_gate;
  decltype(&kge_rope_kvwrite) rope_kvwrite;
  decltype(&kge_rope_cache_host) rope_cache_host;
  decltype(&kge_paged_attn) paged_attn;
  decltype(&kge_argmax) argmax;
  decltype(&kge_sample) sample;
  decltype(&kge_logprobs) logprobs;
  decltype(&kge_tp_reduce_add_norm) tp_reduce_add_norm;
};
extern const Ops kOpsBf16, kOpsF16;     // engine.cu

struct Linear {
  bf16* w = nullptr;     // [N, K] row-major
  bf16* b = nullptr;     // [N] or null
  int N = 0, K = 0;
};

struct LayerW {
  Linear qkv, o, up, down;           // up = gate_up (llama, [2f,d]) or fc1 (opt)
  bf16 *ln1_w = nullptr, *ln1_b = nullptr, *ln2_w = nullptr, *ln2_b = nullptr;
  bf16 *q_norm = nullptr, *k_norm = nullptr;   // qwen3: [hd] each
};

// Per-step metadata, laid out once in pinned host memory and mirrored on the device with the
// same offsets so one cudaMemcpyAsync moves everything.
struct StepMeta {
  int32_t *ids, *pos, *qsl, *seq_lens, *rows, *btab, *top_k;
  int64_t* slots;
  float *temp, *top_p;
  uint64_t* seeds;
};

}  // namespace kge

struct kge_engine {
  kge::ModelSpec spec;
  kge_engine_config cfg;
  int tp_rank = 0, tp_size = 1;
  bool f16 = false;                          // float16 model (--dtype float16); bfloat16 otherwise
  const kge::Ops* op = &kge::kOpsBf16;
  ncclComm_t comm = nullptr;
  // peer-memory exchange (tp_peer.cu): created with the engine when tp_size > 1; h, xn and the gathered logits live in
  // its window at the offsets below (the same on every rank)
  kge_tp_comm* tpc = nullptr;
  int64_t off_h = 0, off_xn = 0, off_logits = 0;
  kge::bf16* logits_loc = nullptr;           // [max_num_seqs, vocab_shard]: this rank's columns of the logits
  kge::bf16* logits_gather = nullptr;        // NCCL path: [tp, max_num_seqs, vocab_shard] before the reorder
  int vocab_shard = 0, vocab_loc = 0, vocab_col0 = 0;   // shard width (multiple of 8), columns really held, first column
  bool tp_peer() const { return tpc && tp_size > 1 && kge_tp_comm_connected(tpc); }
  int H_loc = 0, KV_loc = 0, ffn_loc = 0;
  cudaStream_t stream = nullptr;

  std::vector<void*> allocs;                 // everything cudaMalloc'ed, freed in destroy
  kge::bf16 *embed = nullptr, *pos_embed = nullptr, *lm_head = nullptr;
  kge::bf16 *final_w = nullptr, *final_b = nullptr, *cos_sin = nullptr;
  std::vector<kge::LayerW> layers;
  std::vector<kge::bf16*> kc, vc;
  int num_blocks = 0, max_blocks_per_seq = 0;
  int64_t weight_bytes = 0, kv_bytes = 0;

  kge::bf16 *h = nullptr, *res = nullptr, *xn = nullptr, *qkv = nullptr, *attn = nullptr;
  kge::bf16 *gu = nullptr, *act = nullptr, *xsel = nullptr, *logits = nullptr;
  float* logits_f32 = nullptr;
  int32_t* d_out = nullptr;
  int ld_logits = 0;
  void *ws_gemm = nullptr, *ws_attn = nullptr;
  int64_t ws_gemm_bytes = 0, ws_attn_bytes = 0;

  uint8_t *meta_host = nullptr, *meta_dev = nullptr;
  int64_t meta_bytes = 0;
  int32_t* out_host = nullptr;               // pinned [max_num_seqs]
  // log-probabilities of the last step's sampled rows (only when a request asked): [max_num_seqs, 21] + ranks
  float* lp_vals = nullptr; int32_t* lp_ids = nullptr; int32_t* lp_rank = nullptr;        // device
  float* lp_vals_host = nullptr; int32_t* lp_ids_host = nullptr; int32_t* lp_rank_host = nullptr;   // pinned
  int lp_last_n = -1;                                          // top-N the last step computed, -1 = none
  std::unordered_map<int64_t, std::pair<int, int>> lp_last;    // request id -> (row of the last step, its own N)

  // CUDA graphs of pure decode steps, keyed by what is baked into the launches
  struct DecodeGraph { int B, ns, stride; bool rnd; cudaGraphExec_t exec; int64_t launches; };
  std::vector<DecodeGraph> graphs;
  std::vector<std::pair<int64_t, int>> graph_seen;   // key hash -> eager runs so far
  bool use_graphs = true;

  // min_tokens (vllm:v1/sample/logits_processor/builtin.py MinTokensLogitsProcessor): (logits row, token id) pairs whose
  // logit is set to -inf before sampling in this step - EOS and the stop ids of requests still short of min_tokens
  int32_t* mask_pairs = nullptr;             // device [2 * kMaxMaskPairs]
  int32_t* mask_pairs_host = nullptr;        // pinned
  int step_num_masks = 0;
  static constexpr int kMaxMaskPairs = 4096;

  // mixed steps: the scheduler lists running (decode) requests first; the forward pass sends that
  // prefix to the decode attention kernel and the rest to the prefill kernel
  int step_num_decode = 0, step_dec_max_seq = 1;
  double step_dec_sum_seq = 0;

  // Recording of the scheduler-driven steps of a job (kge_engine_record_steps) for a device-resident replay
  // (kge_engine_replay_recorded): each step's packed metadata as it was shipped, plus, per token row, where its id comes
  // from in a replay - a sampled token of an earlier step (index into the job's sample log) or the recorded prompt id (-1).
  struct RecordedStep {
    int T, B, R, stride, max_q, max_seq, num_decode, dec_max_seq;
    double sum_seq, attn_flops, dec_sum_seq;
    bool decode_only, any_random;
    int64_t bytes;
    int sample0;                               // index of this step's first sampled row in the job's sample log
    std::vector<uint8_t> meta;
    std::vector<int32_t> src;                  // [T]
  };
  bool recording = false;
  std::vector<RecordedStep> recorded;
  std::vector<int32_t> rec_pending_src;                               // handed from kge_engine_step to run_step
  std::unordered_map<int64_t, std::vector<int32_t>> rec_samples;      // request id -> sample-log indices of its outputs
  std::vector<int64_t> rec_sample_req;                                // sample-log index -> request id
  std::unique_ptr<kge::Scheduler> sched;
  int64_t launches = 0;
  int64_t h2d_bytes = 0, d2h_bytes = 0;      // metadata in / sampled ids out, summed over steps
  // optional per-launch CUDA-event timing (kge_engine_set_timing): category totals
  bool timing = false;
  std::vector<cudaEvent_t> ev_pool;
  std::vector<int> ev_cat;                   // category of each recorded (start, stop) pair
  std::vector<double> ev_work_bytes, ev_work_flops;
  kge_kernel_stats stats[KGE_NUM_KERNEL_CATS] = {};
  uint64_t step_count = 0;
  std::string err;
};

namespace kge {
// engine_fwd.cu
int engine_forward(kge_engine* e, int T, int B, int R, int max_blocks, int max_q_len, int max_seq_len,
                   const StepMeta& dm, bool any_random, double sum_seq, double attn_flops);
void collect_timing(kge_engine* e);
inline namespace KGE_ELT_NS { int attn_decode_splits(int B, int KV, int max_seq_len); }   // attn.cu (the same in both element builds)
// A pure decode step (one token per request, metadata at the fixed-stride layout `dm` in meta_dev):
// eager the first time a (B, splits, stride, sampling) key is seen, captured into a CUDA graph the
// second time, replayed afterwards.
int engine_decode_step(kge_engine* e, int B, int stride, int max_seq_len, const StepMeta& dm, bool any_random, double sum_seq);
StepMeta carve_meta(uint8_t* base, int T, int B, int R, int max_blocks, int64_t* total);
}  // namespace kge
