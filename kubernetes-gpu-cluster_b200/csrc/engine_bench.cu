// Measurement helpers of the engine: per-kernel timing read-out, host<->device byte counters, and
// the device-resident replay of one fixed-shape job (bench.py's `value` leg): all step metadata is
// uploaded before the timed region, the sampled ids of one step are the input ids of the next
// without leaving the device, and the stream is synchronised only at the end.
#include <string.h>

#include <algorithm>

#include "engine.h"

using namespace kge;

extern "C" int kge_engine_set_timing(kge_engine* e, int enable) {
  KGE_REQUIRE(e, "kge_engine_set_timing: null engine");
  e->timing = enable != 0;
  if (enable) memset(e->stats, 0, sizeof(e->stats));
  return 0;
}

extern "C" int kge_engine_get_timing(kge_engine* e, kge_kernel_stats* out) {
  KGE_REQUIRE(e && out, "kge_engine_get_timing: null argument");
  memcpy(out, e->stats, sizeof(e->stats));
  return 0;
}

extern "C" int kge_engine_io_bytes(const kge_engine* e, int64_t* h2d, int64_t* d2h) {
  KGE_REQUIRE(e, "kge_engine_io_bytes: null engine");
  if (h2d) *h2d = e->h2d_bytes;
  if (d2h) *d2h = e->d2h_bytes;
  return 0;
}

namespace {
struct ResidentStep {
  int T, B, R, max_q, max_seq;
  double sum_seq, attn_flops;
  int64_t off;        // byte offset of this step's metadata block in the device arena
  bool ids_from_prev; // decode: ids = previous step's sampled ids
  int out_row0;       // where the R sampled ids go in the [steps..] token log
};
}  // namespace

extern "C" int kge_engine_run_resident(kge_engine* e, const int32_t* prompts_host, int32_t B, int32_t prompt_len,
                                       int32_t out_len, int32_t prefill_batch, int32_t* out_tokens_host, float* ms,
                                       float* decode_ms) {
  KGE_REQUIRE(e && prompts_host && out_tokens_host && ms, "kge_engine_run_resident: null argument");
  KGE_REQUIRE(B > 0 && prompt_len > 0 && out_len > 0 && prefill_batch > 0, "kge_engine_run_resident: bad sizes");
  KGE_REQUIRE(!e->sched->has_unfinished(), "kge_engine_run_resident: the scheduler must be idle");
  e->sched->invalidate_prefix_cache();        // this replay writes KV blocks of its own choosing
  KGE_REQUIRE(B <= e->cfg.max_num_seqs, "kge_engine_run_resident: B=%d > max_num_seqs=%d", B, e->cfg.max_num_seqs);
  KGE_REQUIRE(prompt_len + out_len <= e->cfg.max_model_len, "kge_engine_run_resident: sequence exceeds max_model_len");
  KGE_REQUIRE(prefill_batch * prompt_len <= e->cfg.max_num_batched_tokens, "kge_engine_run_resident: prefill step exceeds max_num_batched_tokens");
  const int bs = e->cfg.block_size;
  const int blocks_per_seq = (prompt_len + out_len + bs - 1) / bs;
  KGE_REQUIRE((int64_t)B * blocks_per_seq + 1 <= e->num_blocks, "kge_engine_run_resident: needs %lld KV blocks, pool has %d",
              (long long)B * blocks_per_seq + 1, e->num_blocks);
  KGE_CUDA_OK(cudaSetDevice(e->cfg.device));
  const bool was_timing = e->timing;
  e->timing = false;
  e->step_num_decode = 0;                 // resident steps are pure prefill or pure decode
  e->step_num_masks = 0;

  // ---- build every step's metadata on the host (untimed)
  std::vector<ResidentStep> steps;
  std::vector<uint8_t> arena;
  auto block_of = [&](int b, int j) { return 1 + b * blocks_per_seq + j; };
  auto add_step = [&](int T, int Bs, int R, int max_blocks) -> StepMeta {
    int64_t bytes = 0;
    carve_meta(nullptr, T, Bs, R, max_blocks, &bytes);
    const int64_t off = (int64_t)arena.size();
    arena.resize(arena.size() + (size_t)((bytes + 255) / 256 * 256), 0);
    ResidentStep st{};
    st.T = T; st.B = Bs; st.R = R; st.off = off;
    steps.push_back(st);
    return carve_meta(arena.data() + off, T, Bs, R, max_blocks, nullptr);
  };
  int out_rows = 0;
  for (int b0 = 0; b0 < B; b0 += prefill_batch) {
    const int nb = std::min(prefill_batch, B - b0);
    const int mb = (prompt_len + bs - 1) / bs;
    StepMeta m = add_step(nb * prompt_len, nb, nb, mb);
    ResidentStep& st = steps.back();
    st.max_q = prompt_len; st.max_seq = prompt_len; st.ids_from_prev = false; st.out_row0 = out_rows;
    out_rows += nb;
    m.qsl[0] = 0;
    for (int i = 0; i < nb; ++i) {
      const int b = b0 + i;
      for (int t = 0; t < prompt_len; ++t) {
        m.ids[i * prompt_len + t] = prompts_host[(int64_t)b * prompt_len + t];
        m.pos[i * prompt_len + t] = t;
        m.slots[i * prompt_len + t] = (int64_t)block_of(b, t / bs) * bs + t % bs;
      }
      m.qsl[i + 1] = (i + 1) * prompt_len;
      m.seq_lens[i] = prompt_len;
      m.rows[i] = (i + 1) * prompt_len - 1;
      for (int j = 0; j < mb; ++j) m.btab[i * mb + j] = block_of(b, j);
      st.sum_seq += prompt_len;
      st.attn_flops += 4.0 * prompt_len * 0.5 * (prompt_len + 1);
    }
  }
  const int n_prefill_steps = (int)steps.size();
  for (int s = 1; s < out_len; ++s) {
    const int ctx = prompt_len + s;             // keys visible to the token decoded this step
    const int mb_used = (ctx + bs - 1) / bs;
    const int mb = std::min(e->max_blocks_per_seq, (mb_used + 31) / 32 * 32);   // bucketed stride (graph key)
    StepMeta m = add_step(B, B, B, mb);
    ResidentStep& st = steps.back();
    st.max_q = 1; st.max_seq = ctx; st.ids_from_prev = true; st.out_row0 = out_rows;
    out_rows += B;
    for (int b = 0; b < B; ++b) {
      m.pos[b] = ctx - 1;
      m.slots[b] = (int64_t)block_of(b, (ctx - 1) / bs) * bs + (ctx - 1) % bs;
      m.qsl[b] = b;
      m.seq_lens[b] = ctx;
      m.rows[b] = b;
      for (int j = 0; j < mb_used; ++j) m.btab[b * mb + j] = block_of(b, j);
      st.sum_seq += ctx;
      st.attn_flops += 4.0 * ctx;
    }
    m.qsl[B] = B;
  }
  uint8_t* d_arena = nullptr;
  int32_t* d_log = nullptr;       // every sampled id, step-major
  int32_t* d_first = nullptr;     // first generated token of each request (input of decode step 1)
  KGE_CUDA_OK(cudaMalloc((void**)&d_arena, arena.size()));
  KGE_CUDA_OK(cudaMalloc((void**)&d_log, (size_t)std::max(1, out_rows) * 4));
  KGE_CUDA_OK(cudaMalloc((void**)&d_first, (size_t)B * 4));
  KGE_CUDA_OK(cudaMemcpy(d_arena, arena.data(), arena.size(), cudaMemcpyHostToDevice));
  cudaEvent_t ev0, ev1, ev2;
  cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventCreate(&ev2);
  KGE_CUDA_OK(cudaStreamSynchronize(e->stream));

  // ---- timed region: nothing but kernel launches and device-to-device id hand-offs
  int rc = 0;
  cudaEventRecord(ev0, e->stream);
  for (size_t si = 0; si < steps.size() && rc == 0; ++si) {
    const ResidentStep& st = steps[si];
    if ((int)si == n_prefill_steps) cudaEventRecord(ev1, e->stream);
    if (st.ids_from_prev) {
      // decode: the step's metadata block is copied (device to device) to the engine's fixed slot so the
      // captured graph of this (B, splits, stride) can be replayed; the ids are the previous step's samples
      const int mb = std::min(e->max_blocks_per_seq, ((st.max_seq + bs - 1) / bs + 31) / 32 * 32);
      int64_t bytes = 0;
      StepMeta dm = carve_meta(e->meta_dev, st.T, st.B, st.R, mb, &bytes);
      cudaMemcpyAsync(e->meta_dev, d_arena + st.off, (size_t)bytes, cudaMemcpyDeviceToDevice, e->stream);
      cudaMemcpyAsync(dm.ids, ((int)si == n_prefill_steps) ? d_first : e->d_out, (size_t)st.B * 4, cudaMemcpyDeviceToDevice, e->stream);
      rc = engine_decode_step(e, st.B, mb, st.max_seq, dm, false, st.sum_seq);
    } else {
      const int mb = (prompt_len + bs - 1) / bs;
      StepMeta dm = carve_meta(d_arena + st.off, st.T, st.B, st.R, mb, nullptr);
      rc = engine_forward(e, st.T, st.B, st.R, mb, st.max_q, st.max_seq, dm, false, st.sum_seq, st.attn_flops);
    }
    if (rc) break;
    cudaMemcpyAsync(d_log + st.out_row0, e->d_out, (size_t)st.R * 4, cudaMemcpyDeviceToDevice, e->stream);
    if (!st.ids_from_prev)
      cudaMemcpyAsync(d_first + (int64_t)si * prefill_batch, e->d_out, (size_t)st.R * 4, cudaMemcpyDeviceToDevice, e->stream);
  }
  if (out_len == 1) cudaEventRecord(ev1, e->stream);
  cudaEventRecord(ev2, e->stream);
  cudaError_t serr = cudaStreamSynchronize(e->stream);
  e->timing = was_timing;
  if (rc == 0 && serr != cudaSuccess) { kge::set_error("kge_engine_run_resident: %s", cudaGetErrorString(serr)); rc = -2; }
  if (rc == 0) {
    float t_all = 0.f, t_dec = 0.f;
    cudaEventElapsedTime(&t_all, ev0, ev2);
    cudaEventElapsedTime(&t_dec, ev1, ev2);
    *ms = t_all;
    if (decode_ms) *decode_ms = t_dec;
    std::vector<int32_t> log((size_t)std::max(1, out_rows));
    cudaMemcpy(log.data(), d_log, (size_t)out_rows * 4, cudaMemcpyDeviceToHost);
    for (int b = 0; b < B; ++b) out_tokens_host[(int64_t)b * out_len] = log[b];        // prefill rows are in request order
    for (int s = 1; s < out_len; ++s)
      for (int b = 0; b < B; ++b) out_tokens_host[(int64_t)b * out_len + s] = log[(size_t)B + (size_t)(s - 1) * B + b];
  }
  cudaEventDestroy(ev0); cudaEventDestroy(ev1); cudaEventDestroy(ev2);
  cudaFree(d_arena); cudaFree(d_log); cudaFree(d_first);
  return rc;
}

// ---------------------------------------------------------------------------------------------------------------------
// Record the scheduler-driven steps of a job, then replay exactly those steps device-resident.
namespace {
__global__ void patch_ids_kernel(int32_t* ids, const int32_t* src, const int32_t* log, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < T && src[t] >= 0) ids[t] = log[src[t]];
}
}  // namespace

extern "C" int kge_engine_record_steps(kge_engine* e, int enable) {
  KGE_REQUIRE(e, "kge_engine_record_steps: null engine");
  if (enable) {
    KGE_REQUIRE(!e->sched->has_unfinished(), "kge_engine_record_steps: start recording with an idle scheduler");
    e->recorded.clear(); e->rec_samples.clear(); e->rec_sample_req.clear(); e->rec_pending_src.clear();
  }
  e->recording = enable != 0;
  return 0;
}

extern "C" int kge_engine_replay_recorded(kge_engine* e, int32_t* out_tokens_host, int64_t* out_request_ids_host, int64_t cap,
                                          int64_t* n_samples, int32_t* n_steps, float* ms, float* decode_ms) {
  KGE_REQUIRE(e && out_tokens_host && out_request_ids_host && n_samples && ms, "kge_engine_replay_recorded: null argument");
  KGE_REQUIRE(!e->recording && !e->recorded.empty(), "kge_engine_replay_recorded: record a job first (kge_engine_record_steps)");
  KGE_REQUIRE(!e->sched->has_unfinished(), "kge_engine_replay_recorded: the scheduler must be idle");
  const int64_t ns = (int64_t)e->rec_sample_req.size();
  KGE_REQUIRE(cap >= ns, "kge_engine_replay_recorded: capacity %lld < %lld samples", (long long)cap, (long long)ns);
  e->sched->invalidate_prefix_cache();          // the replay rewrites the KV blocks the recorded job used
  KGE_CUDA_OK(cudaSetDevice(e->cfg.device));
  const bool was_timing = e->timing;
  e->timing = false;
  e->step_num_masks = 0;                  // (a recorded job with min_tokens masks replays without them: bench jobs have none)
  // arena: per step [metadata block | src], 256-byte aligned
  std::vector<int64_t> off_meta, off_src;
  int64_t total = 0;
  for (auto& st : e->recorded) {
    off_meta.push_back(total); total += (st.bytes + 255) / 256 * 256;
    off_src.push_back(total); total += ((int64_t)st.T * 4 + 255) / 256 * 256;
  }
  std::vector<uint8_t> arena((size_t)total, 0);
  for (size_t i = 0; i < e->recorded.size(); ++i) {
    memcpy(arena.data() + off_meta[i], e->recorded[i].meta.data(), (size_t)e->recorded[i].bytes);
    memcpy(arena.data() + off_src[i], e->recorded[i].src.data(), (size_t)e->recorded[i].T * 4);
  }
  size_t first_all_decode = e->recorded.size();
  while (first_all_decode > 0 && e->recorded[first_all_decode - 1].decode_only) --first_all_decode;
  uint8_t* d_arena = nullptr;
  int32_t* d_log = nullptr;
  KGE_CUDA_OK(cudaMalloc((void**)&d_arena, arena.size()));
  KGE_CUDA_OK(cudaMalloc((void**)&d_log, (size_t)std::max<int64_t>(1, ns) * 4));
  KGE_CUDA_OK(cudaMemcpy(d_arena, arena.data(), arena.size(), cudaMemcpyHostToDevice));
  cudaEvent_t ev0, ev1, ev2;
  cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventCreate(&ev2);
  KGE_CUDA_OK(cudaStreamSynchronize(e->stream));
  // ---- timed region: device-to-device metadata moves, the id patch, and the step's kernels
  int rc = 0;
  cudaEventRecord(ev0, e->stream);
  for (size_t i = 0; i < e->recorded.size() && rc == 0; ++i) {
    const kge_engine::RecordedStep& st = e->recorded[i];
    if (i == first_all_decode) cudaEventRecord(ev1, e->stream);
    StepMeta dm = carve_meta(e->meta_dev, st.T, st.B, st.R, st.stride, nullptr);
    cudaMemcpyAsync(e->meta_dev, d_arena + off_meta[i], (size_t)st.bytes, cudaMemcpyDeviceToDevice, e->stream);
    patch_ids_kernel<<<(st.T + 255) / 256, 256, 0, e->stream>>>(dm.ids, (const int32_t*)(d_arena + off_src[i]), d_log, st.T);
    e->step_num_decode = st.num_decode; e->step_dec_max_seq = st.dec_max_seq; e->step_dec_sum_seq = st.dec_sum_seq;
    if (st.decode_only) rc = engine_decode_step(e, st.B, st.stride, st.max_seq, dm, st.any_random, st.sum_seq);
    else rc = engine_forward(e, st.T, st.B, st.R, st.stride, st.max_q, st.max_seq, dm, st.any_random, st.sum_seq, st.attn_flops);
    if (rc) break;
    if (st.R > 0) cudaMemcpyAsync(d_log + st.sample0, e->d_out, (size_t)st.R * 4, cudaMemcpyDeviceToDevice, e->stream);
  }
  if (first_all_decode >= e->recorded.size()) cudaEventRecord(ev1, e->stream);
  cudaEventRecord(ev2, e->stream);
  cudaError_t serr = cudaStreamSynchronize(e->stream);
  e->timing = was_timing;
  if (rc == 0 && serr != cudaSuccess) { kge::set_error("kge_engine_replay_recorded: %s", cudaGetErrorString(serr)); rc = -2; }
  if (rc == 0) {
    float t_all = 0.f, t_dec = 0.f;
    cudaEventElapsedTime(&t_all, ev0, ev2);
    cudaEventElapsedTime(&t_dec, ev1, ev2);
    *ms = t_all;
    if (decode_ms) *decode_ms = t_dec;
    cudaMemcpy(out_tokens_host, d_log, (size_t)ns * 4, cudaMemcpyDeviceToHost);
    memcpy(out_request_ids_host, e->rec_sample_req.data(), (size_t)ns * 8);
    *n_samples = ns;
    if (n_steps) *n_steps = (int32_t)e->recorded.size();
  }
  cudaEventDestroy(ev0); cudaEventDestroy(ev1); cudaEventDestroy(ev2);
  cudaFree(d_arena); cudaFree(d_log);
  return rc;
}
