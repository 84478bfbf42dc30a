// One engine step on the device: embedding -> L x (norm, QKV, RoPE+KV write, paged attention,
// O, norm, MLP) -> final norm -> logits of the requested rows -> sampling.  Plus the C entry
// points kge_step (metadata given by the caller) and kge_engine_step (scheduler-driven).
#include <string.h>

#include <algorithm>

#include <cuda_fp16.h>

#include "engine.h"

using namespace kge;

namespace kge {

static void timer_begin(kge_engine* e, int cat, double bytes, double flops) {
  if (!e->timing) return;
  const size_t idx = e->ev_cat.size() * 2;
  while (e->ev_pool.size() < idx + 2) {
    cudaEvent_t ev;
    cudaEventCreate(&ev);
    e->ev_pool.push_back(ev);
  }
  e->ev_cat.push_back(cat);
  e->ev_work_bytes.push_back(bytes);
  e->ev_work_flops.push_back(flops);
  cudaEventRecord(e->ev_pool[idx], e->stream);
}
static void timer_end(kge_engine* e) {
  if (!e->timing) return;
  cudaEventRecord(e->ev_pool[e->ev_cat.size() * 2 - 1], e->stream);
}
// after the stream has been synchronised
void collect_timing(kge_engine* e) {
  for (size_t i = 0; i < e->ev_cat.size(); ++i) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e->ev_pool[2 * i], e->ev_pool[2 * i + 1]) != cudaSuccess) continue;
    kge_kernel_stats& st = e->stats[e->ev_cat[i]];
    st.launches += 1; st.ms += ms; st.bytes += e->ev_work_bytes[i]; st.flops += e->ev_work_flops[i];
  }
  e->ev_cat.clear(); e->ev_work_bytes.clear(); e->ev_work_flops.clear();
}

#define FWD(cat, bytes, flops, call)       \
  do {                                     \
    timer_begin(e, cat, bytes, flops);     \
    int _rc = (call);                      \
    if (_rc) return _rc;                   \
    timer_end(e);                          \
    ++e->launches;                         \
  } while (0)

enum { C_GEMM_S = 0, C_GEMM_L = 1, C_ATT_DEC = 2, C_ATT_GEN = 3, C_NORM = 4, C_ROPE = 5, C_ACT = 6, C_EMBED = 7, C_SAMPLE = 8, C_AR = 9 };

static int linear_rc(kge_engine* e, bf16* y, const bf16* x, const Linear& l, int T, int act) {
  return e->op->gemm(y, x, l.w, l.b, T, l.N, l.K, act, e->ws_gemm, e->ws_gemm_bytes, e->stream);
}
#define LINEAR(y, x, l, T, act)                                                                   \
  FWD((T) <= 256 ? C_GEMM_S : C_GEMM_L, 2.0 * ((double)(l).N * (l).K + (double)(T) * ((l).K + (l).N)), \
      2.0 * (T) * (double)(l).N * (l).K, linear_rc(e, y, x, l, T, act))

__global__ void mask_logits_kernel(bf16* logits, const int32_t* pairs, int n, int ld, int vocab, bool f16);
int tp_all_reduce(kge_engine* e, bf16* buf, int64_t n);   // tp.cu
int tp_all_gather_logits(kge_engine* e, int R);           // tp.cu (NCCL path)

// The row-parallel projection before this point left its output (under TP: this rank's partial sum) in e->h:
// all-reduce it, add it to the residual stream and normalise into e->xn.  Single GPU: the fused add + norm kernel.
// TP over peer memory: ONE kernel does all three (tp_peer.cu) and keeps the residual for this rank's rows only.
// TP over NCCL (--disable-custom-all-reduce): ncclAllReduce, then the norm kernel.
static int reduce_add_norm(kge_engine* e, int T, const bf16* w, const bf16* b, bool layer_norm, float eps) {
  const int d = e->spec.d;
  if (e->tp_peer())
    return e->op->tp_reduce_add_norm(e->tpc, e->off_xn, e->res, e->off_h, w, b, T, d, eps, layer_norm ? 3 : 1, e->stream);
  if (layer_norm) return e->op->layernorm(e->xn, e->res, e->h, w, b, T, d, eps, e->stream);
  return e->op->add_rmsnorm(e->xn, e->res, e->h, w, T, d, eps, e->stream);
}

int engine_forward(kge_engine* e, int T, int B, int R, int max_blocks, int max_q_len, int max_seq_len,
                   const StepMeta& m, bool any_random, double sum_seq, double attn_flops) {
  const ModelSpec& s = e->spec;
  const int d = s.d, hd = s.hd, Hl = e->H_loc, KVl = e->KV_loc, bs = e->cfg.block_size;
  cudaStream_t st = e->stream;
  const float scale = 1.0f / sqrtf((float)hd);
  const int qkv_n = e->layers[0].qkv.N;
  const double Td2 = (double)T * d * 2;
  const int c_att = max_q_len == 1 ? C_ATT_DEC : C_ATT_GEN;
  const double att_bytes = sum_seq * KVl * hd * 4.0 + (double)T * Hl * hd * 4.0;
  const double att_fl = attn_flops * Hl * hd;
  const int D = (max_q_len > 1 && e->step_num_decode > 0 && e->step_num_decode < B) ? e->step_num_decode : 0;
  const double dec_bytes = e->step_dec_sum_seq * KVl * hd * 4.0 + (double)D * Hl * hd * 4.0;
  // attention of layer i: one launch, or (mixed step) decode kernel for the leading D one-token requests
  // + prefill kernel for the others; query_start_loc entries are absolute rows, so sub-batches only
  // offset the per-request arrays
  auto attend = [&](int i) -> int {
    if (D == 0) {
      FWD(c_att, att_bytes, att_fl, e->op->paged_attn(e->attn, e->qkv, qkv_n, e->kc[i], e->vc[i], m.btab, max_blocks, m.seq_lens, m.qsl, B, T,
                         max_q_len, max_seq_len, Hl, KVl, hd, bs, scale, e->ws_attn, e->ws_attn_bytes, st));
      return 0;
    }
    FWD(C_ATT_DEC, dec_bytes, 4.0 * e->step_dec_sum_seq * Hl * hd,
        e->op->paged_attn(e->attn, e->qkv, qkv_n, e->kc[i], e->vc[i], m.btab, max_blocks, m.seq_lens, m.qsl, D, D, 1,
                       e->step_dec_max_seq, Hl, KVl, hd, bs, scale, e->ws_attn, e->ws_attn_bytes, st));
    FWD(C_ATT_GEN, att_bytes - dec_bytes, att_fl - 4.0 * e->step_dec_sum_seq * Hl * hd,
        e->op->paged_attn(e->attn, e->qkv, qkv_n, e->kc[i], e->vc[i], m.btab + (int64_t)D * max_blocks, max_blocks, m.seq_lens + D,
                       m.qsl + D, B - D, T - D, max_q_len, max_seq_len, Hl, KVl, hd, bs, scale, e->ws_attn, e->ws_attn_bytes, st));
    return 0;
  };
  const double rope_bytes = (double)T * (Hl + 2 * KVl) * hd * 2 * 2;
  const double ar_bytes = Td2;
  const bool peer = e->tp_peer();
  // all-reduce (TP) + residual add + norm; the fused peer kernel is booked under all_reduce
#define REDUCE_NORM(w, b, ln, eps)                                                                      \
  do {                                                                                                  \
    if (e->tp_size > 1 && !peer) FWD(C_AR, ar_bytes, 0, tp_all_reduce(e, e->h, (int64_t)T * d));         \
    FWD(peer ? C_AR : C_NORM, 4 * Td2, 0, reduce_add_norm(e, T, w, b, ln, eps));                        \
  } while (0)
  if (s.is_opt()) {
    FWD(C_EMBED, 2 * Td2, 0, e->op->embed(e->res, m.ids, e->embed, e->pos_embed, m.pos, 2, T, d, st));
    for (int i = 0; i < s.L; ++i) {
      const LayerW& L = e->layers[i];
      // first LN of the layer; for i > 0 it also folds in the previous layer's fc2 output
      if (i == 0) FWD(C_NORM, 2 * Td2, 0, e->op->layernorm(e->xn, e->res, nullptr, L.ln1_w, L.ln1_b, T, d, 1e-5f, st));
      else REDUCE_NORM(L.ln1_w, L.ln1_b, true, 1e-5f);
      LINEAR(e->qkv, e->xn, L.qkv, T, 0);
      FWD(C_ROPE, rope_bytes, 0, e->op->rope_kvwrite(e->qkv, nullptr, nullptr, m.slots, e->kc[i], e->vc[i], T, Hl, KVl, hd, bs, st));
      if (int rc = attend(i)) return rc;
      LINEAR(e->h, e->attn, L.o, T, 0);
      REDUCE_NORM(L.ln2_w, L.ln2_b, true, 1e-5f);
      LINEAR(e->act, e->xn, L.up, T, 1);
      LINEAR(e->h, e->act, L.down, T, 0);
    }
    REDUCE_NORM(e->final_w, e->final_b, true, 1e-5f);
  } else {
    FWD(C_EMBED, 2 * Td2, 0, e->op->embed(e->res, m.ids, e->embed, nullptr, nullptr, 0, T, d, st));
    for (int i = 0; i < s.L; ++i) {
      const LayerW& L = e->layers[i];
      if (i == 0) FWD(C_NORM, 2 * Td2, 0, e->op->rmsnorm(e->xn, e->res, L.ln1_w, T, d, s.eps, st));
      else REDUCE_NORM(L.ln1_w, nullptr, false, s.eps);
      LINEAR(e->qkv, e->xn, L.qkv, T, 0);
      if (s.qk_norm)
        FWD(C_NORM, (double)T * (Hl + KVl) * hd * 4, 0, e->op->head_rmsnorm(e->qkv, L.q_norm, L.k_norm, T, Hl, KVl, hd, s.eps, st));
      FWD(C_ROPE, rope_bytes, 0, e->op->rope_kvwrite(e->qkv, e->cos_sin, m.pos, m.slots, e->kc[i], e->vc[i], T, Hl, KVl, hd, bs, st));
      if (int rc = attend(i)) return rc;
      LINEAR(e->h, e->attn, L.o, T, 0);
      REDUCE_NORM(L.ln2_w, nullptr, false, s.eps);
      if (e->ffn_loc % 64 == 0) {
        // gate_up GEMM with the SiLU*mul in its epilogue: [T,2f] never reaches HBM
        FWD(T <= 256 ? C_GEMM_S : C_GEMM_L, 2.0 * ((double)L.up.N * L.up.K + (double)T * (L.up.K + e->ffn_loc)),
            2.0 * T * (double)L.up.N * L.up.K, e->op->gemm_silu_gate(e->act, e->xn, L.up.w, T, e->ffn_loc, L.up.K, e->ws_gemm, e->ws_gemm_bytes, st));
      } else {
        LINEAR(e->gu, e->xn, L.up, T, 0);
        FWD(C_ACT, (double)T * e->ffn_loc * 2 * 3, 0, e->op->silu_mul(e->act, e->gu, T, e->ffn_loc, st));
      }
      LINEAR(e->h, e->act, L.down, T, 0);
    }
    REDUCE_NORM(e->final_w, nullptr, false, s.eps);
  }
#undef REDUCE_NORM
  if (R > 0) {
    // logits only for the last token of each sampled request (logits_processor.py:54-99)
    FWD(C_EMBED, 2.0 * R * d * 2, 0, e->op->embed(e->xsel, m.rows, e->xn, nullptr, nullptr, 0, R, d, st));
    Linear head;
    head.w = e->lm_head; head.b = nullptr; head.N = e->vocab_loc; head.K = d;
    if (e->tp_size == 1) {
      LINEAR(e->logits, e->xsel, head, R, 0);
    } else {
      // this rank's vocabulary columns, then the all-gather of vllm:model_executor/layers/logits_processor.py:83
      LINEAR(e->logits_loc, e->xsel, head, R, 0);
      if (peer)
        FWD(C_AR, 2.0 * R * s.vocab, 0, kge_tp_all_gather_cols(e->tpc, e->off_logits, e->logits_loc, R, e->vocab_loc, s.vocab, e->vocab_col0, st));
      else
        FWD(C_AR, 2.0 * R * s.vocab, 0, tp_all_gather_logits(e, R));
    }
    if (e->step_num_masks > 0) {
      mask_logits_kernel<<<(e->step_num_masks + 127) / 128, 128, 0, st>>>(e->logits, e->mask_pairs, e->step_num_masks, s.vocab, s.vocab, e->f16);
      KGE_LAUNCH_OK();
      ++e->launches;
      pdl_break_next();
    }
    if (any_random)
      FWD(C_SAMPLE, (double)R * s.vocab * 2, 0, e->op->sample(e->d_out, e->logits, R, s.vocab, s.vocab, m.temp, m.top_k, m.top_p, nullptr, m.seeds,
                     0 /* the step is folded into the per-row seeds */, nullptr, 0, st));
    else
      FWD(C_SAMPLE, (double)R * s.vocab * 2, 0, e->op->argmax(e->d_out, e->logits, R, s.vocab, s.vocab, st));
  }
  return 0;
}

int engine_decode_step(kge_engine* e, int B, int stride, int max_seq_len, const StepMeta& dm, bool any_random, double sum_seq) {
  const double attn_flops = 4.0 * sum_seq;
  const int ns = attn_decode_splits(B, e->KV_loc, max_seq_len);
  if (!e->use_graphs || e->timing || e->step_num_masks > 0)      // a masked step is rare and its mask list is not part of a graph
    return engine_forward(e, B, B, B, stride, 1, max_seq_len, dm, any_random, sum_seq, attn_flops);
  for (auto& g : e->graphs)
    if (g.B == B && g.ns == ns && g.stride == stride && g.rnd == any_random) {
      KGE_CUDA_OK(cudaGraphLaunch(g.exec, e->stream));
      e->launches += g.launches;
      return 0;
    }
  const int64_t key = ((int64_t)B << 40) ^ ((int64_t)ns << 32) ^ ((int64_t)stride << 1) ^ (any_random ? 1 : 0);
  int* seen = nullptr;
  for (auto& kv : e->graph_seen) if (kv.first == key) seen = &kv.second;
  if (!seen) { e->graph_seen.push_back({key, 0}); seen = &e->graph_seen.back().second; }
  if ((*seen)++ == 0)     // first sight: run eagerly so every lazy initialisation happens outside a capture
    return engine_forward(e, B, B, B, stride, 1, max_seq_len, dm, any_random, sum_seq, attn_flops);
  const int64_t l0 = e->launches;
  KGE_CUDA_OK(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
  const int rc = engine_forward(e, B, B, B, stride, 1, max_seq_len, dm, any_random, sum_seq, attn_flops);
  cudaGraph_t graph = nullptr;
  cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
  if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
  KGE_REQUIRE(ce == cudaSuccess && graph, "decode graph capture failed: %s", cudaGetErrorString(ce));
  kge_engine::DecodeGraph g{B, ns, stride, any_random, nullptr, e->launches - l0};
  ce = cudaGraphInstantiate(&g.exec, graph, 0);
  cudaGraphDestroy(graph);
  KGE_REQUIRE(ce == cudaSuccess, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ce));
  e->graphs.push_back(g);
  KGE_CUDA_OK(cudaGraphLaunch(g.exec, e->stream));
  return 0;
}

// logits[row, token] = -inf for each (row, token) pair: tokens a request may not produce yet (min_tokens)
__global__ void mask_logits_kernel(bf16* logits, const int32_t* pairs, int n, int ld, int vocab, bool f16) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int row = pairs[2 * i], tok = pairs[2 * i + 1];
  if (tok >= 0 && tok < vocab) reinterpret_cast<uint16_t*>(logits)[(int64_t)row * ld + tok] = f16 ? (uint16_t)0xFC00u : (uint16_t)0xFF80u;
}

__global__ void elt_to_f32_kernel(float* o, const bf16* x, int64_t n, bool f16) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    o[i] = f16 ? __half2float(reinterpret_cast<const __half*>(x)[i]) : bf2f(x[i]);
}

// Copy the caller's host metadata into the pinned block, ship it, run, read the ids back.
static int run_step(kge_engine* e, const kge_step_batch* b, const float* temp, const int32_t* top_k,
                    const float* top_p, const uint64_t* seeds, int32_t* out_ids_host, float* out_logits_host, int lp_n = -1) {
  const int T = b->T, B = b->B, R = b->R;
  KGE_REQUIRE(T > 0 && B > 0, "kge_step: empty batch");
  KGE_REQUIRE(T <= e->cfg.max_num_batched_tokens, "kge_step: T=%d exceeds max_num_batched_tokens=%d", T, e->cfg.max_num_batched_tokens);
  KGE_REQUIRE(B <= e->cfg.max_num_seqs && R <= e->cfg.max_num_seqs, "kge_step: B=%d / R=%d exceed max_num_seqs=%d", B, R, e->cfg.max_num_seqs);
  KGE_REQUIRE(b->max_blocks > 0 && b->max_blocks <= e->max_blocks_per_seq, "kge_step: max_blocks=%d out of range", b->max_blocks);
  KGE_CUDA_OK(cudaSetDevice(e->cfg.device));
  // pure decode step with every row sampled: block-table rows at a bucketed stride so the captured
  // graph of this (B, splits, stride) can be replayed while contexts grow
  bool decode_only = (T == B && R == B && !out_logits_host);
  for (int i = 0; decode_only && i < B; ++i) decode_only = (b->logits_rows[i] == i);
  const int stride = decode_only ? std::min(e->max_blocks_per_seq, (b->max_blocks + 31) / 32 * 32) : b->max_blocks;
  int64_t bytes = 0;
  StepMeta hm = carve_meta(e->meta_host, T, B, R, stride, &bytes);
  StepMeta dm = carve_meta(e->meta_dev, T, B, R, stride, nullptr);
  memcpy(hm.ids, b->token_ids, (size_t)T * 4);
  memcpy(hm.pos, b->positions, (size_t)T * 4);
  memcpy(hm.qsl, b->query_start_loc, (size_t)(B + 1) * 4);
  memcpy(hm.seq_lens, b->seq_lens, (size_t)B * 4);
  if (stride == b->max_blocks) {
    memcpy(hm.btab, b->block_table, (size_t)B * b->max_blocks * 4);
  } else {
    for (int i = 0; i < B; ++i) {
      memcpy(hm.btab + (size_t)i * stride, b->block_table + (size_t)i * b->max_blocks, (size_t)b->max_blocks * 4);
      memset(hm.btab + (size_t)i * stride + b->max_blocks, 0, (size_t)(stride - b->max_blocks) * 4);
    }
  }
  memcpy(hm.slots, b->slot_mapping, (size_t)T * 8);
  if (R > 0) memcpy(hm.rows, b->logits_rows, (size_t)R * 4);
  bool any_random = false;
  for (int i = 0; i < R; ++i) {
    hm.temp[i] = temp ? temp[i] : 0.f;
    hm.top_k[i] = top_k ? top_k[i] : 0;
    hm.top_p[i] = top_p ? top_p[i] : 1.f;
    hm.seeds[i] = seeds ? seeds[i] : 0;
    any_random |= hm.temp[i] >= 1e-5f;
  }
  int max_q = 1, max_seq = 1;
  double sum_seq = 0, attn_flops = 0;
  e->step_num_decode = 0; e->step_dec_max_seq = 1; e->step_dec_sum_seq = 0;
  bool prefix = true;
  for (int i = 0; i < B; ++i) {
    const int ql = b->query_start_loc[i + 1] - b->query_start_loc[i];
    if (prefix && ql == 1) {
      ++e->step_num_decode;
      e->step_dec_max_seq = std::max(e->step_dec_max_seq, b->seq_lens[i]);
      e->step_dec_sum_seq += b->seq_lens[i];
    } else {
      prefix = false;
    }
    sum_seq += b->seq_lens[i];
    attn_flops += 4.0 * ql * ((double)(b->seq_lens[i] - ql) + 0.5 * (ql + 1));
    KGE_REQUIRE(ql >= 1 && ql <= b->seq_lens[i], "kge_step: request %d has q_len=%d seq_len=%d", i, ql, b->seq_lens[i]);
    KGE_REQUIRE(b->seq_lens[i] <= b->max_blocks * e->cfg.block_size, "kge_step: request %d seq_len=%d exceeds its block table", i, b->seq_lens[i]);
    max_q = std::max(max_q, ql);
    max_seq = std::max(max_seq, b->seq_lens[i]);
  }
  KGE_REQUIRE(b->query_start_loc[B] == T, "kge_step: query_start_loc[B]=%d != T=%d", b->query_start_loc[B], T);
  if (e->recording) {
    kge_engine::RecordedStep rs{};
    rs.T = T; rs.B = B; rs.R = R; rs.stride = stride; rs.max_q = max_q; rs.max_seq = max_seq;
    rs.num_decode = e->step_num_decode; rs.dec_max_seq = e->step_dec_max_seq; rs.dec_sum_seq = e->step_dec_sum_seq;
    rs.sum_seq = sum_seq; rs.attn_flops = attn_flops; rs.decode_only = decode_only && max_q == 1; rs.any_random = any_random;
    rs.bytes = bytes; rs.sample0 = (int)e->rec_sample_req.size();
    rs.meta.assign(e->meta_host, e->meta_host + bytes);
    rs.src = e->rec_pending_src;
    rs.src.resize((size_t)T, -1);
    e->rec_pending_src.clear();                 // a raw kge_step that follows has no sampled-token sources
    e->recorded.push_back(std::move(rs));
  }
  KGE_CUDA_OK(cudaMemcpyAsync(e->meta_dev, e->meta_host, (size_t)bytes, cudaMemcpyHostToDevice, e->stream));
  e->h2d_bytes += bytes;
  e->d2h_bytes += (int64_t)R * 4;
  if (decode_only && max_q == 1) {
    if (int rc = engine_decode_step(e, B, stride, max_seq, dm, any_random, sum_seq)) return rc;
  } else {
    if (int rc = engine_forward(e, T, B, R, stride, max_q, max_seq, dm, any_random, sum_seq, attn_flops)) return rc;
  }
  ++e->step_count;
  e->lp_last_n = -1;
  if (R > 0) {
    KGE_CUDA_OK(cudaMemcpyAsync(e->out_host, e->d_out, (size_t)R * 4, cudaMemcpyDeviceToHost, e->stream));
    if (lp_n >= 0) {
      // raw log-probabilities of the sampled tokens (vllm:v1/sample/sampler.py:300-349), for the rows that asked
      if (int rc = e->op->logprobs(e->lp_vals, e->lp_ids, e->lp_rank, e->logits, e->d_out, R, e->spec.vocab, e->spec.vocab, lp_n, e->stream)) return rc;
      ++e->launches;
      const size_t n = (size_t)R * (lp_n + 1) * 4;
      KGE_CUDA_OK(cudaMemcpyAsync(e->lp_vals_host, e->lp_vals, n, cudaMemcpyDeviceToHost, e->stream));
      KGE_CUDA_OK(cudaMemcpyAsync(e->lp_ids_host, e->lp_ids, n, cudaMemcpyDeviceToHost, e->stream));
      KGE_CUDA_OK(cudaMemcpyAsync(e->lp_rank_host, e->lp_rank, (size_t)R * 4, cudaMemcpyDeviceToHost, e->stream));
      e->d2h_bytes += (int64_t)(2 * n + (size_t)R * 4);
      e->lp_last_n = lp_n;
    }
    if (out_logits_host) {
      // debug/parity path: logits as fp32 through a scratch carved from the (idle) gate_up buffer
      float* tmp = reinterpret_cast<float*>(e->gu);
      const int64_t n = (int64_t)R * e->spec.vocab;
      KGE_REQUIRE(n * 4 <= (int64_t)e->cfg.max_num_batched_tokens * e->layers[0].up.N * 2,
                  "kge_step: logits read-back needs R*vocab*4 <= gate_up buffer");
      elt_to_f32_kernel<<<148 * 4, 256, 0, e->stream>>>(tmp, e->logits, n, e->f16);
      KGE_CUDA_OK(cudaMemcpyAsync(out_logits_host, tmp, (size_t)n * 4, cudaMemcpyDeviceToHost, e->stream));
    }
  }
  KGE_CUDA_OK(cudaStreamSynchronize(e->stream));
  if (e->timing) collect_timing(e);
  if (R > 0 && out_ids_host) memcpy(out_ids_host, e->out_host, (size_t)R * 4);
  return 0;
}

}  // namespace kge

extern "C" int kge_step(kge_engine* e, const kge_step_batch* batch, int32_t* out_ids_host, float* out_logits_host) {
  KGE_REQUIRE(e && batch, "kge_step: null argument");
  e->sched->invalidate_prefix_cache();        // the caller's block tables may overwrite blocks the scheduler has cached
  e->step_num_masks = 0;
  return run_step(e, batch, nullptr, nullptr, nullptr, nullptr, out_ids_host, out_logits_host);
}

static SamplingParams to_sp(const kge_sampling_params* p) {
  SamplingParams sp;
  if (!p) return sp;
  sp.max_tokens = p->max_tokens;
  sp.min_tokens = p->min_tokens;
  sp.ignore_eos = p->ignore_eos != 0;
  sp.temperature = p->temperature;
  sp.top_k = p->top_k;
  sp.top_p = p->top_p;
  sp.seed = p->seed;
  sp.has_seed = p->has_seed != 0;
  if (p->stop_token_ids && p->num_stop_token_ids > 0)
    sp.stop_token_ids.assign(p->stop_token_ids, p->stop_token_ids + p->num_stop_token_ids);
  return sp;
}

extern "C" int kge_add_request(kge_engine* e, int64_t request_id, const int32_t* prompt_ids, int32_t n,
                               const kge_sampling_params* sp) {
  KGE_REQUIRE(e && prompt_ids, "kge_add_request: null argument");
  for (int i = 0; i < n; ++i)
    KGE_REQUIRE(prompt_ids[i] >= 0 && prompt_ids[i] < e->spec.vocab, "kge_add_request: token id %d out of vocabulary", prompt_ids[i]);
  const char* why = nullptr;
  if (!e->sched->add(request_id, prompt_ids, n, to_sp(sp), &why)) {
    kge::set_error("kge_add_request: %s", why ? why : "rejected");
    return -1;
  }
  return 0;
}

extern "C" int kge_request_set_logprobs(kge_engine* e, int64_t request_id, int32_t top_n) {
  KGE_REQUIRE(e, "kge_request_set_logprobs: null engine");
  KGE_REQUIRE(top_n >= -1 && top_n <= 20, "kge_request_set_logprobs: top_n=%d must be in [-1, 20]", top_n);
  Request* r = e->sched->find(request_id);
  KGE_REQUIRE(r, "kge_request_set_logprobs: unknown request %lld", (long long)request_id);
  r->sp.logprobs = top_n;
  return 0;
}

extern "C" int kge_engine_step_logprobs(const kge_engine* e, int64_t request_id, float* vals, int32_t* ids, int32_t* rank,
                                        int32_t cap) {
  KGE_REQUIRE(e && vals && ids, "kge_engine_step_logprobs: null argument");
  auto it = e->lp_last.find(request_id);
  if (it == e->lp_last.end() || e->lp_last_n < 0) return 0;
  const int row = it->second.first, n = std::min(it->second.second, e->lp_last_n) + 1;
  KGE_REQUIRE(cap >= n, "kge_engine_step_logprobs: cap=%d < %d entries", cap, n);
  const int stride = e->lp_last_n + 1;
  for (int k = 0; k < n; ++k) { vals[k] = e->lp_vals_host[(size_t)row * stride + k]; ids[k] = e->lp_ids_host[(size_t)row * stride + k]; }
  if (rank) *rank = e->lp_rank_host[row];
  return n;
}

extern "C" int kge_abort_request(kge_engine* e, int64_t request_id) {
  KGE_REQUIRE(e, "kge_abort_request: null engine");
  return e->sched->abort(request_id) ? 0 : -1;
}

extern "C" int kge_has_unfinished(const kge_engine* e) { return e && e->sched->has_unfinished() ? 1 : 0; }

extern "C" int kge_engine_stats(const kge_engine* e, int32_t* num_running, int32_t* num_waiting, float* kv_usage,
                                int64_t* num_preemptions) {
  KGE_REQUIRE(e, "kge_engine_stats: null engine");
  if (num_running) *num_running = e->sched->num_running();
  if (num_waiting) *num_waiting = e->sched->num_waiting();
  if (kv_usage) *kv_usage = 1.f - (float)e->sched->free_blocks() / (float)std::max(1, e->sched->usable_blocks());
  if (num_preemptions) *num_preemptions = e->sched->num_preemptions();
  return 0;
}

extern "C" int kge_engine_step(kge_engine* e, kge_step_output* outs, int32_t cap) {
  KGE_REQUIRE(e && outs, "kge_engine_step: null argument");
  const std::vector<Scheduled>& sc = e->sched->schedule();
  if (sc.empty()) return 0;
  const int bs = e->cfg.block_size;
  const int B = (int)sc.size();
  KGE_REQUIRE(B <= cap, "kge_engine_step: output capacity %d < %d scheduled requests", cap, B);
  std::vector<int32_t> ids, pos, qsl(1, 0), seq_lens, rows, top_k;
  std::vector<int64_t> slots;
  std::vector<float> temp, top_p;
  std::vector<uint64_t> seeds;
  int lp_n = -1;
  int n_masks = 0;
  e->lp_last.clear();
  int max_blocks = 1;
  for (auto& s : sc) max_blocks = std::max(max_blocks, (int)s.req->blocks.size());
  std::vector<int32_t> btab((size_t)B * max_blocks, 0);
  e->rec_pending_src.clear();
  for (int i = 0; i < B; ++i) {
    Request* r = sc[i].req;
    const int n = sc[i].num_tokens, start = r->num_computed - n;
    for (int t = start; t < start + n; ++t) {
      if (e->recording) {
        const int k = t - (int)r->prompt.size();              // >= 0: the k-th output token of this request
        const auto it = e->rec_samples.find(r->id);
        e->rec_pending_src.push_back(k >= 0 && it != e->rec_samples.end() && k < (int)it->second.size() ? it->second[k] : -1);
      }
      ids.push_back(r->token_at(t));
      pos.push_back(t);
      slots.push_back((int64_t)r->blocks[t / bs] * bs + t % bs);
    }
    qsl.push_back((int)ids.size());
    seq_lens.push_back(start + n);
    std::copy(r->blocks.begin(), r->blocks.end(), btab.begin() + (size_t)i * max_blocks);
    if (Scheduler::needs_sample(sc[i])) {
      rows.push_back((int)ids.size() - 1);
      temp.push_back(r->sp.temperature);
      top_k.push_back(r->sp.top_k);
      top_p.push_back(r->sp.top_p);
      // noise stream of this row: a seeded request draws from (seed, index of the token being sampled) only, so the same
      // seeded request gives the same tokens whatever else is in flight; unseeded ones also mix in the engine step
      uint64_t z = r->sp.has_seed ? r->sp.seed : ((uint64_t)r->id * 0x9E3779B97F4A7C15ull + 1) ^ (e->step_count * 0xD1B54A32D192ED03ull);
      z += (uint64_t)r->output.size() * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull;
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
      z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
      seeds.push_back(z ^ (z >> 31));
      if ((int)r->output.size() < r->sp.min_tokens) {
        // this request may not stop yet: its EOS (unless ignored) and its stop ids cannot be sampled in this step
        const int row = (int)rows.size() - 1;
        auto ban = [&](int32_t tok) {
          if (n_masks < kge_engine::kMaxMaskPairs) { e->mask_pairs_host[2 * n_masks] = row; e->mask_pairs_host[2 * n_masks + 1] = tok; ++n_masks; }
        };
        if (!r->sp.ignore_eos && e->spec.eos >= 0) ban(e->spec.eos);
        for (int32_t t : r->sp.stop_token_ids) ban(t);
      }
      if (r->sp.logprobs >= 0) {
        lp_n = std::max(lp_n, r->sp.logprobs);
        e->lp_last[r->id] = {(int)rows.size() - 1, r->sp.logprobs};
      }
    }
  }
  kge_step_batch b;
  b.T = (int)ids.size(); b.B = B;
  b.token_ids = ids.data(); b.positions = pos.data(); b.query_start_loc = qsl.data();
  b.seq_lens = seq_lens.data(); b.block_table = btab.data(); b.max_blocks = max_blocks;
  b.slot_mapping = slots.data(); b.logits_rows = rows.data(); b.R = (int)rows.size();
  e->step_num_masks = n_masks;
  if (n_masks > 0)
    KGE_CUDA_OK(cudaMemcpyAsync(e->mask_pairs, e->mask_pairs_host, (size_t)n_masks * 8, cudaMemcpyHostToDevice, e->stream));
  std::vector<int32_t> sampled(rows.size());
  if (int rc = run_step(e, &b, temp.data(), top_k.data(), top_p.data(), seeds.data(), sampled.data(), nullptr, lp_n)) return rc;
  if (e->recording)          // the sampled rows of this step, in row order, join the job's sample log
    for (int i = 0; i < B; ++i)
      if (Scheduler::needs_sample(sc[i])) {
        e->rec_samples[sc[i].req->id].push_back((int32_t)e->rec_sample_req.size());
        e->rec_sample_req.push_back(sc[i].req->id);
      }
  std::vector<StepResult> res;
  e->sched->update(sampled, &res);
  for (size_t i = 0; i < res.size(); ++i) {
    outs[i].request_id = res[i].request_id;
    outs[i].token = res[i].token;
    outs[i].finished = res[i].finished;
    outs[i].num_scheduled = res[i].num_scheduled;
  }
  return (int)res.size();
}

extern "C" int kge_engine_run_to_completion(kge_engine* e, kge_step_output* events, int64_t cap, int64_t* n_events,
                                            int32_t* n_steps) {
  KGE_REQUIRE(e && events && n_events, "kge_engine_run_to_completion: null argument");
  std::vector<kge_step_output> step((size_t)std::max(1, e->cfg.max_num_seqs));
  int64_t n = 0;
  int32_t steps = 0;
  while (e->sched->has_unfinished()) {
    const int m = kge_engine_step(e, step.data(), (int32_t)step.size());
    if (m < 0) return m;
    KGE_REQUIRE(m > 0, "kge_engine_run_to_completion: no request can be scheduled (KV pool too small for the waiting request?)");
    ++steps;
    for (int i = 0; i < m; ++i) {
      if (step[i].token < 0 && !step[i].finished) continue;      // a prefill chunk advanced: nothing to report
      KGE_REQUIRE(n < cap, "kge_engine_run_to_completion: more than %lld events", (long long)cap);
      events[n++] = step[i];
    }
  }
  *n_events = n;
  if (n_steps) *n_steps = steps;
  return 0;
}

// ------------------------------------------------------------- scheduler alone (host tests)
struct kge_sched { std::unique_ptr<Scheduler> s; };

extern "C" kge_sched* kge_sched_create(int32_t max_num_seqs, int32_t max_num_batched_tokens, int32_t max_model_len,
                                       int32_t block_size, int32_t num_blocks, int32_t enable_chunked_prefill,
                                       int32_t eos_token_id) {
  SchedConfig c;
  c.max_num_seqs = max_num_seqs; c.max_num_batched_tokens = max_num_batched_tokens;
  c.max_model_len = max_model_len; c.block_size = block_size; c.num_blocks = num_blocks;
  c.enable_chunked_prefill = enable_chunked_prefill != 0; c.eos_token_id = eos_token_id;
  kge_sched* k = new kge_sched();
  k->s.reset(new Scheduler(c));
  return k;
}
extern "C" void kge_sched_destroy(kge_sched* s) { delete s; }
extern "C" void kge_sched_set_reserve_full_isl(kge_sched* s, int32_t on) { if (s) s->s->set_reserve_full_isl(on != 0); }
extern "C" void kge_engine_set_reserve_full_isl(kge_engine* e, int32_t on) { if (e && e->sched) e->sched->set_reserve_full_isl(on != 0); }
extern "C" void kge_sched_set_prefix_caching(kge_sched* s, int32_t on) { if (s) s->s->set_prefix_caching(on != 0); }
extern "C" void kge_engine_set_prefix_caching(kge_engine* e, int32_t on) { if (e && e->sched) e->sched->set_prefix_caching(on != 0); }
extern "C" int32_t kge_sched_cached_tokens(kge_sched* s, int64_t request_id) {
  Request* r = s ? s->s->find(request_id) : nullptr;
  return r ? r->num_cached_tokens : -1;
}
extern "C" int kge_engine_prefix_cache_stats(const kge_engine* e, int64_t* queries, int64_t* hits) {
  KGE_REQUIRE(e && e->sched, "kge_engine_prefix_cache_stats: null engine");
  if (queries) *queries = e->sched->prefix_cache_queries();
  if (hits) *hits = e->sched->prefix_cache_hits();
  return 0;
}
extern "C" int kge_sched_add(kge_sched* s, int64_t request_id, const int32_t* prompt_ids, int32_t n,
                             const kge_sampling_params* sp) {
  KGE_REQUIRE(s && prompt_ids, "kge_sched_add: null argument");
  const char* why = nullptr;
  if (!s->s->add(request_id, prompt_ids, n, to_sp(sp), &why)) { kge::set_error("kge_sched_add: %s", why ? why : "rejected"); return -1; }
  return 0;
}
extern "C" int kge_sched_schedule(kge_sched* s, int64_t* ids, int32_t* num_tokens, int32_t cap) {
  KGE_REQUIRE(s && ids && num_tokens, "kge_sched_schedule: null argument");
  const auto& sc = s->s->schedule();
  KGE_REQUIRE((int)sc.size() <= cap, "kge_sched_schedule: capacity %d < %d", cap, (int)sc.size());
  for (size_t i = 0; i < sc.size(); ++i) { ids[i] = sc[i].req->id; num_tokens[i] = sc[i].num_tokens; }
  return (int)sc.size();
}
extern "C" int kge_sched_update(kge_sched* s, const int64_t* ids, const int32_t* tokens, int32_t n,
                                kge_step_output* outs, int32_t cap) {
  KGE_REQUIRE(s, "kge_sched_update: null argument");
  (void)ids;
  std::vector<int32_t> sampled(tokens, tokens + n);
  std::vector<StepResult> res;
  s->s->update(sampled, &res);
  KGE_REQUIRE((int)res.size() <= cap, "kge_sched_update: capacity %d < %d", cap, (int)res.size());
  for (size_t i = 0; i < res.size(); ++i) {
    outs[i].request_id = res[i].request_id; outs[i].token = res[i].token;
    outs[i].finished = res[i].finished; outs[i].num_scheduled = res[i].num_scheduled;
  }
  return (int)res.size();
}
extern "C" int kge_sched_counts(const kge_sched* s, int32_t* running, int32_t* waiting, int32_t* free_blocks) {
  KGE_REQUIRE(s, "kge_sched_counts: null argument");
  if (running) *running = s->s->num_running();
  if (waiting) *waiting = s->s->num_waiting();
  if (free_blocks) *free_blocks = s->s->free_blocks();
  return 0;
}
