/* kge.h - C ABI of the B200 serving-engine hot path (libkge.so).
 *
 * The reference (alikhabazian/Kubernetes-gpu-cluster) has no FFI of its own: its Helm values
 * start a container (`vllm serve <model> ...`, ref:values-01-minimal-example.yaml:4-7,
 * ref:old_README.md:1089-1095) whose engine runs the LLM forward pass.  This header is the
 * boundary between that engine's host side (scheduler / HTTP server, C++ or Python) and the
 * sm_100a kernels.  Every entry point cites the step of the deployed engine it replaces
 * ("vllm:" = vLLM 0.22.0 sources, the engine the values files pin by image tag).
 *
 * Conventions: plain pointers and sizes only; all tensor pointers are DEVICE pointers unless
 * the name says host; bf16 tensors are passed as void*; `stream` is a cudaStream_t passed as
 * void* (NULL = default stream); return 0 = ok, negative = error with text in
 * kge_last_error() (thread-local).  No entry point falls back to the CPU: without a CUDA
 * device every compute call fails with -2.
 */
#ifndef KGE_H_
#define KGE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define KGE_ABI_VERSION 1

const char* kge_last_error(void);
int kge_abi_version(void);
/* 0 = no usable CUDA device, else the SM count of the current device. */
int kge_device_sm_count(void);
/* Launch-shape knobs for tuning and diagnostics (no counterpart in the reference): they change how work is
 * partitioned over the 148 SMs, never the results.  Process-wide; set them before creating an engine
 * (captured decode graphs keep the shape they were captured with).
 *   "pdl" 0/1                 programmatic dependent launch between the kernels of a step
 *   "gemm_even_grid" 0/1      persistent weight-streaming GEMM: size the grid so every CTA gets the same tile count
 *   "attn_decode_splits" n    force the split-KV factor of decode attention (0 = heuristic)
 *   "attn_split_balance" 0/1  heuristic picks the split factor that fills whole waves of resident CTAs
 *   "attn_cluster_merge" 0/1  decode attention: 1 = the split-KV CTAs of a (request, kv head) form a thread-block
 *                             cluster and merge over distributed shared memory instead of workspace + merge kernel
 *                             (same bits; measured slower, so off by default)
 *   "attn_tma" 1/0            attention K/V tiles fetched by the TMA unit (one producer thread, cp.async.bulk.tensor boxes
 *                             of one KV block) instead of per-thread 16-byte cp.async gathers; 0 = the round-1 kernels, kept
 *                             for same-box comparison and for block sizes below 8
 *   "gemm_ws" 0/1             1 = decode-sized projections split K over ordinary CTAs, two per SM, reduced through the caller's
 *                             workspace by the last CTA to arrive (faster alone, 19.5 vs 25.5 us on the 8B qkv shape, but not
 *                             inside the decode step, so off by default); 0 = the cluster / distributed-shared-memory split
 *   "smem_carveout" 1/0       1 = every kernel prefers the all-shared-memory L1 split, so consecutive kernels never make an
 *                             SM reconfigure (and drain) between them; 0 = small kernels keep the driver's default split
 *   "gemm_l2_hints" 0/1       1 = the GEMM's TMA loads carry L2 eviction hints (weights evict_first, activations evict_last);
 *                             measured slower on the prefill step (weight tiles are re-read by other CTA pairs), so off
 *   "gemm_wide" 1/0           GEMMs of up to 64 tokens: 1 = where it keeps as many CTAs streaming as the 128-row plan, tiles of
 *                             256 weight rows with the tokens as the MMA's M side and the weights as its N side (a tcgen05.mma
 *                             fed from shared memory costs ~165 cycles whatever its shape, so twice the weight bytes per MMA);
 *                             0 = always 128-row tiles with the tokens as N.  C3 decode step 6.54-6.58 vs 6.64-6.71 ms
 *   "attn_persistent" 1/0     decode attention with 64-token KV blocks or larger: 1 = 2 x 148 resident CTAs each take an equal,
 *                             contiguous share of all key tiles of the launch (a (request, kv head) group is cut where shares
 *                             end; pieces merged by the last CTA of the group to arrive) - one round whatever the batch and
 *                             the context lengths are (launches of up to 192 requests with a workspace of at least three partial slots per
 *                             head; others use the other kernel); 0 = one CTA per (request, kv head, split) with equal splits.
 *                             64 x 2176 keys: 91.6-93.3 us against 95.6; 64 x 1024: 49.7 against 56.5; 8 x 8192: 53.0 against 60.2
 *                             (profiles/r02_summary.md section 6)
 * Returns 0, or -1 for an unknown name. */
int kge_set_tunable(const char* name, int value);

/* Measurement hook (no counterpart in the reference): while `device_buffer` is non-null every following GEMM launch
 * writes 16 uint64 per CTA there - %globaltimer at kernel entry, after set-up, at the first tile's arrival, after the last
 * MMA, after the split exchange, at exit, and the SM id (slot list in tools/gemm_trace.py).  Null switches it off. */
void kge_gemm_set_trace(void* device_buffer);

/* ---------------------------------------------------------------- per-op entry points --- */

/* Token (+ learned position) embedding gather.
 * vllm:model_executor/models/llama.py:392-393,407; opt.py:61-70,276-281 (offset 2).
 * out[t,:] = table[ids[t],:] (+ pos_table[positions[t]+pos_offset,:], one bf16 rounding). */
int kge_embed(void* out, const int32_t* ids, const void* table, const void* pos_table,
              const int32_t* positions, int pos_offset, int T, int d, void* stream);

/* RMSNorm, vllm:ir/ops/layernorm.py:10-21.  y = bf16(bf16(x*rsqrt(mean(x^2)+eps)) * w). */
int kge_rmsnorm(void* y, const void* x, const void* w, int T, int d, float eps, void* stream);

/* Fused residual-add + RMSNorm, vllm:ir/ops/layernorm.py:40-58.
 * s = fp32(x)+fp32(residual); residual <- bf16(s); y = norm(s) (un-rounded s). */
int kge_add_rmsnorm(void* y, void* residual, const void* x, const void* w, int T, int d,
                    float eps, void* stream);

/* Per-head RMSNorm of the q and k parts of a packed qkv buffer [T,(H+2KV)*hd], in place, before RoPE:
 * Qwen3's q_norm / k_norm (vllm:model_executor/models/qwen3.py:142-160), weights [hd] each, same
 * arithmetic as kge_rmsnorm over head_dim. */
int kge_head_rmsnorm(void* qkv, const void* q_weight, const void* k_weight, int T, int H, int KV, int hd,
                     float eps, void* stream);

/* LayerNorm (OPT), vllm:model_executor/models/opt.py:148-150,166-168,250-253.
 * If `add` is non-NULL: h = bf16(fp32(residual)+fp32(add)) is written back to `residual`
 * first (opt.py:181,194) and the norm is taken of h; y may be NULL (add only). */
int kge_layernorm(void* y, void* residual, const void* add, const void* w, const void* b, int T,
                  int d, float eps, void* stream);

/* y[T,f] = silu(x[:, :f]) * x[:, f:], vllm:model_executor/layers/activation.py:137-141. */
int kge_silu_mul(void* y, const void* x, int T, int f, void* stream);

/* Dense projection y[T,N] = act(x[T,K] * w[N,K]^T + bias[N]); bf16 in/out, fp32 accumulate,
 * one rounding.  act: 0 none, 1 relu.  Replaces F.linear in QKV/O/MLP/LM-head
 * (vllm:model_executor/models/llama.py:94-121,164-180; opt.py:92-106,151-165;
 * layers/logits_processor.py:89-99).  tcgen05/TMEM/TMA kernel; K % 8 == 0 (16-byte rows; a partial last
 * 64-wide k-block is zero-filled by TMA), x and w 16-byte aligned.
 * `workspace`: optional device scratch of >= kge_gemm_workspace_bytes(T, N, K) bytes, 16-byte aligned, whose first
 * 1024 bytes are ZERO before the first call that uses it (arrival counters; every call leaves them zero again).  With it,
 * decode-sized shapes split K over two CTAs per SM and reduce through the scratch; without it (NULL, or too small) they
 * split K inside a thread-block cluster.  Same results either way up to fp32 summation order.  Calls that share one
 * workspace must not overlap on the GPU. */
int kge_gemm_bf16(void* y, const void* x, const void* w, const void* bias, int T, int N, int K,
                  int act, void* workspace, int64_t workspace_bytes, void* stream);
int64_t kge_gemm_workspace_bytes(int T, int N, int K);

/* Fused SwiGLU front half: y[T,f] = silu(x * w[:f]^T) * (x * w[f:]^T) with w = [gate | up] stacked
 * [2f, K] (vllm:model_executor/models/llama.py:94-110 gate_up_proj + activation.py:137-141), same
 * cast points as the two separate ops (both GEMM outputs and the SiLU rounded to bf16 before the
 * product).  One kernel: each 128-row MMA tile holds 64 gate rows and their 64 up rows, so the
 * [T,2f] intermediate never goes to HBM.  f % 64 == 0. */
int kge_gemm_silu_gate_bf16(void* y, const void* x, const void* w_gate_up, int T, int f, int K,
                            void* stream);
/* the same with the optional split-K workspace of kge_gemm_bf16 (kge_gemm_workspace_bytes(T, f, K) bytes) */
int kge_gemm_silu_gate_ws_bf16(void* y, const void* x, const void* w_gate_up, int T, int f, int K, void* workspace,
                               int64_t workspace_bytes, void* stream);

/* RoPE (neox pairing, Llama-3 scaled cache) on the q and k parts of a packed qkv buffer
 * [T,(H+2KV)*hd], in place, then scatter of k and v rows into the paged cache.
 * vllm:model_executor/layers/rotary_embedding/common.py:160-183 (each op rounds in bf16);
 * vllm:v1/attention/backends/flash_attn.py:850-874 (reshape_and_cache).
 * cos_sin [max_pos, hd] bf16 = cos(hd/2) | sin(hd/2) per row, or NULL for no rotation (OPT).
 * Cache layout (this engine's, HBM-resident): [num_blocks, KV, block_size, hd] bf16 for K and
 * for V; slot = block*block_size + offset; slot < 0 skips the row. */
int kge_rope_kvwrite(void* qkv, const void* cos_sin, const int32_t* positions,
                     const int64_t* slot_mapping, void* k_cache, void* v_cache, int T, int H,
                     int KV, int hd, int block_size, void* stream);

/* Build the cos|sin cache on the host exactly as
 * vllm:model_executor/layers/rotary_embedding/base.py:69-93 + llama3_rope.py:33-54 and round
 * to bf16.  out_host: [max_pos, hd] uint16 bf16 bit patterns.  factor == 0 => no scaling. */
int kge_rope_cache_host(uint16_t* out_host, int max_pos, int hd, double theta, double factor,
                        double low_freq_factor, double high_freq_factor, int orig_max_pos);

/* Causal paged attention over a flattened batch (prefill chunks and decode rows mixed):
 * vllm:v1/attention/backends/flash_attn.py:667; semantics vllm:v1/attention/backends/
 * cpu_attn.py:292-398.  Request b owns query rows [qsl[b], qsl[b+1]) which are the LAST
 * rows of its seq_lens[b]-long context; q is read from the packed qkv buffer (row stride
 * q_stride elements); out [T, H*hd].  scale = hd^-0.5 applied in fp32; softmax in fp32.
 * `workspace` >= kge_attn_workspace_bytes() of device scratch (split-KV partials). */
int kge_paged_attn(void* out, const void* q, int q_stride, const void* k_cache,
                   const void* v_cache, const int32_t* block_table, int max_blocks_per_seq,
                   const int32_t* seq_lens, const int32_t* query_start_loc, int B, int T,
                   int max_q_len, int max_seq_len, int H, int KV, int hd, int block_size,
                   float scale, void* workspace, int64_t workspace_bytes, void* stream);
int64_t kge_attn_workspace_bytes(int B, int H, int hd, int max_seq_len);

/* Decode attention (one query token per request) for hosts that cannot allocate device memory.
 * No counterpart in the reference or in the engine it deploys: this entry point exists only here.
 *
 * Use it when your process has no way to hand kge_paged_attn a workspace: the engine's KV pool is
 * sized from --gpu-memory-utilization and may have left nothing for a later cudaMalloc, and a
 * serving loop that must keep every stream moving cannot afford the device-wide stall cudaMalloc
 * causes.  You still pass the inputs and the output; the split-KV partial results go to an array
 * compiled into libkge.so, kge_paged_attn_decode_noalloc_max_bytes() (8 MiB) of device memory that
 * every GPU using the library holds until the process exits.  The call itself performs no
 * cudaMalloc / cudaFree and calls nothing that allocates.
 *
 * Limits and rules:
 *  - a batch whose partials do not fit (B * H * splits * (hd + 2) * 4 bytes, splits as chosen by
 *    kge_paged_attn for the same B / KV / max_seq_len) is rejected with -1 and a message saying so;
 *  - the FIRST call on each device looks up the array's device address (with that device current).
 *    Make that first call during start-up, before any stream capture begins;
 *  - all calls on a device share the one array, which keeps whatever the last call left in it.
 *    Nothing in it needs clearing (each launch overwrites every slot it later reads), but two
 *    calls that run on the GPU at the same time - different streams, different threads - corrupt
 *    each other's partials: serialise them;
 *  - same arguments and results as kge_paged_attn with max_q_len == 1 (H / KV <= 16, hd 64 or 128). */
int kge_paged_attn_decode_noalloc(void* out, const void* q, int q_stride, const void* k_cache,
                                  const void* v_cache, const int32_t* block_table,
                                  int max_blocks_per_seq, const int32_t* seq_lens,
                                  const int32_t* query_start_loc, int B, int max_seq_len, int H, int KV,
                                  int hd, int block_size, float scale, void* stream);
int64_t kge_paged_attn_decode_noalloc_max_bytes(void);

/* Greedy sampling, vllm:v1/sample/sampler.py:91,235-236: argmax over fp32(bf16 logits),
 * lowest index wins ties.  logits [B, ld] bf16 (ld >= V: padded vocab columns ignored). */
int kge_argmax(int32_t* out_ids, const void* logits, int B, int V, int ld, void* stream);

/* Temperature / top-k / top-p sampling, vllm:v1/sample/sampler.py:254-296 and
 * vllm:v1/sample/ops/topk_topp_sampler.py:328-369,395-417: rows with temperature < 1e-5 are
 * greedy; others draw argmax(softmax(mask(logits/T)) / q), q ~ Exp(1).  `exp_noise` [B,V]
 * fp32 supplies q (NULL: Philox noise from seeds[b] and step).  top_k[b] <= 0 or >= V and
 * top_p[b] >= 1 disable the filters.  `workspace` >= kge_sample_workspace_bytes(B,V). */
int kge_sample(int32_t* out_ids, const void* logits, int B, int V, int ld,
               const float* temperature, const int32_t* top_k, const float* top_p,
               const float* exp_noise, const uint64_t* seeds, uint64_t step, void* workspace,
               int64_t workspace_bytes, void* stream);
int64_t kge_sample_workspace_bytes(int B, int V);

/* Log-probabilities of sampled tokens, vllm:v1/sample/sampler.py:300-349 (compute_logprobs, gather_logprobs) and
 * vllm:v1/sample/ops/logprobs.py:11-27: logprobs = log_softmax(fp32(logits)).  Per row: out_vals/out_ids
 * [R, top_n+1] = the sampled token first, then the top_n most likely tokens (descending; equal values by ascending
 * id); out_rank [R] = number of vocabulary entries whose logit is >= the sampled token's (1 = it was the argmax).
 * logits [R, ld] bf16, sampled_ids [R] on the device; 0 <= top_n <= 20. */
int kge_logprobs(float* out_vals, int32_t* out_ids, int32_t* out_rank, const void* logits,
                 const int32_t* sampled_ids, int R, int V, int ld, int top_n, void* stream);

/* ------------------------------------------------------------ float16 twins of the ops ---
 * `--dtype float16` (ref:values-01-minimal-example8.yaml:30-31, ...example9.yaml:31-32): the same kernels compiled for
 * IEEE half storage - same arguments, same cast points (every "rounded to the model dtype" step rounds to float16;
 * accumulation, softmax and norm statistics stay fp32), tcgen05.mma kind::f16 with F16 operands.  Buffers the bfloat16
 * entry points describe as bf16 are float16 here. */
int kge_embed_f16(void* out, const int32_t* ids, const void* table, const void* pos_table, const int32_t* positions,
                  int pos_offset, int T, int d, void* stream);
int kge_rmsnorm_f16(void* y, const void* x, const void* w, int T, int d, float eps, void* stream);
int kge_add_rmsnorm_f16(void* y, void* residual, const void* x, const void* w, int T, int d, float eps, void* stream);
int kge_head_rmsnorm_f16(void* qkv, const void* q_weight, const void* k_weight, int T, int H, int KV, int hd, float eps,
                         void* stream);
int kge_layernorm_f16(void* y, void* residual, const void* add, const void* w, const void* b, int T, int d, float eps,
                      void* stream);
int kge_silu_mul_f16(void* y, const void* x, int T, int f, void* stream);
int kge_gemm_f16(void* y, const void* x, const void* w, const void* bias, int T, int N, int K, int act, void* workspace,
                 int64_t workspace_bytes, void* stream);
int kge_gemm_silu_gate_f16(void* y, const void* x, const void* w_gate_up, int T, int f, int K, void* stream);
int kge_gemm_silu_gate_ws_f16(void* y, const void* x, const void* w_gate_up, int T, int f, int K, void* workspace,
                              int64_t workspace_bytes, void* stream);
int kge_rope_kvwrite_f16(void* qkv, const void* cos_sin, const int32_t* positions, const int64_t* slot_mapping,
                         void* k_cache, void* v_cache, int T, int H, int KV, int hd, int block_size, void* stream);
int kge_rope_cache_host_f16(uint16_t* out_host, int max_pos, int hd, double theta, double factor,
                            double low_freq_factor, double high_freq_factor, int orig_max_pos);
int kge_paged_attn_f16(void* out, const void* q, int q_stride, const void* k_cache, const void* v_cache,
                       const int32_t* block_table, int max_blocks_per_seq, const int32_t* seq_lens,
                       const int32_t* query_start_loc, int B, int T, int max_q_len, int max_seq_len, int H, int KV, int hd,
                       int block_size, float scale, void* workspace, int64_t workspace_bytes, void* stream);
int kge_argmax_f16(int32_t* out_ids, const void* logits, int B, int V, int ld, void* stream);
int kge_sample_f16(int32_t* out_ids, const void* logits, int B, int V, int ld, const float* temperature,
                   const int32_t* top_k, const float* top_p, const float* exp_noise, const uint64_t* seeds, uint64_t step,
                   void* workspace, int64_t workspace_bytes, void* stream);
int kge_logprobs_f16(float* out_vals, int32_t* out_ids, int32_t* out_rank, const void* logits, const int32_t* sampled_ids,
                     int R, int V, int ld, int top_n, void* stream);
/* (kge_tp_reduce_add_norm_f16 is declared with the tensor-parallel exchange below) */

/* -------------------------------------------------------------------- engine (model) --- */

typedef struct kge_engine kge_engine;

typedef struct {
  int32_t max_num_seqs;           /* vllm --max-num-seqs                                  */
  int32_t max_num_batched_tokens; /* vllm --max-num-batched-tokens (per-step token budget) */
  int32_t max_model_len;          /* vllm --max-model-len                                 */
  int32_t block_size;             /* KV block size in tokens (0: 64; 16 and 32 accepted)  */
  int32_t num_blocks;             /* 0: size from gpu_memory_utilization                  */
  float gpu_memory_utilization;   /* vllm --gpu-memory-utilization                        */
  int32_t tp_rank, tp_size;       /* vllm --tensor-parallel-size                          */
  int32_t enable_chunked_prefill; /* vllm --enable-chunked-prefill                        */
  int32_t device;                 /* CUDA device ordinal                                   */
  int32_t dtype;                  /* vllm --dtype: 0 auto (config.json torch_dtype: float16 -> float16, else bfloat16),
                                     1 bfloat16, 2 float16 (ref:values-01-minimal-example8.yaml:30-31)                  */
} kge_engine_config;

/* One engine step's device-side metadata - what vLLM hands its attention backend
 * (vllm:v1/attention/backends/cpu_attn.py:107-118).  HOST pointers; kge_step copies them. */
typedef struct {
  int32_t T, B;
  const int32_t* token_ids;       /* [T]   */
  const int32_t* positions;       /* [T]   */
  const int32_t* query_start_loc; /* [B+1] */
  const int32_t* seq_lens;        /* [B]   */
  const int32_t* block_table;     /* [B, max_blocks] */
  int32_t max_blocks;
  const int64_t* slot_mapping;    /* [T]   */
  const int32_t* logits_rows;     /* [R] rows of the flattened batch that need logits */
  int32_t R;
} kge_step_batch;

/* Load an HF-layout directory (config.json + *.safetensors), the same thing the values
 * files point the engine at (ref:values-01-minimal-example3.yaml:8,22-30), or, when
 * `dummy_seed` >= 0 and the directory holds only config.json, random weights
 * (--load-format dummy).  Allocates weights, activations and the paged KV pool. */
int kge_engine_create(kge_engine** out, const char* model_dir, const kge_engine_config* cfg,
                      int64_t dummy_seed);
void kge_engine_destroy(kge_engine* e);
int kge_engine_num_blocks(const kge_engine* e);
int kge_engine_dtype(const kge_engine* e);        /* 1 bfloat16, 2 float16: what the engine settled on */
int kge_engine_vocab(const kge_engine* e);
/* eos_token_id of config.json (-1 if none): what ends a request unless ignore_eos (vllm:v1/core/sched/utils.py:94-117). */
int kge_engine_eos_token_id(const kge_engine* e);
/* The context length in force: --max-model-len, or min(max_position_embeddings, 8192) when it was left out. */
int kge_engine_max_model_len(const kge_engine* e);
int64_t kge_engine_weight_bytes(const kge_engine* e);
int64_t kge_engine_launches(const kge_engine* e);

/* Host-only view of a model directory - what kge_engine_create reads, without a GPU (tests, tooling):
 * config.json -> the model description as JSON; *.safetensors tensors (all shards, BF16 / F16 / F32) sliced and
 * converted to bf16 bits with round-to-nearest-even, exactly as the engine uploads them.  r1 / c1 = -1 mean "to the end";
 * returns the number of elements written, -1 on error. */
typedef struct kge_model_dir kge_model_dir;
kge_model_dir* kge_model_open(const char* dir);
void kge_model_close(kge_model_dir* m);
int kge_model_spec_json(const kge_model_dir* m, char* out, int cap);
int64_t kge_model_read_bf16(const kge_model_dir* m, const char* name, int64_t r0, int64_t r1, int64_t c0, int64_t c1,
                            uint16_t* out, int64_t cap);
/* the same slice as IEEE half bits: what a float16 engine (--dtype float16) uploads */
int64_t kge_model_read_f16(const kge_model_dir* m, const char* name, int64_t r0, int64_t r1, int64_t c0, int64_t c1,
                           uint16_t* out, int64_t cap);

/* The model directory's tokenizer (tokenizer.json byte-level BPE of the GPT-2 / Llama-3 / Qwen families, or GPT-2's
 * vocab.json + merges.txt), for front ends that take and return text like the engine they replace
 * (vllm:entrypoints/llm.py accepts str prompts).  Host only.  encode / decode return the length and copy only when the
 * result fits `cap` (call with cap = 0 to size a buffer); decode skips special tokens and replaces invalid UTF-8 the way
 * HF's byte-level decoder does. */
typedef struct kge_tokenizer kge_tokenizer;
kge_tokenizer* kge_tokenizer_open(const char* model_dir);
void kge_tokenizer_close(kge_tokenizer* t);
int kge_tokenizer_vocab_size(const kge_tokenizer* t);
int64_t kge_tokenizer_encode(const kge_tokenizer* t, const char* text, int64_t text_bytes, int add_special_tokens,
                             int32_t* out_ids, int64_t cap);
int64_t kge_tokenizer_decode(const kge_tokenizer* t, const int32_t* ids, int64_t n, char* out, int64_t cap);

/* One forward pass + greedy sampling of the R requested rows (the GPUModelRunner.execute_model
 * + Sampler path, vllm:v1/worker/gpu_model_runner.py:3955, vllm:v1/sample/sampler.py:67).
 * out_ids_host [R]; out_logits_host [R, vocab] fp32 or NULL.  Synchronous. */
int kge_step(kge_engine* e, const kge_step_batch* batch, int32_t* out_ids_host,
             float* out_logits_host);

/* Per-kernel timing for the roofline report: when enabled, every launch of the forward pass is
 * bracketed by CUDA events on the engine's stream and accumulated per category together with its
 * ALGORITHMIC work (bytes that must cross HBM once / flops), DESIGN.md "kernels".  Categories:
 * 0 gemm (T<=256: weight streaming), 1 gemm (T>256: tensor bound), 2 decode attention,
 * 3 general (prefill / mixed) attention, 4 norms, 5 rope+kv write, 6 activation, 7 embed/gather,
 * 8 sampling, 9 all-reduce.  Timing adds event overhead: never enable it for a headline number. */
#define KGE_NUM_KERNEL_CATS 10
typedef struct {
  int64_t launches;
  double ms;          /* summed device time of the launches            */
  double bytes;       /* summed algorithmic HBM bytes                  */
  double flops;       /* summed algorithmic flops                      */
} kge_kernel_stats;
int kge_engine_set_timing(kge_engine* e, int enable);
int kge_engine_get_timing(kge_engine* e, kge_kernel_stats* out /* [KGE_NUM_KERNEL_CATS] */);
/* Host<->device traffic of the serving loop so far: per-step metadata in, sampled ids out. */
int kge_engine_io_bytes(const kge_engine* e, int64_t* h2d, int64_t* d2h);

/* Device-resident replay of one fixed-shape job (the bench's `value` leg): B requests of
 * prompt_len tokens each are prefilled whole, `prefill_batch` requests per step, then decoded
 * out_len-1 more steps greedily; all step metadata is uploaded BEFORE the timed region, sampled ids
 * feed the next step on the device, and nothing is synchronised until the end.  prompts_host
 * [B, prompt_len]; out_tokens_host [B, out_len]; *ms = device time of the whole job (CUDA events on
 * the engine's stream); *decode_ms = the decode part.  Uses KV blocks 1..; the scheduler must be idle. */
int kge_engine_run_resident(kge_engine* e, const int32_t* prompts_host, int32_t B, int32_t prompt_len,
                            int32_t out_len, int32_t prefill_batch, int32_t* out_tokens_host, float* ms,
                            float* decode_ms);

/* The same measurement on the steps the SCHEDULER makes: record a job that is driven through kge_add_request /
 * kge_engine_step (or kge_engine_run_to_completion), then replay exactly those steps - same batches, same block tables,
 * same kernels - with all metadata resident in HBM and every sampled id handed to its next step on the device.  The
 * replay's sampled tokens come back in sampling order with the request each belongs to; they equal the recorded job's.
 * *decode_ms = device time of the trailing run of pure decode steps. */
int kge_engine_record_steps(kge_engine* e, int enable);
int kge_engine_replay_recorded(kge_engine* e, int32_t* out_tokens_host, int64_t* out_request_ids_host, int64_t cap,
                               int64_t* n_samples, int32_t* n_steps, float* ms, float* decode_ms);

/* Tensor parallelism (one engine = one rank = one GPU): all ranks share a 128-byte NCCL
 * unique id made by rank 0; after kge_engine_init_tp the forward pass sums the row-parallel
 * projections with a bf16 ncclAllReduce (vllm:model_executor/layers/linear.py:1559;
 * ref:values-01-minimal-example8.yaml:32-38 --disable-custom-all-reduce --tensor-parallel-size 2).
 * Every rank must then call kge_step / kge_engine_step with identical batches. */
int kge_nccl_unique_id(char* out128);
int kge_engine_init_tp(kge_engine* e, const char* id128);

/* Tensor-parallel exchange over peer memory (NVLink / NVSwitch) - the default TP data path; the NCCL path above is what
 * --disable-custom-all-reduce selects (ref:values-01-minimal-example8.yaml:32).
 *
 * Every rank owns one window of device memory; all ranks map all windows (CUDA IPC between processes, peer access
 * between threads of one process).  kge_tp_comm_info fills the record a rank publishes; after every rank has collected
 * all `world` records (any transport: a file, torch.distributed, MPI) it calls kge_tp_comm_connect.  Offsets below are
 * relative to kge_tp_comm_data() and must be the same on every rank.
 *
 * kge_tp_reduce_add_norm: x[T,d] (bf16 partial sums, one per rank, at x_off of each window) are summed in fp32 in rank
 *   order and rounded once to bf16 - what ncclAllReduce(sum, bf16) of vllm:model_executor/layers/linear.py:1559 returns -
 *   then, fused, what always follows it:
 *     mode 1  fused add + RMSNorm  (vllm:ir/ops/layernorm.py:40-58; llama.py:323-331,430)
 *     mode 3  residual add + LayerNorm of the rounded sum  (vllm:model_executor/models/opt.py:181-196)
 *     mode 0  nothing (plain all-reduce; y_off may equal x_off)
 *   The result y[T,d] lands at y_off of EVERY rank's window.  Rows are partitioned over ranks: rank r reduces and
 *   normalises rows [r*ceil(T/N), (r+1)*ceil(T/N)) only, so `residual` (a local [T,d] buffer, read and updated) is
 *   maintained for a rank's OWN rows only; the other rows of it are never read.  Every rank must make the same call.
 * kge_tp_all_gather_cols: dst[r, col0 + c] at dst_off of every rank's window <- src[r, c] (local [R, cols]); the
 *   all-gather of vocab-sharded logits (vllm:model_executor/layers/logits_processor.py:83).
 * A peer that never arrives makes the kernel trap after ~20 s (an error on the stream) instead of hanging the GPU. */
typedef struct kge_tp_comm kge_tp_comm;
typedef struct {
  char ipc_handle[64];   /* cudaIpcMemHandle_t of the window                                   */
  uint64_t ptr;          /* its device address in the owning process (used by same-process peers) */
  int64_t pid;
  int64_t bytes;
  int32_t device;
  int32_t rank;
} kge_tp_peer_info;
int kge_tp_comm_create(kge_tp_comm** out, int rank, int world, int device, int64_t data_bytes);
void kge_tp_comm_destroy(kge_tp_comm* c);
int kge_tp_comm_info(kge_tp_comm* c, kge_tp_peer_info* out);
int kge_tp_comm_connect(kge_tp_comm* c, const kge_tp_peer_info* all /* [world], indexed by rank */);
int kge_tp_comm_connected(const kge_tp_comm* c);
void* kge_tp_comm_data(kge_tp_comm* c);
int kge_tp_reduce_add_norm(kge_tp_comm* c, int64_t y_off, void* residual, int64_t x_off, const void* w, const void* b,
                           int T, int d, float eps, int mode, void* stream);
int kge_tp_reduce_add_norm_f16(kge_tp_comm* c, int64_t y_off, void* residual, int64_t x_off, const void* w, const void* b,
                               int T, int d, float eps, int mode, void* stream);      /* float16 twin */
int kge_tp_all_gather_cols(kge_tp_comm* c, int64_t dst_off, const void* src, int R, int cols, int ld_dst, int col0,
                           void* stream);
/* The engine's own communicator (created with the engine when tp_size > 1): publish / connect as above.  Once
 * connected, the forward pass uses the fused exchange, programmatic dependent launch and decode CUDA graphs; an
 * engine that was only given kge_engine_init_tp keeps the NCCL all-reduce path. */
int kge_engine_tp_info(kge_engine* e, kge_tp_peer_info* out);
int kge_engine_tp_connect(kge_engine* e, const kge_tp_peer_info* all);

/* ----------------------------------------------- scheduler + serving loop (host side) --- */

typedef struct {
  int32_t max_tokens;
  int32_t min_tokens;
  int32_t ignore_eos;
  float temperature;     /* < 1e-5 => greedy (vllm:sampling_params.py:25) */
  int32_t top_k;         /* <= 0 => off */
  float top_p;           /* >= 1 => off */
  uint64_t seed;         /* used when has_seed != 0 (0 is a valid seed) */
  const int32_t* stop_token_ids;
  int32_t num_stop_token_ids;
  int32_t has_seed;      /* request carries its own seed: its random tokens depend only on (seed, output index),
                            not on batch composition or arrival time (vllm:v1/sample/ops/topk_topp_sampler.py:180-184) */
} kge_sampling_params;

typedef struct {
  int64_t request_id;
  int32_t token;         /* sampled token, or -1 if the step only advanced a prefill chunk */
  int32_t finished;      /* 0 running, 1 "stop", 2 "length"                                 */
  int32_t num_scheduled; /* tokens of this request run this step                            */
} kge_step_output;

/* Continuous batching: vllm:v1/core/sched/scheduler.py:329-830,1283 (RUNNING first under the
 * token budget, tail preemption + recompute, WAITING admission), stop rules
 * vllm:v1/core/sched/utils.py:94-130, block manager vllm:v1/core/kv_cache_manager.py:236-430. */
int kge_add_request(kge_engine* e, int64_t request_id, const int32_t* prompt_ids, int32_t n,
                    const kge_sampling_params* sp);
int kge_abort_request(kge_engine* e, int64_t request_id);
/* SamplingParams.logprobs (vllm:sampling_params.py): after kge_add_request, ask for the top_n log-probabilities of
 * every token the request samples (0 = only the sampled token's, -1 = none).  After each kge_engine_step,
 * kge_engine_step_logprobs copies what that step produced for the request: entry 0 is the sampled token, then the
 * top_n; returns the number of entries (0 when the request sampled nothing in that step). */
int kge_request_set_logprobs(kge_engine* e, int64_t request_id, int32_t top_n);
int kge_engine_step_logprobs(const kge_engine* e, int64_t request_id, float* vals, int32_t* ids, int32_t* rank,
                             int32_t cap);
int kge_has_unfinished(const kge_engine* e);
/* Schedule + forward + sample + update.  Writes up to cap outputs; returns the count (>= 0)
 * or a negative error. */
int kge_engine_step(kge_engine* e, kge_step_output* outs, int32_t cap);
/* kge_engine_step in a loop until no request is left (what LLM.generate does, vllm:entrypoints/llm.py:422, without
 * a host-language round trip per step).  Appends one event per sampled token / finished request, in step order;
 * cap >= sum of max_tokens + number of requests is always enough.  Returns 0 or a negative error. */
int kge_engine_run_to_completion(kge_engine* e, kge_step_output* events, int64_t cap, int64_t* n_events,
                                 int32_t* n_steps);
/* Counters for /metrics (vllm:v1/metrics/loggers.py:452-788). */
int kge_engine_stats(const kge_engine* e, int32_t* num_running, int32_t* num_waiting,
                     float* kv_usage, int64_t* num_preemptions);

/* The scheduler alone (no device), for host-logic tests: same policy object the engine uses. */
typedef struct kge_sched kge_sched;
kge_sched* kge_sched_create(int32_t max_num_seqs, int32_t max_num_batched_tokens,
                            int32_t max_model_len, int32_t block_size, int32_t num_blocks,
                            int32_t enable_chunked_prefill, int32_t eos_token_id);
void kge_sched_destroy(kge_sched* s);
/* vllm --scheduler-reserve-full-isl / --no-scheduler-reserve-full-isl (vllm:config/scheduler.py, default on;
 * the admission gate of vllm:v1/core/kv_cache_manager.py:346-360): a waiting request is admitted only when
 * free blocks cover its whole known sequence.  The engine's scheduler has the same switch. */
void kge_sched_set_reserve_full_isl(kge_sched* s, int32_t on);
void kge_engine_set_reserve_full_isl(kge_engine* e, int32_t on);
/* vllm --enable-prefix-caching / --no-enable-prefix-caching (vllm:v1/core/block_pool.py, kv_cache_manager.py:194-235):
 * full KV blocks stay addressable by the hash of their token prefix; a later request (or a preempted one) whose
 * tokens start the same way takes them from the pool instead of recomputing them.  The C default is off; the
 * `vllm serve` front end and the Python mirror switch it on unless told otherwise, as vLLM V1 does.  Switch before
 * the first request.  The stats are vllm:prefix_cache_queries / _hits (tokens). */
void kge_sched_set_prefix_caching(kge_sched* s, int32_t on);
void kge_engine_set_prefix_caching(kge_engine* e, int32_t on);
int32_t kge_sched_cached_tokens(kge_sched* s, int64_t request_id);     /* at its last admission; -1 = unknown id */
int kge_engine_prefix_cache_stats(const kge_engine* e, int64_t* queries, int64_t* hits);
int kge_sched_add(kge_sched* s, int64_t request_id, const int32_t* prompt_ids, int32_t n,
                  const kge_sampling_params* sp);
/* Returns the number of scheduled requests; ids/num_tokens [cap]. */
int kge_sched_schedule(kge_sched* s, int64_t* ids, int32_t* num_tokens, int32_t cap);
/* Feed sampled tokens for requests that reached their last known token this step. */
int kge_sched_update(kge_sched* s, const int64_t* ids, const int32_t* tokens, int32_t n,
                     kge_step_output* outs, int32_t cap);
int kge_sched_counts(const kge_sched* s, int32_t* running, int32_t* waiting, int32_t* free_blocks);

#ifdef __cplusplus
}
#endif
#endif /* KGE_H_ */
