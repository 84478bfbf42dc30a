// Log-probabilities of sampled tokens, vllm:v1/sample/sampler.py:300-349 (compute_logprobs + gather_logprobs):
// logprobs = log_softmax(fp32(logits)); per row the sampled token's logprob, the top-N logprobs (torch.topk order:
// value descending; equal values by ascending index here) and the sampled token's rank = #{j : x_j >= x_sampled}
// (vllm:v1/sample/ops/logprobs.py:11-27).  One CTA per row; the row (vocab x 2 B) is read 2 + N times, from L2 after
// the first pass.  Only requests that ask for logprobs pay for it.
#include <math.h>

#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

constexpr int kLpThreads = 1024;
constexpr int kLpMaxTop = 20;       // vLLM's default max_logprobs

struct MaxSum { float m, s; };
__device__ __forceinline__ MaxSum ms_merge(MaxSum a, MaxSum b) {
  if (b.m == -INFINITY) return a;
  if (a.m == -INFINITY) return b;
  const float m = fmaxf(a.m, b.m);
  return {m, a.s * expf(a.m - m) + b.s * expf(b.m - m)};
}

// (value desc, index asc) order: is (v, i) before (w, j)?
__device__ __forceinline__ bool lp_before(float v, int i, float w, int j) { return v > w || (v == w && i < j); }

__global__ void __launch_bounds__(kLpThreads)
logprobs_kernel(float* __restrict__ out_vals, int32_t* __restrict__ out_ids, int32_t* __restrict__ out_rank,
                const bf16* __restrict__ logits, const int32_t* __restrict__ sampled, int V, int ld, int top_n) {
  __shared__ float sh_m[32], sh_s[32], sh_v[32];
  __shared__ int sh_i[32], sh_c[32];
  __shared__ float bc_logz, bc_v;
  __shared__ int bc_i;
  pdl_launch_dependents();
  pdl_wait();
  const int row_id = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
  const bf16* row = logits + (int64_t)row_id * ld;
  const int tok = sampled[row_id];
  const float x_tok = (tok >= 0 && tok < V) ? bf2f(row[tok]) : -INFINITY;

  // pass 1: log-sum-exp (online, fp32) and the rank count
  MaxSum acc{-INFINITY, 0.f};
  int cnt = 0;
  for (int j = tid; j < V; j += blockDim.x) {
    const float x = bf2f(row[j]);
    cnt += x >= x_tok;
    if (x == -INFINITY) continue;                            // a masked logit adds nothing to the sum
    if (x > acc.m) { acc.s = acc.s * expf(acc.m - x) + 1.f; acc.m = x; }
    else acc.s += expf(x - acc.m);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSum other{__shfl_xor_sync(0xffffffffu, acc.m, o), __shfl_xor_sync(0xffffffffu, acc.s, o)};
    acc = ms_merge(acc, other);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) { sh_m[wid] = acc.m; sh_s[wid] = acc.s; sh_c[wid] = cnt; }
  __syncthreads();
  if (wid == 0) {
    acc = lane < nw ? MaxSum{sh_m[lane], sh_s[lane]} : MaxSum{-INFINITY, 0.f};
    cnt = lane < nw ? sh_c[lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      MaxSum other{__shfl_xor_sync(0xffffffffu, acc.m, o), __shfl_xor_sync(0xffffffffu, acc.s, o)};
      acc = ms_merge(acc, other);
      cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if (lane == 0) {
      bc_logz = acc.m + logf(acc.s);
      out_rank[row_id] = cnt;
      out_vals[(int64_t)row_id * (top_n + 1)] = x_tok - bc_logz;
      out_ids[(int64_t)row_id * (top_n + 1)] = tok;
    }
  }
  __syncthreads();
  const float logz = bc_logz;

  // passes 2..N+1: the next element in (value desc, index asc) order after the previous pick
  float last_v = INFINITY;
  int last_i = -1;
  for (int k = 1; k <= top_n; ++k) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int j = tid; j < V; j += blockDim.x) {
      const float x = bf2f(row[j]);
      if (lp_before(last_v, last_i, x, j) && lp_before(x, j, bv, bi)) { bv = x; bi = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (lp_before(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { sh_v[wid] = bv; sh_i[wid] = bi; }
    __syncthreads();
    if (wid == 0) {
      bv = lane < nw ? sh_v[lane] : -INFINITY;
      bi = lane < nw ? sh_i[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (lp_before(ov, oi, bv, bi)) { bv = ov; bi = oi; }
      }
      if (lane == 0) {
        bc_v = bv; bc_i = bi;
        const bool have = bi != 0x7fffffff;                   // fewer than N elements in the row
        out_vals[(int64_t)row_id * (top_n + 1) + k] = have ? bv - logz : -INFINITY;
        out_ids[(int64_t)row_id * (top_n + 1) + k] = have ? bi : -1;
      }
    }
    __syncthreads();
    last_v = bc_v; last_i = bc_i;
    __syncthreads();
  }
}

}  // namespace KGE_ELT_NS
}  // namespace kge

using namespace kge;

extern "C" int KGE_API(kge_logprobs)(float* out_vals, int32_t* out_ids, int32_t* out_rank, const void* logits,
                            const int32_t* sampled_ids, int R, int V, int ld, int top_n, void* stream) {
  KGE_REQUIRE(out_vals && out_ids && out_rank && logits && sampled_ids, "kge_logprobs: null pointer");
  KGE_REQUIRE(V > 0 && ld >= V, "kge_logprobs: V=%d ld=%d", V, ld);
  KGE_REQUIRE(top_n >= 0 && top_n <= kLpMaxTop, "kge_logprobs: top_n=%d must be in [0, %d]", top_n, kLpMaxTop);
  if (R <= 0) return 0;
  KGE_CUDA_OK(launch_pdl(logprobs_kernel, dim3(R), dim3(kLpThreads), 0, (cudaStream_t)stream, out_vals, out_ids, out_rank,
                         (const bf16*)logits, sampled_ids, V, ld, top_n));
  return 0;
}
