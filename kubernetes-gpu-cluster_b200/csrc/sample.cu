// Temperature / top-k / top-p sampling, one CTA per row, no sort: the two thresholds are found
// by 4-pass 8-bit radix selection over the order-preserving integer image of the fp32 logits
// (counts for top-k, softmax mass for top-p); the draw is the Gumbel-max form the reference uses,
// argmax(p_i / q_i) with q ~ Exp(1).  Rows are re-read from L2 (V*2 bytes per pass); HBM traffic
// is one read of the bf16 logits.
#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

constexpr int kSampleThreads = 1024;

__device__ __forceinline__ uint32_t fkey(float f) {   // monotone: a < b  <=>  fkey(a) < fkey(b)
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Philox-4x32-10 -> one uniform in (0,1] per (seed, step, index)
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t step, uint32_t idx) {
  uint32_t c0 = idx, c1 = (uint32_t)step, c2 = (uint32_t)(step >> 32), c3 = 0x9E3779B9u;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return ((float)(c0 >> 8) + 1.0f) * (1.0f / 16777216.0f);
}

template <typename T>
__device__ __forceinline__ T block_reduce(T v, T* sh, bool is_max) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const T other = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? (other > v ? other : v) : v + other;
  }
  __syncthreads();
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  T t = sh[0];
  for (int i = 1; i < (int)(blockDim.x >> 5); ++i) t = is_max ? (sh[i] > t ? sh[i] : t) : t + sh[i];
  return t;
}

__global__ void __launch_bounds__(kSampleThreads)
sample_kernel(int32_t* __restrict__ out, const bf16* __restrict__ logits, int V, int ld,
              const float* __restrict__ temperature, const int32_t* __restrict__ top_k,
              const float* __restrict__ top_p, const float* __restrict__ exp_noise,
              const uint64_t* __restrict__ seeds, uint64_t step) {
  __shared__ float shf[32];
  __shared__ float hist[256];
  __shared__ uint32_t s_prefix;
  __shared__ float s_below;
  __shared__ int sidx[32];
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x, tid = threadIdx.x;
  const bf16* row = logits + (int64_t)b * ld;
  const float temp = temperature ? temperature[b] : 0.f;
  const bool greedy = temp < 1e-5f;
  const float inv_t = greedy ? 1.f : 1.f / temp;

  // pass 1: max of x = logit / T
  float mx = -INFINITY;
  for (int i = tid; i < V; i += blockDim.x) mx = fmaxf(mx, bf2f(row[i]) * inv_t);
  mx = block_reduce<float>(mx, shf, true);

  uint32_t thr_key = 0;   // keep x with fkey(x) >= thr_key
  if (!greedy) {
    const int k = top_k ? top_k[b] : 0;
    if (k > 0 && k < V) {
      // radix select on counts: largest key v with count(key >= v) >= k
      uint32_t prefix = 0, mask = 0;
      int want = k;
      for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = tid; i < 256; i += blockDim.x) hist[i] = 0.f;
        __syncthreads();
        for (int i = tid; i < V; i += blockDim.x) {
          const uint32_t key = fkey(bf2f(row[i]) * inv_t);
          if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1.0f);
        }
        __syncthreads();
        if (tid == 0) {
          int acc = 0, d = 255;
          for (; d > 0; --d) {
            const int c = (int)hist[d];
            if (acc + c >= want) break;
            acc += c;
          }
          s_prefix = prefix | ((uint32_t)d << shift);
          s_below = (float)(want - acc);
        }
        __syncthreads();
        prefix = s_prefix; want = (int)s_below; mask |= 255u << shift;
        __syncthreads();
      }
      thr_key = prefix;
    }
    const float p = top_p ? top_p[b] : 1.f;
    if (p < 1.f) {
      // softmax mass over the top-k survivors, then radix select on mass from the bottom:
      // smallest key v with mass(key <= v) > (1-p) Z  (ascending cumsum <= 1-p is dropped)
      float z = 0.f;
      for (int i = tid; i < V; i += blockDim.x) {
        const float x = bf2f(row[i]) * inv_t;
        if (fkey(x) >= thr_key) z += expf(x - mx);
      }
      z = block_reduce<float>(z, shf, false);
      const float target = (1.f - p) * z;
      uint32_t prefix = 0, mask = 0;
      float below = 0.f;   // mass strictly below the current prefix bucket
      for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = tid; i < 256; i += blockDim.x) hist[i] = 0.f;
        __syncthreads();
        for (int i = tid; i < V; i += blockDim.x) {
          const float x = bf2f(row[i]) * inv_t;
          const uint32_t key = fkey(x);
          if (key >= thr_key && (key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255], expf(x - mx));
        }
        __syncthreads();
        if (tid == 0) {
          float acc = below;
          int d = 0;
          for (; d < 255; ++d) {
            if (acc + hist[d] > target) break;
            acc += hist[d];
          }
          s_prefix = prefix | ((uint32_t)d << shift);
          s_below = acc;
        }
        __syncthreads();
        prefix = s_prefix; below = s_below; mask |= 255u << shift;
        __syncthreads();
      }
      if (prefix > thr_key) thr_key = prefix;
    }
  }

  // final pass: greedy argmax, or argmax of exp(x - max) / q over the kept set
  float best = -INFINITY;
  int bi = 0x7fffffff;
  const uint32_t mx_key = fkey(mx);
  for (int i = tid; i < V; i += blockDim.x) {
    const float x = bf2f(row[i]) * inv_t;
    float score;
    if (greedy) {
      score = x;
    } else {
      const uint32_t key = fkey(x);
      if (key < thr_key && key != mx_key) continue;     // the maximum is always kept
      const float q = exp_noise ? exp_noise[(int64_t)b * V + i]
                                : -logf(philox_uniform(seeds ? seeds[b] : 0, step, (uint32_t)i));
      score = expf(x - mx) / q;
    }
    if (score > best || (score == best && i < bi)) { best = score; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  const int lane = tid & 31, wid = tid >> 5;
  __syncthreads();
  if (lane == 0) { shf[wid] = best; sidx[wid] = bi; }
  __syncthreads();
  if (tid == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); ++i)
      if (shf[i] > best || (shf[i] == best && sidx[i] < bi)) { best = shf[i]; bi = sidx[i]; }
    out[b] = bi == 0x7fffffff ? 0 : bi;
  }
}

}  // namespace KGE_ELT_NS
}  // namespace kge

using namespace kge;

#ifndef KGE_HALF
extern "C" int64_t kge_sample_workspace_bytes(int B, int V) {
  (void)B; (void)V;
  return 0;   // thresholds are found in shared memory; kept in the ABI for a future sort path
}

#endif

extern "C" int KGE_API(kge_sample)(int32_t* out_ids, const void* logits, int B, int V, int ld,
                          const float* temperature, const int32_t* top_k, const float* top_p,
                          const float* exp_noise, const uint64_t* seeds, uint64_t step, void* workspace,
                          int64_t workspace_bytes, void* stream) {
  (void)workspace; (void)workspace_bytes;
  KGE_REQUIRE(out_ids && logits && V > 0 && ld >= V, "kge_sample: bad arguments");
  if (B <= 0) return 0;
  KGE_CUDA_OK(launch_pdl(sample_kernel, dim3(B), dim3(kSampleThreads), 0, (cudaStream_t)stream, out_ids, (const bf16*)logits, V,
                         ld, temperature, top_k, top_p, exp_noise, seeds, step));
  return 0;
}
