// HBM-bound row kernels: embedding gather, (add+)RMSNorm, LayerNorm, SiLU*mul, RoPE+KV-write,
// greedy argmax.  All use 128-bit accesses (8 bf16 per thread per access); rows stay in
// registers between the reduction and the store so every input byte is read once.
#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

constexpr int kMaxVecPerThread = 4;   // up to 4 x (8 bf16) per thread: d <= 32 * blockDim

static inline int row_threads(int d) {
  int v = d / 8;
  int t = 32;
  while (t < v && t < 256) t <<= 1;
  while (t < 512 && (v + t - 1) / t > kMaxVecPerThread) t <<= 1;
  return t;
}

// ------------------------------------------------------------------------ embedding
__global__ void embed_kernel(bf16* __restrict__ out, const int32_t* __restrict__ ids,
                             const bf16* __restrict__ table, const bf16* __restrict__ pos_table,
                             const int32_t* __restrict__ positions, int pos_offset, int d) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const int64_t row = ids[t];
  const uint4* src = reinterpret_cast<const uint4*>(table + row * d);
  const uint4* psrc = nullptr;
  if (pos_table) psrc = reinterpret_cast<const uint4*>(pos_table + (int64_t)(positions[t] + pos_offset) * d);
  uint4* dst = reinterpret_cast<uint4*>(out + (int64_t)t * d);
  for (int i = threadIdx.x; i < d / 8; i += blockDim.x) {
    uint4 a = ld_nc_v4(src + i);
    if (psrc) {
      uint4 p = ld_nc_v4(psrc + i);
      a.x = pack_bf2(bf_lo(a.x) + bf_lo(p.x), bf_hi(a.x) + bf_hi(p.x));
      a.y = pack_bf2(bf_lo(a.y) + bf_lo(p.y), bf_hi(a.y) + bf_hi(p.y));
      a.z = pack_bf2(bf_lo(a.z) + bf_lo(p.z), bf_hi(a.z) + bf_hi(p.z));
      a.w = pack_bf2(bf_lo(a.w) + bf_lo(p.w), bf_hi(a.w) + bf_hi(p.w));
    }
    dst[i] = a;
  }
}

// ------------------------------------------------------------------------ norms
// MODE 0: rmsnorm(x)            MODE 1: s=x+res (fp32), res<-bf16(s), rmsnorm(s)
// MODE 2: layernorm(res)        MODE 3: h=bf16(res+add), res<-h, layernorm(h)  (y may be null)
template <int MODE>
__global__ void __launch_bounds__(512)
norm_kernel(bf16* __restrict__ y, bf16* __restrict__ residual, const bf16* __restrict__ x,
            const bf16* __restrict__ w, const bf16* __restrict__ b, int d, float eps) {
  __shared__ float red[33];
  pdl_launch_dependents();
  pdl_wait();
  const int64_t base = (int64_t)blockIdx.x * d;
  const int nvec = d / 8;
  float v[kMaxVecPerThread][8];
  float acc = 0.f;
#pragma unroll
  for (int j = 0; j < kMaxVecPerThread; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < nvec) {
      uint4 a;
      if (MODE == 0 || MODE == 1) a = *reinterpret_cast<const uint4*>(x + base + i * 8);
      else a = *reinterpret_cast<const uint4*>(residual + base + i * 8);
      float* f = v[j];
      f[0] = bf_lo(a.x); f[1] = bf_hi(a.x); f[2] = bf_lo(a.y); f[3] = bf_hi(a.y);
      f[4] = bf_lo(a.z); f[5] = bf_hi(a.z); f[6] = bf_lo(a.w); f[7] = bf_hi(a.w);
      if (MODE == 1 || MODE == 3) {
        uint4 r;
        if (MODE == 1) r = *reinterpret_cast<const uint4*>(residual + base + i * 8);
        else r = *reinterpret_cast<const uint4*>(x + base + i * 8);   // `x` carries `add`
        f[0] += bf_lo(r.x); f[1] += bf_hi(r.x); f[2] += bf_lo(r.y); f[3] += bf_hi(r.y);
        f[4] += bf_lo(r.z); f[5] += bf_hi(r.z); f[6] += bf_lo(r.w); f[7] += bf_hi(r.w);
        uint4 o;
        o.x = pack_bf2(f[0], f[1]); o.y = pack_bf2(f[2], f[3]);
        o.z = pack_bf2(f[4], f[5]); o.w = pack_bf2(f[6], f[7]);
        *reinterpret_cast<uint4*>(residual + base + i * 8) = o;
        if (MODE == 3) {   // OPT normalises the ROUNDED sum (opt.py:181-186)
          f[0] = bf_lo(o.x); f[1] = bf_hi(o.x); f[2] = bf_lo(o.y); f[3] = bf_hi(o.y);
          f[4] = bf_lo(o.z); f[5] = bf_hi(o.z); f[6] = bf_lo(o.w); f[7] = bf_hi(o.w);
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) acc += (MODE <= 1) ? f[k] * f[k] : f[k];
    }
  }
  if (MODE == 3 && y == nullptr) return;
  float tot = block_sum(acc, red);
  float mean = 0.f, rstd;
  if (MODE <= 1) {
    rstd = 1.0f / sqrtf(tot / (float)d + eps);
  } else {
    mean = tot / (float)d;
    float a2 = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxVecPerThread; ++j) {
      const int i = threadIdx.x + j * blockDim.x;
      if (i < nvec) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { float c = v[j][k] - mean; a2 += c * c; }
      }
    }
    float var = block_sum(a2, red) / (float)d;
    rstd = 1.0f / sqrtf(var + eps);
  }
#pragma unroll
  for (int j = 0; j < kMaxVecPerThread; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < nvec) {
      uint4 ww = *reinterpret_cast<const uint4*>(w + i * 8);
      float wf[8] = {bf_lo(ww.x), bf_hi(ww.x), bf_lo(ww.y), bf_hi(ww.y),
                     bf_lo(ww.z), bf_hi(ww.z), bf_lo(ww.w), bf_hi(ww.w)};
      float o[8];
      if (MODE <= 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rbf(v[j][k] * rstd) * wf[k];   // weight mul in bf16
      } else {
        uint4 bb = *reinterpret_cast<const uint4*>(b + i * 8);
        float bfv[8] = {bf_lo(bb.x), bf_hi(bb.x), bf_lo(bb.y), bf_hi(bb.y),
                        bf_lo(bb.z), bf_hi(bb.z), bf_lo(bb.w), bf_hi(bb.w)};
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = (v[j][k] - mean) * rstd * wf[k] + bfv[k];
      }
      uint4 ov;
      ov.x = pack_bf2(o[0], o[1]); ov.y = pack_bf2(o[2], o[3]);
      ov.z = pack_bf2(o[4], o[5]); ov.w = pack_bf2(o[6], o[7]);
      *reinterpret_cast<uint4*>(y + base + i * 8) = ov;
    }
  }
}

// ------------------------------------------------------------------------ silu * mul
__device__ __forceinline__ float silu_bf(float g) {
  // F.silu on a bf16 tensor: fp32 x / (1 + exp(-x)), rounded to bf16
  return rbf(g / (1.0f + expf(-g)));
}
__global__ void silu_mul_kernel(bf16* __restrict__ y, const bf16* __restrict__ x, int f, int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  const int fv = f / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / fv;
    const int c = (int)(i % fv);
    const uint4 g = ld_nc_v4(x + t * 2 * f + c * 8);
    const uint4 u = ld_nc_v4(x + t * 2 * f + f + c * 8);
    uint4 o;
    o.x = pack_bf2(silu_bf(bf_lo(g.x)) * bf_lo(u.x), silu_bf(bf_hi(g.x)) * bf_hi(u.x));
    o.y = pack_bf2(silu_bf(bf_lo(g.y)) * bf_lo(u.y), silu_bf(bf_hi(g.y)) * bf_hi(u.y));
    o.z = pack_bf2(silu_bf(bf_lo(g.z)) * bf_lo(u.z), silu_bf(bf_hi(g.z)) * bf_hi(u.z));
    o.w = pack_bf2(silu_bf(bf_lo(g.w)) * bf_lo(u.w), silu_bf(bf_hi(g.w)) * bf_hi(u.w));
    *reinterpret_cast<uint4*>(y + t * f + c * 8) = o;
  }
}

// ------------------------------------------------------------------------ rope + kv write
// One CTA per token.  Work items are 8-element vectors of the first half of every q/k head
// (rotated together with their partner in the second half) and of every v head.
__global__ void rope_kvwrite_kernel(bf16* __restrict__ qkv, const bf16* __restrict__ cos_sin,
                                    const int32_t* __restrict__ positions,
                                    const int64_t* __restrict__ slot_mapping,
                                    bf16* __restrict__ k_cache, bf16* __restrict__ v_cache,
                                    int H, int KV, int hd, int block_size) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const int half = hd / 2, hv = half / 8;          // vectors per half head
  bf16* row = qkv + (int64_t)t * (H + 2 * KV) * hd;
  const int64_t slot = slot_mapping[t];
  const int64_t blk = slot >= 0 ? slot / block_size : 0;
  const int off = slot >= 0 ? (int)(slot % block_size) : 0;
  const bf16* cs = cos_sin ? cos_sin + (int64_t)positions[t] * hd : nullptr;
  const int n_rot = (H + KV) * hv;                 // q heads then k heads
  const int n_v = KV * (hd / 8);
  for (int w = threadIdx.x; w < n_rot + n_v; w += blockDim.x) {
    if (w < n_rot) {
      const int head = w / hv, c = w % hv;
      bf16* p1 = row + head * hd + c * 8;
      bf16* p2 = p1 + half;
      uint4 a = *reinterpret_cast<uint4*>(p1), b = *reinterpret_cast<uint4*>(p2);
      if (cs) {
        const uint4 cv = *reinterpret_cast<const uint4*>(cs + c * 8);
        const uint4 sv = *reinterpret_cast<const uint4*>(cs + half + c * 8);
        const uint32_t* au = &a.x; const uint32_t* bu = &b.x;
        const uint32_t* cu = &cv.x; const uint32_t* su = &sv.x;
        uint32_t o1[4], o2[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float x1l = bf_lo(au[k]), x1h = bf_hi(au[k]), x2l = bf_lo(bu[k]), x2h = bf_hi(bu[k]);
          float cl = bf_lo(cu[k]), ch = bf_hi(cu[k]), sl = bf_lo(su[k]), sh = bf_hi(su[k]);
          // o1 = x1*cos - x2*sin ; o2 = x2*cos + x1*sin ; every op rounds to bf16
          o1[k] = pack_bf2(rbf(x1l * cl) - rbf(x2l * sl), rbf(x1h * ch) - rbf(x2h * sh));
          o2[k] = pack_bf2(rbf(x2l * cl) + rbf(x1l * sl), rbf(x2h * ch) + rbf(x1h * sh));
        }
        a = make_uint4(o1[0], o1[1], o1[2], o1[3]);
        b = make_uint4(o2[0], o2[1], o2[2], o2[3]);
        *reinterpret_cast<uint4*>(p1) = a;
        *reinterpret_cast<uint4*>(p2) = b;
      }
      if (head >= H && slot >= 0) {
        const int kh = head - H;
        bf16* dst = k_cache + ((blk * KV + kh) * block_size + off) * (int64_t)hd;
        *reinterpret_cast<uint4*>(dst + c * 8) = a;
        *reinterpret_cast<uint4*>(dst + half + c * 8) = b;
      }
    } else if (slot >= 0) {
      const int u = w - n_rot;
      const int kh = u / (hd / 8), c = u % (hd / 8);
      const uint4 a = *reinterpret_cast<const uint4*>(row + (H + KV + kh) * hd + c * 8);
      bf16* dst = v_cache + ((blk * KV + kh) * block_size + off) * (int64_t)hd;
      *reinterpret_cast<uint4*>(dst + c * 8) = a;
    }
  }
}

// ------------------------------------------------------------------------ per-head RMSNorm (Qwen3)
// One warp per (token, q or k head): rms_norm over head_dim with vLLM's cast points, in place.
__global__ void head_rmsnorm_kernel(bf16* __restrict__ qkv, const bf16* __restrict__ qw, const bf16* __restrict__ kw,
                                    int T, int H, int KV, int hd, float eps) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= (int64_t)T * (H + KV)) return;
  const int t = (int)(item / (H + KV)), head = (int)(item % (H + KV));
  bf16* row = qkv + (int64_t)t * (H + 2 * KV) * hd + head * hd;
  const bf16* w = head < H ? qw : kw;
  const int per = hd / 32;                          // 2 (hd 64) or 4 (hd 128) elements per lane
  float v[4], ss = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (i < per) { v[i] = bf2f(row[lane * per + i]); ss += v[i] * v[i]; }
  ss = warp_sum(ss);
  const float rstd = 1.0f / sqrtf(ss / (float)hd + eps);
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (i < per) row[lane * per + i] = f2bf(rbf(v[i] * rstd) * bf2f(w[lane * per + i]));
}

// ------------------------------------------------------------------------ greedy argmax
// One CTA per row; 128-bit loads; ties resolve to the lowest index (torch.argmax semantics).
__global__ void __launch_bounds__(1024)
argmax_kernel(int32_t* __restrict__ out, const bf16* __restrict__ logits, int V, int ld) {
  __shared__ float sval[32];
  __shared__ int sidx[32];
  pdl_launch_dependents();
  pdl_wait();
  const bf16* row = logits + (int64_t)blockIdx.x * ld;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  const int nvec = V / 8;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 a = ld_nc_v4(row + i * 8);
    const float f[8] = {bf_lo(a.x), bf_hi(a.x), bf_lo(a.y), bf_hi(a.y),
                        bf_lo(a.z), bf_hi(a.z), bf_lo(a.w), bf_hi(a.w)};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      // NaN is treated as the maximum by torch.argmax; logits here are finite by construction
      if (f[k] > best) { best = f[k]; bi = i * 8 + k; }
    }
  }
  for (int i = nvec * 8 + threadIdx.x; i < V; i += blockDim.x) {
    const float f = bf2f(row[i]);
    if (f > best || (f == best && i < bi)) { best = f; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sval[wid] = best; sidx[wid] = bi; }
  __syncthreads();
  if (wid == 0) {
    const int nw = blockDim.x >> 5;
    best = lane < nw ? sval[lane] : -INFINITY;
    bi = lane < nw ? sidx[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (lane == 0) out[blockIdx.x] = (bi == 0x7fffffff) ? 0 : bi;
  }
}

}  // namespace KGE_ELT_NS
}  // namespace kge

using namespace kge;

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int KGE_API(kge_embed)(void* out, const int32_t* ids, const void* table, const void* pos_table,
                         const int32_t* positions, int pos_offset, int T, int d, void* stream) {
  KGE_REQUIRE(T >= 0 && d > 0 && d % 8 == 0, "kge_embed: d=%d must be a positive multiple of 8", d);
  KGE_REQUIRE(aligned16(out) && aligned16(table) && aligned16(pos_table), "kge_embed: pointers must be 16-byte aligned");
  KGE_REQUIRE(!pos_table || positions, "kge_embed: pos_table given without positions");
  if (T == 0) return 0;
  int th = d / 8 < 256 ? ((d / 8 + 31) / 32) * 32 : 256;
  KGE_CUDA_OK(launch_pdl(embed_kernel, dim3(T), dim3(th), 0, (cudaStream_t)stream, (bf16*)out, ids, (const bf16*)table,
                         (const bf16*)pos_table, positions, pos_offset, d));
  return 0;
}

static int norm_common(int mode, void* y, void* residual, const void* x, const void* w,
                       const void* b, int T, int d, float eps, void* stream) {
  KGE_REQUIRE(T >= 0 && d > 0 && d % 8 == 0 && d <= 16384,
              "kge norm: d=%d must be a multiple of 8 and <= 16384", d);
  KGE_REQUIRE(aligned16(y) && aligned16(residual) && aligned16(x) && aligned16(w) && aligned16(b),
              "kge norm: pointers must be 16-byte aligned");
  if (T == 0) return 0;
  const int th = row_threads(d);
  KGE_REQUIRE((d / 8 + th - 1) / th <= kMaxVecPerThread, "kge norm: d=%d too large", d);
  cudaStream_t s = (cudaStream_t)stream;
  auto kern = mode == 0 ? norm_kernel<0> : mode == 1 ? norm_kernel<1> : mode == 2 ? norm_kernel<2> : norm_kernel<3>;
  KGE_CUDA_OK(launch_pdl(kern, dim3(T), dim3(th), 0, s, (bf16*)y, (bf16*)residual, (const bf16*)x, (const bf16*)w,
                         (const bf16*)b, d, eps));
  return 0;
}

extern "C" int KGE_API(kge_rmsnorm)(void* y, const void* x, const void* w, int T, int d, float eps, void* stream) {
  KGE_REQUIRE(y && x && w, "kge_rmsnorm: null pointer");
  return norm_common(0, y, nullptr, x, w, nullptr, T, d, eps, stream);
}
extern "C" int KGE_API(kge_add_rmsnorm)(void* y, void* residual, const void* x, const void* w, int T, int d,
                               float eps, void* stream) {
  KGE_REQUIRE(y && residual && x && w, "kge_add_rmsnorm: null pointer");
  return norm_common(1, y, residual, x, w, nullptr, T, d, eps, stream);
}
extern "C" int KGE_API(kge_layernorm)(void* y, void* residual, const void* add, const void* w, const void* b,
                             int T, int d, float eps, void* stream) {
  KGE_REQUIRE(residual && (y || add), "kge_layernorm: null pointer");
  KGE_REQUIRE(!y || (w && b), "kge_layernorm: weight and bias required");
  return norm_common(add ? 3 : 2, y, residual, add, w, b, T, d, eps, stream);
}

extern "C" int KGE_API(kge_silu_mul)(void* y, const void* x, int T, int f, void* stream) {
  KGE_REQUIRE(y && x && f > 0 && f % 8 == 0, "kge_silu_mul: f=%d must be a positive multiple of 8", f);
  KGE_REQUIRE(aligned16(y) && aligned16(x), "kge_silu_mul: pointers must be 16-byte aligned");
  if (T <= 0) return 0;
  const int64_t nvec = (int64_t)T * (f / 8);
  int64_t blocks = (nvec + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  KGE_CUDA_OK(launch_pdl(silu_mul_kernel, dim3((int)blocks), dim3(256), 0, (cudaStream_t)stream, (bf16*)y, (const bf16*)x, f, nvec));
  return 0;
}

extern "C" int KGE_API(kge_rope_kvwrite)(void* qkv, const void* cos_sin, const int32_t* positions,
                                const int64_t* slot_mapping, void* k_cache, void* v_cache, int T,
                                int H, int KV, int hd, int block_size, void* stream) {
  KGE_REQUIRE(qkv && slot_mapping && k_cache && v_cache, "kge_rope_kvwrite: null pointer");
  KGE_REQUIRE(hd % 16 == 0 && hd <= 256, "kge_rope_kvwrite: head_dim=%d must be a multiple of 16", hd);
  KGE_REQUIRE(!cos_sin || positions, "kge_rope_kvwrite: cos_sin given without positions");
  KGE_REQUIRE(aligned16(qkv) && aligned16(cos_sin) && aligned16(k_cache) && aligned16(v_cache),
              "kge_rope_kvwrite: pointers must be 16-byte aligned");
  if (T <= 0) return 0;
  const int items = (H + KV) * (hd / 16) + KV * (hd / 8);
  int th = ((items + 31) / 32) * 32;
  if (th > 256) th = 256;
  KGE_CUDA_OK(launch_pdl(rope_kvwrite_kernel, dim3(T), dim3(th), 0, (cudaStream_t)stream, (bf16*)qkv, (const bf16*)cos_sin,
                         positions, slot_mapping, (bf16*)k_cache, (bf16*)v_cache, H, KV, hd, block_size));
  return 0;
}

extern "C" int KGE_API(kge_head_rmsnorm)(void* qkv, const void* q_weight, const void* k_weight, int T, int H, int KV, int hd,
                                float eps, void* stream) {
  KGE_REQUIRE(qkv && q_weight && k_weight, "kge_head_rmsnorm: null pointer");
  KGE_REQUIRE(hd == 64 || hd == 128, "kge_head_rmsnorm: head_dim=%d unsupported (64, 128)", hd);
  if (T <= 0) return 0;
  const int64_t items = (int64_t)T * (H + KV);
  KGE_CUDA_OK(launch_pdl(head_rmsnorm_kernel, dim3((unsigned)((items + 7) / 8)), dim3(256), 0, (cudaStream_t)stream, (bf16*)qkv,
                         (const bf16*)q_weight, (const bf16*)k_weight, T, H, KV, hd, eps));
  return 0;
}

extern "C" int KGE_API(kge_argmax)(int32_t* out_ids, const void* logits, int B, int V, int ld, void* stream) {
  KGE_REQUIRE(out_ids && logits && V > 0 && ld >= V, "kge_argmax: bad arguments");
  KGE_REQUIRE(aligned16(logits) && ld % 8 == 0, "kge_argmax: logits rows must be 16-byte aligned");
  if (B <= 0) return 0;
  KGE_CUDA_OK(launch_pdl(argmax_kernel, dim3(B), dim3(1024), 0, (cudaStream_t)stream, out_ids, (const bf16*)logits, V, ld));
  return 0;
}
