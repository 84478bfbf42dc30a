// Causal paged attention over the engine's HND cache ([block][kv_head][slot][hd], bf16).
//
// Decode steps (one query token per request) run here; batches with longer query chunks go to the
// tcgen05 kernel in attn_tc.cu.  cp.async-staged, XOR-swizzled K/V tiles of 64 keys, fp32 online
// softmax, GQA packing (the M rows of a tile are the q heads of one kv head, so each K/V tile
// fetched from HBM serves all of them):
//   attn_decode_kernel   split-KV across CTAs, warps split the keys of a tile; partial (m, l, O)
//                        go to workspace and attn_merge_kernel combines them.
// HBM-bound at decode: algorithmic bytes = sum_b ctx_b * KV * hd * 2 (K) * 2 (V) * 2 B per layer.
#include <stdlib.h>

#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

constexpr int BN = 64;            // keys per tile
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;   // src-size 0 => zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32." KGE_MMA_ELT "." KGE_MMA_ELT ".f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// thread-block cluster helpers (split-KV partials are merged through distributed shared memory)
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
  return v;
}

// byte offset of 16-byte chunk `c` of row `r` in a [rows][HD] bf16 tile, XOR swizzled
template <int HD>
__device__ __forceinline__ uint32_t swz(int r, int c) {
  static_assert(HD % 64 == 0, "rows must span whole 128-byte lines for this swizzle");
  return (uint32_t)(r * HD * 2 + ((c ^ (r & 7)) << 4));
}

// Stage one 64-key tile of K AND V for (request, kv head) into shared memory.
// Thread t owns chunk column c = t % CH of rows t / CH + j * (NTHREADS / CH); the swizzled
// destination offsets are tile-invariant (dst[], computed once per kernel) and the source element
// offset of a row (one block-table lookup, broadcast across the lanes that share the row) is
// computed once and used for both K and V.
template <int HD, int NTHREADS>
struct KvLoader {
  static constexpr int CH = HD / 8;
  static constexpr int RP = NTHREADS / CH;       // rows per pass
  static constexpr int NP = BN / RP;             // passes per tile
  uint32_t dst[NP];
  int row0, c8;
  const int32_t* btab;
  int64_t head_off;                              // kvh * block_size * HD
  int64_t blk_stride;                            // KV * block_size * HD
  int bs_shift, bs_mask;

  __device__ __forceinline__ void init(const int32_t* bt, int kvh, int KV, int shift, int tid) {
    btab = bt; bs_shift = shift; bs_mask = (1 << shift) - 1;
    row0 = tid / CH;
    const int c = tid % CH;
    c8 = c * 8;
    head_off = ((int64_t)kvh << shift) * HD;
    blk_stride = ((int64_t)KV << shift) * HD;
#pragma unroll
    for (int j = 0; j < NP; ++j) dst[j] = swz<HD>(row0 + j * RP, c);
  }
  __device__ __forceinline__ void load(uint32_t sK, uint32_t sV, const bf16* __restrict__ kc,
                                       const bf16* __restrict__ vc, int key0, int kv_len) const {
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const int key = key0 + row0 + j * RP;
      const bool ok = key < kv_len;
      int64_t off = 0;
      if (ok) off = (int64_t)btab[key >> bs_shift] * blk_stride + head_off + (key & bs_mask) * HD + c8;
      cp_async16(sK + dst[j], kc + off, ok);
      cp_async16(sV + dst[j], vc + off, ok);
    }
  }
};

// --------------------------------------------------------------------------- decode
// grid (splits, KV, B); 128 threads.  One query token per request; the G (<=16) q heads of the
// kv head are the M rows.  Warp w handles keys 16w..16w+15 of every 64-key tile.
// Partial results: part_o [B, H, splits, HD] fp32 (un-normalised), part_ml [B, H, splits, 2].
// CLUSTER: the `splits` CTAs of one (request, kv head) form a thread-block cluster; each leaves its partial in
// its own shared memory and every rank finishes a slice of the output columns by reading its peers' partials
// over DSMEM - same arithmetic and summation order as attn_merge_kernel, no workspace, no second kernel.
template <int HD, int STAGES, bool CLUSTER>
__global__ void __launch_bounds__(128)
attn_decode_kernel(bf16* __restrict__ out, float* __restrict__ part_o, float* __restrict__ part_ml,
                   const bf16* __restrict__ q, int q_stride, const bf16* __restrict__ k_cache,
                   const bf16* __restrict__ v_cache, const int32_t* __restrict__ block_table,
                   int max_blocks, const int32_t* __restrict__ seq_lens,
                   const int32_t* __restrict__ qsl, int H, int KV, int bs_shift, float scale_log2) {
  constexpr int CH = HD / 8, NT = 128;
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.z, kvh = blockIdx.y, split = blockIdx.x, nsplit = gridDim.x;
  const int G = H / KV;
  const int kv_len = seq_lens[b];
  const int ntiles_all = (kv_len + BN - 1) / BN;
  const int tps = (ntiles_all + nsplit - 1) / nsplit;
  const int t_begin = split * tps, t_end = min(ntiles_all, t_begin + tps);
  const int32_t* btab = block_table + (int64_t)b * max_blocks;
  const int qrow = qsl[b];

  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_u32(smem);                 // 16 x HD
  const uint32_t sK = sQ + 16 * HD * 2;
  const uint32_t sV = sK + STAGES * BN * HD * 2;
  float* sRed = reinterpret_cast<float*>(smem + 16 * HD * 2);  // [4][16][HD+2], aliases the K stages after the loop
  float* sPart = reinterpret_cast<float*>(smem + 16 * HD * 2 + STAGES * BN * HD * 2);   // CLUSTER: [16][HD+2] in the V stages
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, tg = lane & 3;

  for (int i = tid; i < 16 * CH; i += NT) {
    const int r = i / CH, c = i % CH;
    const bool ok = r < G;
    const bf16* src = ok ? q + (int64_t)qrow * q_stride + (kvh * G + r) * HD + c * 8 : q;
    cp_async16(sQ + swz<HD>(r, c), src, ok);
  }
  KvLoader<HD, NT> ld;
  ld.init(btab, kvh, KV, bs_shift, tid);
  // prologue: STAGES-1 tiles in flight
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (t_begin + s < t_end)
      ld.load(sK + s * BN * HD * 2, sV + s * BN * HD * 2, k_cache, v_cache, (t_begin + s) * BN, kv_len);
    cp_async_commit();
  }

  uint32_t qf[HD / 16][4];
  float o[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  for (int kt = t_begin; kt < t_end; ++kt) {
    const int it = kt - t_begin, st = it % STAGES;
    const int nxt = kt + STAGES - 1;
    cp_async_wait<STAGES - 2>();
    __syncthreads();             // tile `it` landed for all threads; stage (it-1)%STAGES is free
    if (nxt < t_end) {
      const int ns = (it + STAGES - 1) % STAGES;
      ld.load(sK + ns * BN * HD * 2, sV + ns * BN * HD * 2, k_cache, v_cache, nxt * BN, kv_len);
    }
    cp_async_commit();
    if (it == 0) {
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk)
        ldsm_x4(qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3], sQ + swz<HD>(lane & 15, kk * 2 + (lane >> 4)));
    }
    const uint32_t kb = sK + st * BN * HD * 2, vb = sV + st * BN * HD * 2;
    float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4(b0, b1, b2, b3, kb + swz<HD>(warp * 16 + (lane & 7) + ((lane >> 4) << 3), kk * 2 + ((lane >> 3) & 1)));
      mma_bf16(s[0], qf[kk], b0, b1);
      mma_bf16(s[1], qf[kk], b2, b3);
    }
    const int key_base = kt * BN + warp * 16 + 2 * tg;
    float mx0 = m0, mx1 = m1;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k0 = key_base + j * 8;
      s[j][0] = (k0 < kv_len) ? s[j][0] * scale_log2 : -INFINITY;
      s[j][1] = (k0 + 1 < kv_len) ? s[j][1] * scale_log2 : -INFINITY;
      s[j][2] = (k0 < kv_len) ? s[j][2] * scale_log2 : -INFINITY;
      s[j][3] = (k0 + 1 < kv_len) ? s[j][3] * scale_log2 : -INFINITY;
      mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
      mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float base0 = (mx0 == -INFINITY) ? 0.f : mx0, base1 = (mx1 == -INFINITY) ? 0.f : mx1;
    const float c0 = exp2f(m0 - base0), c1 = exp2f(m1 - base1);
    m0 = mx0; m1 = mx1;
    uint32_t pf[4];
    float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float p0 = exp2f(s[j][0] - base0), p1 = exp2f(s[j][1] - base0);
      const float p2 = exp2f(s[j][2] - base1), p3 = exp2f(s[j][3] - base1);
      rs0 += p0 + p1; rs1 += p2 + p3;
      pf[j * 2 + 0] = pack_bf2(p0, p1);
      pf[j * 2 + 1] = pack_bf2(p2, p3);
    }
    l0 = l0 * c0 + rs0; l1 = l1 * c1 + rs1;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) { o[i][0] *= c0; o[i][1] *= c0; o[i][2] *= c1; o[i][3] *= c1; }
#pragma unroll
    for (int nn = 0; nn < HD / 16; ++nn) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4_t(b0, b1, b2, b3, vb + swz<HD>(warp * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), nn * 2 + (lane >> 4)));
      mma_bf16(o[2 * nn], pf, b0, b1);
      mma_bf16(o[2 * nn + 1], pf, b2, b3);
    }
  }
  cp_async_wait<0>();
  __syncthreads();               // every warp is done with the K/V stages that sRed aliases
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

  // merge the four warps' (m, l, O) through shared memory: sRed[w][row][HD] + ml at [HD], [HD+1]
  constexpr int RS = HD + 2;
  float* mine = sRed + warp * 16 * RS;
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) {
    mine[g * RS + i * 8 + 2 * tg] = o[i][0];
    mine[g * RS + i * 8 + 2 * tg + 1] = o[i][1];
    mine[(g + 8) * RS + i * 8 + 2 * tg] = o[i][2];
    mine[(g + 8) * RS + i * 8 + 2 * tg + 1] = o[i][3];
  }
  if (tg == 0) {
    mine[g * RS + HD] = m0; mine[g * RS + HD + 1] = l0;
    mine[(g + 8) * RS + HD] = m1; mine[(g + 8) * RS + HD + 1] = l1;
  }
  __syncthreads();
  // thread -> (row, column slice); rows < G only
  for (int i = tid; i < G * HD; i += NT) {
    const int r = i / HD, c = i % HD;
    float mm = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) mm = fmaxf(mm, sRed[(w * 16 + r) * RS + HD]);
    float acc = 0.f, ll = 0.f;
    const float basem = (mm == -INFINITY) ? 0.f : mm;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float sc = exp2f(sRed[(w * 16 + r) * RS + HD] - basem);
      acc += sRed[(w * 16 + r) * RS + c] * sc;
      ll += sRed[(w * 16 + r) * RS + HD + 1] * sc;
    }
    const int head = kvh * G + r;
    if (nsplit == 1) {
      out[(int64_t)qrow * H * HD + head * HD + c] = f2bf(ll > 0.f ? acc / ll : 0.f);
    } else if (CLUSTER) {
      sPart[r * RS + c] = acc;
      if (c == 0) { sPart[r * RS + HD] = mm; sPart[r * RS + HD + 1] = ll; }
    } else {
      const int64_t pi = ((int64_t)b * H + head) * nsplit + split;
      part_o[pi * HD + c] = acc;
      if (c == 0) { part_ml[pi * 2] = mm; part_ml[pi * 2 + 1] = ll; }
    }
  }
  if (CLUSTER && nsplit > 1) {
    cluster_sync_all();                                   // every rank's partial is in its shared memory
    const uint32_t sp = smem_u32(sPart);
    const int total = G * HD, per = (total + nsplit - 1) / nsplit;
    const int e_end = min(total, (split + 1) * per);
    for (int i = split * per + tid; i < e_end; i += NT) {
      const int r = i / HD, c = i % HD;
      float mm = -INFINITY;
      for (int sidx = 0; sidx < nsplit; ++sidx) mm = fmaxf(mm, ld_dsmem_f32(sp + (r * RS + HD) * 4, sidx));
      const float basem = (mm == -INFINITY) ? 0.f : mm;
      float acc = 0.f, ll = 0.f;
      for (int sidx = 0; sidx < nsplit; ++sidx) {
        const float sc = exp2f(ld_dsmem_f32(sp + (r * RS + HD) * 4, sidx) - basem);
        acc += ld_dsmem_f32(sp + (r * RS + c) * 4, sidx) * sc;
        ll += ld_dsmem_f32(sp + (r * RS + HD + 1) * 4, sidx) * sc;
      }
      out[(int64_t)qrow * H * HD + (kvh * G + r) * HD + c] = f2bf(ll > 0.f ? acc / ll : 0.f);
    }
    cluster_sync_all();                                   // nobody exits while a peer still reads its partial
  }
}

// grid (H, B); HD threads.
__global__ void attn_merge_kernel(bf16* __restrict__ out, const float* __restrict__ part_o,
                                  const float* __restrict__ part_ml, const int32_t* __restrict__ qsl,
                                  int H, int HD, int nsplit) {
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.y, head = blockIdx.x, c = threadIdx.x;
  const int64_t p0 = ((int64_t)b * H + head) * nsplit;
  float mm = -INFINITY;
  for (int s = 0; s < nsplit; ++s) mm = fmaxf(mm, part_ml[(p0 + s) * 2]);
  const float basem = (mm == -INFINITY) ? 0.f : mm;
  float acc = 0.f, ll = 0.f;
  for (int s = 0; s < nsplit; ++s) {
    const float sc = exp2f(part_ml[(p0 + s) * 2] - basem);
    acc += part_o[(p0 + s) * HD + c] * sc;
    ll += part_ml[(p0 + s) * 2 + 1] * sc;
  }
  out[(int64_t)qsl[b] * H * HD + head * HD + c] = f2bf(ll > 0.f ? acc / ll : 0.f);
}

constexpr int kDecodeStages = 3;
template <int HD> constexpr int decode_smem() {
  static_assert(4 * 16 * (HD + 2) * 4 <= kDecodeStages * BN * HD * 2, "sRed must fit in the K stages");
  return 16 * HD * 2 + 2 * kDecodeStages * BN * HD * 2;
}

int attn_decode_splits(int B, int KV, int max_seq_len);
bool decode_tma_supported(int B, int KV, int bs);          // attn_decode_tma.cu
int paged_attn_decode_tma(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt, int max_blocks,
                          const int32_t* sl, const int32_t* qsl, int B, int H, int KV, int hd, int bs, int ns, float scale_log2,
                          float* part_o, float* part_ml, void* ws, int64_t ws_bytes, int max_seq_len, cudaStream_t st);
static int decode_splits(int B, int KV, int max_seq_len) { return attn_decode_splits(B, KV, max_seq_len); }
// Split-KV factor of a decode launch; also part of the key of the engine's captured decode graphs.
int attn_decode_splits(int B, int KV, int max_seq_len) {
  const int target = 148 * 4;
  int s = (target + B * KV - 1) / (B * KV);
  int max_s = (max_seq_len + 4 * BN - 1) / (4 * BN);   // >= 256 keys per split
  if (max_s > 32) max_s = 32;
  if (max_s < 1) max_s = 1;
  if (const int forced = tunable(TUN_ATTN_DECODE_SPLITS)) return forced < max_s ? forced : max_s;
  if (s > max_s) s = max_s;
  if (s < 1) s = 1;
  if (tunable(TUN_ATTN_SPLIT_BALANCE)) {
    // Equal-length requests make equal CTAs, and the kernel is issue-bound per CTA, so the launch runs in
    // rounds of 148 x 2 resident CTAs: among s..s+3 take the factor whose last round is fullest.
    const int slots = 148 * 2;
    int best = s; double best_eff = 0;
    for (int c = s; c <= s + 3 && c <= max_s; ++c) {
      const int n = B * KV * c;
      const double eff = (double)n / (double)(((n + slots - 1) / slots) * slots);
      if (eff > best_eff + 0.02) { best_eff = eff; best = c; }
    }
    s = best;
  }
  return s;
}

template <int HD>
static int launch_attn(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc,
                       const int32_t* bt, int max_blocks, const int32_t* sl, const int32_t* qsl, int B,
                       int max_q_len, int max_seq_len, int H, int KV, int bs, float scale, void* ws,
                       int64_t ws_bytes, cudaStream_t st) {
  const float sl2 = scale * kLog2e;
  int bs_shift = 0;
  while ((1 << bs_shift) < bs) ++bs_shift;
  const int G = H / KV;
  if (max_q_len == 1 && G <= 16) {
    const int ns = decode_splits(B, KV, max_seq_len);
    if (ns > 1 && ns <= 8 && tunable(TUN_ATTN_CLUSTER_MERGE)) {
      // opt-in: split-KV CTAs of one (request, kv head) as a cluster, merged over DSMEM - no workspace and one
      // kernel, but measured SLOWER than the merge kernel behind programmatic dependent launch (C3 decode step
      // 7.64 vs 7.38 ms, profiles/r01_knob_sweep.md): cluster co-scheduling costs more than the extra launch
      auto kern = attn_decode_kernel<HD, kDecodeStages, true>;
      KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, decode_smem<HD>()));
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(ns, KV, B); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = decode_smem<HD>(); cfg.stream = st;
      cudaLaunchAttribute attr[2];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = ns; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[1].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr; cfg.numAttrs = pdl_for_this_launch() ? 2 : 1;
      KGE_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, out, (float*)nullptr, (float*)nullptr, q, q_stride, kc, vc, bt, max_blocks, sl, qsl,
                                     H, KV, bs_shift, sl2));
      return 0;
    }
    float* po = nullptr; float* pml = nullptr;
    if (ns > 1) {
      const int64_t need = (int64_t)B * H * ns * (HD + 2) * 4;
      KGE_REQUIRE(ws && ws_bytes >= need, "kge_paged_attn: workspace %lld < %lld bytes", (long long)ws_bytes, (long long)need);
      po = (float*)ws;
      pml = po + (int64_t)B * H * ns * HD;
    }
    // default: K/V tiles staged by the TMA unit, splits merged by the last CTA to finish (attn_decode_tma.cu)
    // (measured, profiles/r02_attn_block_size.jsonl: with 16-token blocks a TMA box is 2 KB and the TMA unit's request
    // rate, not HBM, bounds the kernel - 107 us against 102 us for the cp.async gathers; from 32-token blocks on the
    // boxes are large enough and TMA wins, 96 us at 64)
    if (tunable(TUN_ATTN_TMA) && bs >= 32 && decode_tma_supported(B, KV, bs))
      return paged_attn_decode_tma(out, q, q_stride, kc, vc, bt, max_blocks, sl, qsl, B, H, KV, HD, bs, ns, sl2, po, pml, ws, ws_bytes,
                                   max_seq_len, st);
    auto kern = attn_decode_kernel<HD, kDecodeStages, false>;
    KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, decode_smem<HD>()));
    KGE_CUDA_OK(launch_pdl(kern, dim3(ns, KV, B), dim3(128), decode_smem<HD>(), st, out, po, pml, q, q_stride, kc, vc, bt,
                           max_blocks, sl, qsl, H, KV, bs_shift, sl2));
    if (ns > 1)
      KGE_CUDA_OK(launch_pdl(attn_merge_kernel, dim3(H, B), dim3(HD), 0, st, out, (const float*)po, (const float*)pml, qsl, H, HD, ns));
    return 0;
  }
  kge::set_error("kge_paged_attn: internal: batch with query chunks reached the decode launcher");
  return -1;
}

int paged_attn_prefill_tc(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt,
                          int max_blocks, const int32_t* sl, const int32_t* qsl, int B, int max_q_len, int H, int KV,
                          int hd, int bs, float scale, cudaStream_t st);   // attn_tc.cu

}  // namespace KGE_ELT_NS
}  // namespace kge

using namespace kge;

#ifndef KGE_HALF
extern "C" int64_t kge_attn_workspace_bytes(int B, int H, int hd, int max_seq_len) {
  (void)max_seq_len;
  return (int64_t)B * H * 32 * (hd + 2) * 4;
}

#endif

extern "C" int KGE_API(kge_paged_attn)(void* out, const void* q, int q_stride, const void* k_cache,
                              const void* v_cache, const int32_t* block_table, int max_blocks_per_seq,
                              const int32_t* seq_lens, const int32_t* query_start_loc, int B, int T,
                              int max_q_len, int max_seq_len, int H, int KV, int hd, int block_size,
                              float scale, void* workspace, int64_t workspace_bytes, void* stream) {
  KGE_REQUIRE(out && q && k_cache && v_cache && block_table && seq_lens && query_start_loc,
              "kge_paged_attn: null pointer");
  KGE_REQUIRE(H > 0 && KV > 0 && H % KV == 0, "kge_paged_attn: H=%d must be a multiple of KV=%d", H, KV);
  KGE_REQUIRE(block_size > 0 && (block_size & (block_size - 1)) == 0, "kge_paged_attn: block_size=%d must be a power of two", block_size);
  KGE_REQUIRE(q_stride % 8 == 0, "kge_paged_attn: q_stride must be a multiple of 8");
  if (B <= 0 || T <= 0) return 0;
  KGE_REQUIRE(max_q_len >= 1 && max_seq_len >= 1, "kge_paged_attn: max_q_len/max_seq_len must be >= 1");
  cudaStream_t st = (cudaStream_t)stream;
  KGE_REQUIRE(hd == 64 || hd == 128, "kge_paged_attn: head_dim=%d unsupported (64, 128)", hd);
  if (!(max_q_len == 1 && H / KV <= 16))
    return paged_attn_prefill_tc((bf16*)out, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache, block_table,
                                 max_blocks_per_seq, seq_lens, query_start_loc, B, max_q_len, H, KV, hd, block_size, scale, st);
  if (hd == 64)
    return launch_attn<64>((bf16*)out, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache,
                           block_table, max_blocks_per_seq, seq_lens, query_start_loc, B, max_q_len,
                           max_seq_len, H, KV, block_size, scale, workspace, workspace_bytes, st);
  if (hd == 128)
    return launch_attn<128>((bf16*)out, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache,
                            block_table, max_blocks_per_seq, seq_lens, query_start_loc, B, max_q_len,
                            max_seq_len, H, KV, block_size, scale, workspace, workspace_bytes, st);
  kge::set_error("kge_paged_attn: head_dim=%d unsupported (64, 128)", hd);
  return -1;
}

#ifndef KGE_HALF      // bfloat16 only: one scratch array per library image
// ---------------------------------------------------------------------------------------------
// Decode attention for callers that must not allocate.
//
// Who this is for: a host that has handed the whole device to the engine up front - the KV pool is
// sized from gpu_memory_utilization, so a later cudaMalloc can simply fail - or one that drives
// decode steps from a latency-critical thread where cudaMalloc's implicit device-wide
// synchronisation would stall every stream.  Such a caller can still pass inputs and outputs it
// already owns, but has nowhere to get the split-KV scratch from.  Here that scratch lives in the
// library image itself: a file-scope device array that every GPU which loads libkge.so carries
// (kNoAllocScratchBytes of its memory, from load until the process exits).
//
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address - passing
// it to a kernel would fault (or, where the GPU can reach host memory, quietly scribble somewhere
// wrong).  The device address is looked up once per device with cudaGetSymbolAddress, with that
// device current, and is what the kernels receive.
constexpr int64_t kNoAllocScratchBytes = 8ll << 20;     // 8 MiB per device
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace kge {
constexpr int kMaxDevices = 64;
static void* g_noalloc_ptr[kMaxDevices];                // device address of the array, per device
static bool g_noalloc_ready[kMaxDevices];

// One-time set-up for the CURRENT device.  Not thread-safe against a concurrent first call on the
// same device; later calls only read.
static int noalloc_scratch(void** out) {
  int dev = -1;
  KGE_CUDA_OK(cudaGetDevice(&dev));
  KGE_REQUIRE(dev >= 0 && dev < kMaxDevices, "kge_paged_attn_decode_noalloc: device ordinal %d unsupported", dev);
  if (!g_noalloc_ready[dev]) {
    void* ptr = nullptr;
    cudaError_t err = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (err != cudaSuccess || ptr == nullptr) {
      kge::set_error("kge_paged_attn_decode_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                     cudaGetErrorString(err));
      return -2;
    }
    g_noalloc_ptr[dev] = ptr;
    g_noalloc_ready[dev] = true;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace kge

extern "C" int64_t kge_paged_attn_decode_noalloc_max_bytes(void) { return kNoAllocScratchBytes; }

extern "C" int kge_paged_attn_decode_noalloc(void* out, const void* q, int q_stride, const void* k_cache,
                                             const void* v_cache, const int32_t* block_table,
                                             int max_blocks_per_seq, const int32_t* seq_lens,
                                             const int32_t* query_start_loc, int B, int max_seq_len, int H,
                                             int KV, int hd, int block_size, float scale, void* stream) {
  KGE_REQUIRE(out && q && k_cache && v_cache && block_table && seq_lens && query_start_loc,
              "kge_paged_attn_decode_noalloc: null pointer");
  KGE_REQUIRE(H > 0 && KV > 0 && H % KV == 0 && H / KV <= 16,
              "kge_paged_attn_decode_noalloc: H=%d KV=%d (group must be <= 16)", H, KV);
  KGE_REQUIRE(hd == 64 || hd == 128, "kge_paged_attn_decode_noalloc: head_dim=%d unsupported (64, 128)", hd);
  KGE_REQUIRE(block_size > 0 && (block_size & (block_size - 1)) == 0,
              "kge_paged_attn_decode_noalloc: block_size=%d must be a power of two", block_size);
  KGE_REQUIRE(q_stride % 8 == 0, "kge_paged_attn_decode_noalloc: q_stride must be a multiple of 8");
  if (B <= 0) return 0;
  KGE_REQUIRE(max_seq_len >= 1, "kge_paged_attn_decode_noalloc: max_seq_len must be >= 1");
  // same split choice as kge_paged_attn; the partials of all splits must fit the fixed array
  const int ns = decode_splits(B, KV, max_seq_len);
  const int64_t need = ns > 1 ? (int64_t)B * H * ns * (hd + 2) * 4 : 0;
  KGE_REQUIRE(need <= kNoAllocScratchBytes,
              "kge_paged_attn_decode_noalloc: this batch needs %lld bytes of split-KV scratch but the built-in array "
              "holds %lld (B=%d H=%d splits=%d head_dim=%d); use kge_paged_attn with a caller workspace",
              (long long)need, (long long)kNoAllocScratchBytes, B, H, ns, hd);
  void* ws = nullptr;
  if (int rc = noalloc_scratch(&ws)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // every (request, head, split) slot of the scratch is written by this launch before the merge
  // kernel reads it, so nothing has to be cleared between calls
  if (hd == 64)
    return launch_attn<64>((bf16*)out, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache, block_table,
                           max_blocks_per_seq, seq_lens, query_start_loc, B, 1, max_seq_len, H, KV, block_size, scale,
                           ws, kNoAllocScratchBytes, st);
  return launch_attn<128>((bf16*)out, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache, block_table,
                          max_blocks_per_seq, seq_lens, query_start_loc, B, 1, max_seq_len, H, KV, block_size, scale,
                          ws, kNoAllocScratchBytes, st);
}
#endif
