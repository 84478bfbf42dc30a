// Decode attention (one query token per request) with TMA-staged K/V tiles.
//
// Same arithmetic as attn_decode_kernel in attn.cu (64-key tiles, the G <= 16 q heads of one kv head as the M rows of
// mma.sync.m16n8k16, fp32 online softmax in the exp2 domain, four warps taking 16 keys each of a tile, split-KV across
// CTAs) - what changes is how the bytes arrive and how the splits are merged:
//   - A (block, kv head) run of the HND cache is `block_size` contiguous rows of head_dim bf16.  The cache of one layer is
//     described to the TMA unit as a 2-D tensor [rows, head_dim] and ONE producer thread fetches every tile with
//     cp.async.bulk.tensor.2d boxes of {64 columns, block_size rows} (128-byte swizzle), completion counted by an
//     mbarrier (expect_tx).  The four compute warps issue no loads and no address arithmetic at all: they wait on the
//     stage's "full" barrier, read the swizzled tile with ldmatrix and release the stage with one arrive per warp.
//     (The cp.async version spent 42 % of its issue slots, mostly on per-16-byte address generation, and reached
//     84 % of HBM peak; profiles/r01_ncu_summary.md.)
//   - The split-KV partials of a (request, kv head) are merged by whichever of its CTAs finishes LAST (arrival counter
//     after __threadfence), in split order - the same sums as attn_merge_kernel, without a second launch.
// Keys past the end of a sequence: the producer always fetches whole tiles; block-table entries past the table are read
// as block 0.  Those keys are masked to -inf before the softmax, so their P is exactly 0; their V rows only have to be
// finite (the engine keeps block 0 zero and never hands out uninitialised blocks).
// HBM-bound: algorithmic bytes = sum_b ctx_b * KV * hd * 2 (K) * 2 (V) * 2 B per layer.
#include <cuda.h>

#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

int make_tmap(CUtensorMap* m, const void* ptr, int64_t rows, int64_t cols, int box_rows);   // gemm_tc.cu (cached)

namespace adt {

constexpr int BN = 64;                       // keys per tile
constexpr int kStages = 3;
constexpr uint32_t kSpin = 1u << 26;
constexpr int kMaxGroups = 1 << 16;          // (request, kv head) pairs one launch may have

// arrival counters of the split-KV merge, one per (request, kv head); zero at load, reset by the CTA that merges
__device__ unsigned int g_split_arrivals[kMaxGroups];

__device__ __forceinline__ uint32_t s_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok, spins = 0;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (!ok && ++spins > kSpin) __trap();
  } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, uint64_t pol) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32." KGE_MMA_ELT "." KGE_MMA_ELT ".f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void bar_compute() { asm volatile("bar.sync 1, 128;" ::: "memory"); }   // the four compute warps

// Q tile [16][HD] as the compute warps store it: 16-byte chunks XOR-swizzled by row % 8
template <int HD>
__device__ __forceinline__ uint32_t swz_q(int r, int c) { return (uint32_t)(r * HD * 2 + ((c ^ (r & 7)) << 4)); }
// K / V tile as TMA writes it: [64-column half][64 keys][128 B], chunks XOR-swizzled by key % 8 (SWIZZLE_128B)
__device__ __forceinline__ uint32_t swz_kv(int key, int c /* 16-byte chunk of the full row */) {
  return (uint32_t)((c >> 3) * (BN * 128) + key * 128 + (((c & 7) ^ (key & 7)) << 4));
}

// grid (splits, KV, B); 160 threads: warps 0-3 compute, warp 4 = TMA producer.
template <int HD>
__global__ void __launch_bounds__(160, 2)
attn_decode_tma_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                       bf16* __restrict__ out, float* __restrict__ part_o, float* __restrict__ part_ml,
                       const bf16* __restrict__ q, int q_stride, const int32_t* __restrict__ block_table, int max_blocks,
                       const int32_t* __restrict__ seq_lens, const int32_t* __restrict__ qsl, int H, int KV, int bs_shift,
                       float scale_log2) {
  constexpr int CH = HD / 8;
  constexpr uint32_t TILE = BN * HD * 2;                  // one K or V tile
  constexpr uint32_t Q_BYTES = 16 * HD * 2;
  static_assert(Q_BYTES % 1024 == 0 && TILE % 1024 == 0, "swizzled TMA boxes need 1024-byte aligned stage bases");
  const int b = blockIdx.z, kvh = blockIdx.y, split = blockIdx.x, nsplit = gridDim.x;
  const int G = H / KV;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sQ = s_addr(smem);
  const uint32_t sK = sQ + Q_BYTES;
  const uint32_t sV = sK + kStages * TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Q_BYTES + 2 * kStages * TILE);
  const uint32_t bar0 = s_addr(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (kStages + s); };
  float* sRed = reinterpret_cast<float*>(smem + Q_BYTES);     // [4][16][HD+2], aliases the K stages after the loop
  volatile int* s_last = reinterpret_cast<volatile int*>(bars + 2 * kStages);   // (no static shared memory: the dynamic window must start 1024-aligned)

  if (tid == 0) {
    if ((sQ & 1023u) != 0) __trap();
    for (int s = 0; s < kStages; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
  }
  __syncthreads();                              // barriers initialised (before anything of the previous kernel is needed)
  pdl_launch_dependents();
  pdl_wait();                                   // q, the KV rows of this step and the metadata come from earlier kernels
  const int kv_len = seq_lens[b];
  const int ntiles_all = (kv_len + BN - 1) / BN;
  const int tps = (ntiles_all + nsplit - 1) / nsplit;
  const int t_begin = split * tps, t_end = min(ntiles_all, t_begin + tps);
  const int qrow = qsl[b];
  if (warp < 4) {
    // the Q tile arrives while the producer warp already has the first K/V tiles in flight
    for (int i = tid; i < 16 * CH; i += 128) {
      const int r = i / CH, c = i % CH;
      const bool ok = r < G;
      const bf16* src = ok ? q + (int64_t)qrow * q_stride + (kvh * G + r) * HD + c * 8 : q;
      cp_async16(sQ + swz_q<HD>(r, c), src, ok);
    }
    asm volatile("cp.async.commit_group;");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    bar_compute();                              // Q tile visible to every compute warp
  }

  float o[HD / 8][4];
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const int g = lane >> 2, tg = lane & 3;

  if (warp == 4) {
    // ---------------------------------------------------------------- producer (the whole warp)
    // Block ids are read 32 at a time, one per lane (a coalesced 128-byte load, the next 32 already in flight), and
    // handed out by shuffle; lane j issues the boxes of block j of the tile.  A single thread that loads one table entry,
    // waits for it and issues one block's boxes, eight times per tile, cannot keep three stages full.
    const int32_t* btab = block_table + (int64_t)b * max_blocks;
    const int bs = 1 << bs_shift;
    const int box_rows = bs < BN ? bs : BN;
    const int boxes = BN / box_rows;
    auto ld_blk = [&](int bi) { return (bi >= 0 && bi < max_blocks) ? btab[bi] : 0; };   // past the table: block 0 (masked keys)
    int cbase = (t_begin * BN) >> bs_shift;
    int v = ld_blk(cbase + lane), v_next = ld_blk(cbase + 32 + lane);
    for (int kt = t_begin; kt < t_end; ++kt) {
      const int it = kt - t_begin, st = it % kStages;
      const int key = kt * BN + (lane < boxes ? lane : 0) * box_rows;      // first key of this lane's box
      const int bi = key >> bs_shift;
      const int bi_last = (kt * BN + (boxes - 1) * box_rows) >> bs_shift;
      if (bi_last >= cbase + 32) {                                         // warp-uniform: the tile leaves the held chunk
        const int nb = (kt * BN) >> bs_shift;
        if (nb == cbase + 32) { v = v_next; } else { v = ld_blk(nb + lane); }
        cbase = nb;
        v_next = ld_blk(cbase + 32 + lane);
      }
      const int blk = __shfl_sync(0xffffffffu, v, (bi - cbase) & 31);
      if (lane == 0) {
        if (it >= kStages) mbar_wait(empty(st), (uint32_t)((it / kStages - 1) & 1));
        mbar_expect_tx(full(st), 2 * TILE);
      }
      __syncwarp();
      if (lane < boxes) {
        const int row = (blk * KV + kvh) * bs + (key & (bs - 1));
#pragma unroll
        for (int h = 0; h < HD / 64; ++h) {
          const uint32_t off = (uint32_t)(h * (BN * 128) + lane * box_rows * 128);
          tma_load_2d(sK + st * TILE + off, &tm_k, full(st), h * 64, row);
          tma_load_2d(sV + st * TILE + off, &tm_v, full(st), h * 64, row);
        }
      }
    }
  } else {
    // ---------------------------------------------------------------- compute
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk)
      ldsm_x4(qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3], sQ + swz_q<HD>(lane & 15, kk * 2 + (lane >> 4)));
    for (int kt = t_begin; kt < t_end; ++kt) {
      const int it = kt - t_begin, st = it % kStages;
      mbar_wait(full(st), (uint32_t)((it / kStages) & 1));
      const uint32_t kb = sK + st * TILE, vb = sV + st * TILE;
      float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(b0, b1, b2, b3, kb + swz_kv(warp * 16 + (lane & 7) + ((lane >> 4) << 3), kk * 2 + ((lane >> 3) & 1)));
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
        ldsm_x4_t(b0, b1, b2, b3, vb + swz_kv(warp * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), nn * 2 + (lane >> 4)));
        mma_bf16(o[2 * nn], pf, b0, b1);
        mma_bf16(o[2 * nn + 1], pf, b2, b3);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty(st));      // this warp has read its keys of the stage
    }
    bar_compute();                                // every compute warp is done with the stages that sRed aliases
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
    bar_compute();
    for (int i = tid; i < G * HD; i += 128) {
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
      } else {
        const int64_t pi = ((int64_t)b * H + head) * nsplit + split;
        part_o[pi * HD + c] = acc;
        if (c == 0) { part_ml[pi * 2] = mm; part_ml[pi * 2 + 1] = ll; }
      }
    }
    if (nsplit > 1) {
      // the CTA that arrives last merges every split of this (request, kv head), in split order
      // (one thread publishes for the CTA: the barrier orders the others' stores before its fence, as a grid-wide
      // barrier does - 128 fences while the producer's next tiles are in flight each cost a memory round trip)
      bar_compute();
      if (tid == 0) {
        __threadfence();
        const unsigned prev = atomicAdd(&g_split_arrivals[b * KV + kvh], 1u);
        __threadfence();
        *s_last = (prev == (unsigned)nsplit - 1u) ? 1 : 0;
      }
      bar_compute();
      if (*s_last) {
        for (int i = tid; i < G * HD; i += 128) {
          const int r = i / HD, c = i % HD;
          const int64_t p0 = ((int64_t)b * H + kvh * G + r) * nsplit;
          float mm = -INFINITY;
          for (int sidx = 0; sidx < nsplit; ++sidx) mm = fmaxf(mm, __ldcg(part_ml + (p0 + sidx) * 2));
          const float basem = (mm == -INFINITY) ? 0.f : mm;
          float acc = 0.f, ll = 0.f;
          for (int sidx = 0; sidx < nsplit; ++sidx) {
            const float sc = exp2f(__ldcg(part_ml + (p0 + sidx) * 2) - basem);
            acc += __ldcg(part_o + (p0 + sidx) * HD + c) * sc;
            ll += __ldcg(part_ml + (p0 + sidx) * 2 + 1) * sc;
          }
          out[(int64_t)qrow * H * HD + (kvh * G + r) * HD + c] = f2bf(ll > 0.f ? acc / ll : 0.f);
        }
        if (tid == 0) g_split_arrivals[b * KV + kvh] = 0;      // ready for the next launch
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- persistent variant
// The same tile arithmetic, cut differently.  With one CTA per (request, kv head, split) the launch runs in rounds of
// 2 x 148 resident CTAs and the last round is part empty: at 64 requests x 2176 keys (1024 CTAs, 3.46 rounds) the SMs
// are busy 84 us of a 101 us kernel, and while they are busy HBM runs at 6.8 TB/s (profiles/r02_summary.md).  Here the
// key tiles of all (request, kv head) groups form ONE sequence of U units, and CTA c of G takes the contiguous share
// [c * chunk, (c + 1) * chunk): every CTA streams the same number of tiles, in one round.  A share covers the end of
// one group, whole groups, and the start of another; each piece ("segment") is an ordinary split-KV partial, merged by
// the last CTA of its group to arrive, in split order.  The producer warp runs on into the next segment while the
// compute warps merge the current one (the merge buffer has its own shared memory, four q-head rows at a time, and the
// next segment's Q tile arrives in a second buffer).
constexpr int kMaxReq = 1024;                // requests one launch may have (prefix of their tile counts in shared memory)

struct Segment { int b, kvh, t_begin, t_end, split, nsplit, kv_len; };

template <int HD>
__global__ void __launch_bounds__(160, 2)
attn_decode_tma_persistent_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                                  bf16* __restrict__ out, float* __restrict__ part_o, float* __restrict__ part_ml,
                                  const bf16* __restrict__ q, int q_stride, const int32_t* __restrict__ block_table, int max_blocks,
                                  const int32_t* __restrict__ seq_lens, const int32_t* __restrict__ qsl, int B, int H, int KV,
                                  int bs_shift, float scale_log2, int slots /* partial slots per (request, head) */, int min_chunk) {
  constexpr int CH = HD / 8;
  constexpr uint32_t TILE = BN * HD * 2;
  constexpr uint32_t Q_BYTES = 16 * HD * 2;
  constexpr int RS = HD + 2;
  const int G = H / KV;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sQ = s_addr(smem);              // two Q tiles: the next segment's arrives during this segment's merge
  const uint32_t sK = sQ + 2 * Q_BYTES;
  const uint32_t sV = sK + kStages * TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * Q_BYTES + 2 * kStages * TILE);
  const uint32_t bar0 = s_addr(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (kStages + s); };
  volatile int* s_last = reinterpret_cast<volatile int*>(bars + 2 * kStages);
  float* sRed = reinterpret_cast<float*>(smem + 2 * Q_BYTES + 2 * kStages * TILE + 2 * kStages * 8 + 64);   // [4 warps][4 rows][HD+2]
  int* s_pre = reinterpret_cast<int*>(sRed + 4 * 4 * RS);      // [B + 1] tiles of the requests before b (per kv head)

  if (tid == 0) {
    if ((sQ & 1023u) != 0) __trap();
    for (int s = 0; s < kStages; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
  }
  pdl_launch_dependents();
  pdl_wait();
  // tiles per request (a request without keys still gets one, fully masked, tile: its output rows are written as zeros)
  for (int b = tid; b < B; b += 160) { const int n = (seq_lens[b] + BN - 1) / BN; s_pre[b + 1] = n > 0 ? n : 1; }
  if (tid == 0) s_pre[0] = 0;
  __syncthreads();
  if (warp == 0) {                                 // inclusive scan, each lane a contiguous run of requests
    const int per = (B + 31) / 32;
    const int lo = 1 + lane * per, hi = min(B + 1, lo + per);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += s_pre[i];
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
    int run = incl - sum;
    for (int i = lo; i < hi; ++i) { run += s_pre[i]; s_pre[i] = run; }
  }
  __syncthreads();
  const int total_tiles = s_pre[B];                // per kv head
  const int64_t U = (int64_t)total_tiles * KV;
  int chunk = (int)((U + gridDim.x - 1) / gridDim.x);
  if (chunk < min_chunk) chunk = min_chunk;        // bounds the segments of one group by the partial slots there are
  const int64_t u0 = (int64_t)blockIdx.x * chunk, u1 = min(U, u0 + chunk);

  // The walk over my share: units are ordered request-major, then kv head, then tile.
  struct Walk {
    int64_t u, u1;
    int b, chunk, KV;
    const int* pre;
    const int32_t* lens;          // (global: two or three loads per CTA; a shared copy would cost the second resident CTA from 98 requests on)
    __device__ __forceinline__ bool next(Segment& sg) {
      if (u >= u1) return false;
      while ((int64_t)pre[b + 1] * KV <= u) ++b;
      const int nt = pre[b + 1] - pre[b];
      const int64_t base = (int64_t)pre[b] * KV;
      const int rel = (int)(u - base);
      sg.b = b; sg.kvh = rel / nt; sg.t_begin = rel % nt;
      const int64_t left = u1 - u;
      sg.t_end = (left < (int64_t)(nt - sg.t_begin)) ? sg.t_begin + (int)left : nt;
      const int64_t g0 = base + (int64_t)sg.kvh * nt, g1 = g0 + nt;      // the group's units
      const int c_first = (int)(g0 / chunk), c_last = (int)((g1 - 1) / chunk);
      sg.split = (int)(u / chunk) - c_first;
      sg.nsplit = c_last - c_first + 1;
      sg.kv_len = lens[b];
      u += sg.t_end - sg.t_begin;
      return true;
    }
  };
  auto walk = [&]() {
    Walk w{u0, u1, 0, chunk, KV, s_pre, seq_lens};
    if (u0 < u1) {                                 // first request of my share: binary search in the prefix
      int lo = 0, hi = B - 1;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if ((int64_t)s_pre[mid + 1] * KV <= u0) lo = mid + 1; else hi = mid; }
      w.b = lo;
    }
    return w;
  };

  if (warp == 4) {
    // ---------------------------------------------------------------- producer (the whole warp), as in the kernel above
    const int bs = 1 << bs_shift;
    const int box_rows = bs < BN ? bs : BN;
    const int boxes = BN / box_rows;
    int it = 0;                                                            // tiles issued by this CTA so far
    uint64_t pol;                                                          // K/V are read once per step: first out of L2 (+1.7 %)
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    Walk w = walk();
    Segment sg;
    while (w.next(sg)) {
      const int kvh = sg.kvh;
      const int32_t* btab = block_table + (int64_t)sg.b * max_blocks;
      const bool has_keys = sg.kv_len > 0;
      auto ld_blk = [&](int bi) { return (has_keys && bi >= 0 && bi < max_blocks) ? btab[bi] : 0; };
      int cbase = (sg.t_begin * BN) >> bs_shift;
      int v = ld_blk(cbase + lane), v_next = ld_blk(cbase + 32 + lane);
      for (int kt = sg.t_begin; kt < sg.t_end; ++kt, ++it) {
        const int st = it % kStages;
        const int key = kt * BN + (lane < boxes ? lane : 0) * box_rows;
        const int bi = key >> bs_shift;
        const int bi_last = (kt * BN + (boxes - 1) * box_rows) >> bs_shift;
        if (bi_last >= cbase + 32) {
          const int nb = (kt * BN) >> bs_shift;
          if (nb == cbase + 32) { v = v_next; } else { v = ld_blk(nb + lane); }
          cbase = nb;
          v_next = ld_blk(cbase + 32 + lane);
        }
        const int blk = __shfl_sync(0xffffffffu, v, (bi - cbase) & 31);
        if (lane == 0) {
          if (it >= kStages) mbar_wait(empty(st), (uint32_t)((it / kStages - 1) & 1));
          mbar_expect_tx(full(st), 2 * TILE);
        }
        __syncwarp();
        if (lane < boxes) {
          const int row = (blk * KV + kvh) * bs + (key & (bs - 1));
#pragma unroll
          for (int h = 0; h < HD / 64; ++h) {
            const uint32_t off = (uint32_t)(h * (BN * 128) + lane * box_rows * 128);
            tma_load_2d_hint(sK + st * TILE + off, &tm_k, full(st), h * 64, row, pol);
            tma_load_2d_hint(sV + st * TILE + off, &tm_v, full(st), h * 64, row, pol);
          }
        }
      }
    }
    return;
  }

  // ------------------------------------------------------------------ compute warps
  const int g = lane >> 2, tg = lane & 3;
  int it = 0;                                                              // tiles consumed by this CTA so far
  auto load_q = [&](int buf, int qrow, int kvh) {                          // Q tile of a segment into buffer `buf`
    for (int i = tid; i < 16 * CH; i += 128) {
      const int r = i / CH, c = i % CH;
      const bool ok = r < G;
      const bf16* src = ok ? q + (int64_t)qrow * q_stride + (kvh * G + r) * HD + c * 8 : q;
      cp_async16(sQ + buf * Q_BYTES + swz_q<HD>(r, c), src, ok);
    }
    asm volatile("cp.async.commit_group;");
  };
  Walk w = walk();
  Segment sg, sg_next;
  bool have = w.next(sg);
  int qrow = 0, qrow_next = 0, qbuf = 0;
  if (have) { qrow = qsl[sg.b]; load_q(0, qrow, sg.kvh); }
  for (; have; qbuf ^= 1) {
    const bool have_next = w.next(sg_next);
    if (have_next) qrow_next = qsl[sg_next.b];
    const int b = sg.b, kvh = sg.kvh, split = sg.split, nsplit = sg.nsplit, kv_len = sg.kv_len;
    if (nsplit > slots) __trap();                 // max_seq_len understated by the caller: a group has more pieces than slots
    asm volatile("cp.async.wait_group 0;" ::: "memory");                   // this segment's Q (requested one segment ago)
    bar_compute();
    float o[HD / 8][4];
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk)
      ldsm_x4(qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3], sQ + qbuf * Q_BYTES + swz_q<HD>(lane & 15, kk * 2 + (lane >> 4)));
    for (int kt = sg.t_begin; kt < sg.t_end; ++kt, ++it) {
      const int st = it % kStages;
      mbar_wait(full(st), (uint32_t)((it / kStages) & 1));
      const uint32_t kb = sK + st * TILE, vb = sV + st * TILE;
      float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(b0, b1, b2, b3, kb + swz_kv(warp * 16 + (lane & 7) + ((lane >> 4) << 3), kk * 2 + ((lane >> 3) & 1)));
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
        ldsm_x4_t(b0, b1, b2, b3, vb + swz_kv(warp * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), nn * 2 + (lane >> 4)));
        mma_bf16(o[2 * nn], pf, b0, b1);
        mma_bf16(o[2 * nn + 1], pf, b2, b3);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty(st));
    }
    // the next segment's Q goes to the other buffer (last read at the start of the previous segment)
    if (have_next) load_q(qbuf ^ 1, qrow_next, sg_next.kvh);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    // merge the four warps' (m, l, O), four q-head rows per pass: sRed[w][row - r0][HD] + ml at [HD], [HD+1]
    float* mine = sRed + warp * 4 * RS;
    for (int r0 = 0; r0 < G; r0 += 4) {
      if (g >= r0 && g < r0 + 4) {
        float* dst = mine + (g - r0) * RS;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) { dst[i * 8 + 2 * tg] = o[i][0]; dst[i * 8 + 2 * tg + 1] = o[i][1]; }
        if (tg == 0) { dst[HD] = m0; dst[HD + 1] = l0; }
      }
      if (g + 8 >= r0 && g + 8 < r0 + 4) {
        float* dst = mine + (g + 8 - r0) * RS;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) { dst[i * 8 + 2 * tg] = o[i][2]; dst[i * 8 + 2 * tg + 1] = o[i][3]; }
        if (tg == 0) { dst[HD] = m1; dst[HD + 1] = l1; }
      }
      bar_compute();
      for (int i = tid; i < 4 * HD; i += 128) {
        const int rl = i / HD, c = i % HD, r = r0 + rl;
        if (r >= G) break;
        float mm = -INFINITY;
#pragma unroll
        for (int w2 = 0; w2 < 4; ++w2) mm = fmaxf(mm, sRed[(w2 * 4 + rl) * RS + HD]);
        float acc = 0.f, ll = 0.f;
        const float basem = (mm == -INFINITY) ? 0.f : mm;
#pragma unroll
        for (int w2 = 0; w2 < 4; ++w2) {
          const float sc = exp2f(sRed[(w2 * 4 + rl) * RS + HD] - basem);
          acc += sRed[(w2 * 4 + rl) * RS + c] * sc;
          ll += sRed[(w2 * 4 + rl) * RS + HD + 1] * sc;
        }
        const int head = kvh * G + r;
        if (nsplit == 1) {
          out[(int64_t)qrow * H * HD + head * HD + c] = f2bf(ll > 0.f ? acc / ll : 0.f);
        } else {
          const int64_t pi = ((int64_t)b * H + head) * slots + split;
          part_o[pi * HD + c] = acc;
          if (c == 0) { part_ml[pi * 2] = mm; part_ml[pi * 2 + 1] = ll; }
        }
      }
      bar_compute();                              // the rows are read: the next pass (or segment) may overwrite them
    }
    if (nsplit > 1) {
      if (tid == 0) {                             // (the barrier above ordered every thread's partial stores before this)
        unsigned prev;                              // release my CTA's partials, acquire the others': one instruction
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(prev) : "l"(&g_split_arrivals[b * KV + kvh]) : "memory");
        *s_last = (prev == (unsigned)nsplit - 1u) ? 1 : 0;
      }
      bar_compute();
      if (*s_last) {
        // Every load here is an L2 round trip, and all CTAs reach this point together at the end of the kernel: first the
        // (m, l) pairs of every (row, split) into shared memory in one trip, then the partial outputs eight loads at a time.
        float* sMl = sRed;                          // [G][nsplit][2] (the merge passes are done with sRed)
        const int64_t grp0 = ((int64_t)b * H + kvh * G) * slots;
        for (int i = tid; i < G * nsplit * 2; i += 128) {
          const int r = i / (nsplit * 2), rest = i % (nsplit * 2);
          sMl[i] = __ldcg(part_ml + (grp0 + (int64_t)r * slots) * 2 + rest);
        }
        bar_compute();
        // item = (row r, column c); the (item, split) pairs of this thread form one flat sequence, consumed in order
        const int n_my = (G * HD - tid + 127) / 128;          // items tid, tid + 128, ...
        const int n_flat = n_my * nsplit;
        float acc = 0.f, ll = 0.f, basem = 0.f;
        for (int f0 = 0; f0 < n_flat; f0 += 8) {
          float v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int f = f0 + k;
            const int item = tid + 128 * (f / nsplit), sidx = f % nsplit;
            const int r = item / HD, c = item % HD;
            v[k] = f < n_flat ? __ldcg(part_o + (grp0 + (int64_t)r * slots + sidx) * HD + c) : 0.f;
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int f = f0 + k;
            if (f < n_flat) {
              const int item = tid + 128 * (f / nsplit), sidx = f % nsplit;
              const int r = item / HD, c = item % HD;
              const float* ml = sMl + (r * nsplit) * 2;
              if (sidx == 0) {
                float mm = -INFINITY;
                for (int s2 = 0; s2 < nsplit; ++s2) mm = fmaxf(mm, ml[s2 * 2]);
                basem = (mm == -INFINITY) ? 0.f : mm;
                acc = 0.f; ll = 0.f;
              }
              const float sc = exp2f(ml[sidx * 2] - basem);
              acc += v[k] * sc;
              ll += ml[sidx * 2 + 1] * sc;
              if (sidx == nsplit - 1) out[(int64_t)qrow * H * HD + (kvh * G + r) * HD + c] = f2bf(ll > 0.f ? acc / ll : 0.f);
            }
          }
        }
        bar_compute();                              // sMl is read: the next segment's merge passes may overwrite it
        if (tid == 0) g_split_arrivals[b * KV + kvh] = 0;
      }
    }
    have = have_next; sg = sg_next; qrow = qrow_next;
  }
}

template <int HD> constexpr int smem_bytes_persistent(int B) {
  return 2 * 16 * HD * 2 + 2 * kStages * BN * HD * 2 + 2 * kStages * 8 + 64 + 4 * 4 * (HD + 2) * 4 + (B + 1) * 4;
}

template <int HD> constexpr int smem_bytes() {
  static_assert(4 * 16 * (HD + 2) * 4 <= kStages * BN * HD * 2, "sRed must fit in the K stages");
  return 16 * HD * 2 + 2 * kStages * BN * HD * 2 + 2 * kStages * 8 + 64;
}

}  // namespace adt

// Launcher used by kge_paged_attn / kge_paged_attn_decode_noalloc (attn.cu).  ws: split-KV partials (ns > 1).
template <int HD>
static int launch_decode_tma(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt, int max_blocks,
                             const int32_t* sl, const int32_t* qsl, int B, int H, int KV, int bs, int ns, float scale_log2,
                             float* part_o, float* part_ml, void* ws, int64_t ws_bytes, int max_seq_len, cudaStream_t st) {
  int bs_shift = 0;
  while ((1 << bs_shift) < bs) ++bs_shift;
  CUtensorMap tk, tv;
  const int box_rows = bs < adt::BN ? bs : adt::BN;
  // rows: the pool's real extent is not known here and only bounds TMA's out-of-range handling, which is never relied on
  const int64_t rows = (int64_t)1 << 30;
  if (int rc = make_tmap(&tk, kc, rows, HD, box_rows)) return rc;
  if (int rc = make_tmap(&tv, vc, rows, HD, box_rows)) return rc;
  // persistent variant (knob attn_persistent): needs partial slots for the segments of a group in the caller's workspace
  if (tunable(TUN_ATTN_PERSISTENT) && ws && B <= 192) {
    int64_t slots = ws_bytes / ((int64_t)B * H * (HD + 2) * 4);
    if (slots > 32) slots = 32;
    if (slots >= 3) {
      const int max_tiles = max_seq_len > 0 ? (max_seq_len + adt::BN - 1) / adt::BN : 1;
      const int min_chunk = (max_tiles + (int)slots - 2) / ((int)slots - 1);
      auto kern = adt::attn_decode_tma_persistent_kernel<HD>;
      const int smem = adt::smem_bytes_persistent<HD>(B);
      KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      prefer_max_shared(kern);
      // the grid is 2 x 148 CTAs that must all be resident: the per-request prefix grows the footprint with B
      static int occ_of[193];                      // 0: not asked yet
      if (occ_of[B] == 0) {
        int o = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kern, 160, smem) != cudaSuccess) { (void)cudaGetLastError(); o = 1; }
        occ_of[B] = o > 0 ? o : 1;
      }
      const int64_t units = (int64_t)B * KV * max_tiles;
      // (with less than two tiles per CTA the per-CTA prefix scan is not paid back)
      if (occ_of[B] >= 2 && units >= 2 * 2 * 148) {
        const int grid = 2 * 148;
        float* po = (float*)ws;
        float* pml = po + (int64_t)B * H * slots * HD;
        KGE_CUDA_OK(launch_pdl(kern, dim3(grid), dim3(160), smem, st, tk, tv, out, po, pml, q, q_stride, bt, max_blocks, sl, qsl, B, H, KV,
                               bs_shift, scale_log2, (int)slots, min_chunk));
        return 0;
      }
    }
  }
  auto kern = adt::attn_decode_tma_kernel<HD>;
  KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, adt::smem_bytes<HD>()));
  KGE_CUDA_OK(launch_pdl(kern, dim3(ns, KV, B), dim3(160), adt::smem_bytes<HD>(), st, tk, tv, out, part_o, part_ml, q, q_stride, bt,
                         max_blocks, sl, qsl, H, KV, bs_shift, scale_log2));
  return 0;
}

bool decode_tma_supported(int B, int KV, int bs) { return bs >= 8 && (int64_t)B * KV <= adt::kMaxGroups; }

int paged_attn_decode_tma(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt, int max_blocks,
                          const int32_t* sl, const int32_t* qsl, int B, int H, int KV, int hd, int bs, int ns, float scale_log2,
                          float* part_o, float* part_ml, void* ws, int64_t ws_bytes, int max_seq_len, cudaStream_t st) {
  if (hd == 128)
    return launch_decode_tma<128>(out, q, q_stride, kc, vc, bt, max_blocks, sl, qsl, B, H, KV, bs, ns, scale_log2, part_o, part_ml, ws, ws_bytes,
                                  max_seq_len, st);
  return launch_decode_tma<64>(out, q, q_stride, kc, vc, bt, max_blocks, sl, qsl, B, H, KV, bs, ns, scale_log2, part_o, part_ml, ws, ws_bytes,
                               max_seq_len, st);
}

}  // namespace KGE_ELT_NS
}  // namespace kge
