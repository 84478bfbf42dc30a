// Dense projection  y[T,N] = act(x[T,K] * w[N,K]^T + bias)  on the 5th-gen tensor cores.
//
// Weight-stationary orientation ("swap-AB"): the 128-row MMA M dimension runs over WEIGHT rows
// (N is always large: 768..128256) and the MMA N dimension (16..256) over TOKENS, so a decode
// step with T = 16..256 tokens wastes no tensor-core rows and streams every weight byte exactly
// once, while prefill uses 128x256 tiles.  D[weight row, token] lives in TMEM (fp32); the
// epilogue reads it back with tcgen05.ld, adds bias (one scalar per TMEM lane), applies the
// activation, rounds once to bf16 and stores y transposed.
//
// Persistent, warp-specialised: warp 0 = TMA producer (cp.async.bulk.tensor, 128B swizzle),
// warp 1 = tcgen05.mma issuer (one elected lane), warp 2 = TMEM allocator, warps 4-7 = epilogue.
// smem ring of STAGES {W tile 128x64, X tile TOKNx64}; two TMEM accumulator stages so the
// epilogue of tile i overlaps the main loop of tile i+1.
// Split-K (decode, when N/128 tiles cannot fill 148 SMs): the S CTAs of a thread-block CLUSTER
// share one output tile and each take 1/S of K.  Their fp32 partial accumulators are exchanged
// through distributed shared memory (st.shared::cluster): CTA r receives column slice r of every
// peer, sums the S partials in rank order (deterministic), adds bias, rounds once and stores.
// No global scratch, no atomics, no serial tail; the cluster guarantees co-scheduling.
// Stream-K through a workspace (WS; decode-sized projections when the caller provides scratch): the (tile, k-block) units
// of the whole problem are laid end to end and cut into one contiguous share per CTA, one CTA per SM - every SM streams
// exactly 1/148 of the weights through ONE uninterrupted TMA ring, whatever the tile count (48 tiles x split 2 used 96 SMs,
// 224 tiles used 112).  A CTA's share covers at most one partial tile at its start and one at its end: those fp32 partials
// go to the workspace (they stay in L2) and the CTA that arrives LAST on the tile's counter sums the tile's partials in
// k order (deterministic), applies bias / activation / SiLU-gate, rounds once and stores; whole tiles in between take the
// ordinary epilogue.  The accumulators are double-buffered in TMEM, so all of this overlaps the next segment's main loop.
//
// Roofline: decode T<=256 -> HBM, algorithmic bytes = N*K*2 (weights) + T*(K+N)*2;
//           prefill        -> tensor, 2*T*N*K flop.
#include <cuda.h>
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

constexpr int BLK_M = 128;   // weight rows per tile (UMMA M)
constexpr int BLK_K = 64;    // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 256;
constexpr uint32_t kSpinLimit = 1u << 24;   // seconds at most: a lost arrival traps instead of hanging the GPU

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

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
    if (!ok && ++spins > kSpinLimit) __trap();   // a lost arrival becomes an error, not a hang
  } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// The same loads with an L2 eviction-priority hint (knob gemm_l2_hints): weights are read once per launch (evict_first),
// the activation tile is re-read by every CTA (evict_last), so the prefill GEMM stops re-fetching X from HBM behind its
// own output stream (ncu, round 1: 2.42 GB of DRAM traffic for 0.77 GB of algorithmic bytes on the gate_up shape).
__device__ __forceinline__ uint64_t l2_policy(bool keep) {
  uint64_t pol;
  if (keep) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  else asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, uint64_t pol) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(pol) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
               ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128B-swizzled operand tile (rows of 128 B, 8-row atoms 1024 B apart).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;                    // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;          // stride byte offset between 8-row atoms
  d |= (uint64_t)1 << 46;                    // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
  return d;
}

struct GemmParams {
  bf16* y;
  const bf16* bias;
  float* ws_part;            // WS: [tile][split][token column][128 rows] fp32 partial accumulators
  unsigned int* ws_count;    // WS: per-tile arrival counters (zero before the first launch; the reducing CTA resets its own)
  int T, N, K, act;
  int l2_hints;              // 1: weights evict_first, activations evict_last, output stored streaming
  int m_tiles, n_tiles, splits, kb_total, kb_per_split;
  int rows;                  // gemm_wide_kernel: outputs per tile
  unsigned long long* trace; // phase timestamps, 16 per CTA (tools/gemm_trace.py); null outside that tool
};

// One %globaltimer stamp of this CTA's phase `slot` (see tools/gemm_trace.py for the slot list).
__device__ __forceinline__ void trace_stamp(const GemmParams& p, int slot) {
  if (p.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory");
    p.trace[(size_t)blockIdx.x * 16 + slot] = t;
  }
}

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_nctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire;" ::: "memory"); }
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}

// exchange buffer of the split variant: [src rank][local column][128 rows] fp32
__host__ __device__ constexpr int xbuf_bytes(int tokn) { return (tokn + 8) * BLK_M * 4; }

// GATED: the 128 MMA rows of a tile are 64 gate rows (TMEM lanes 0-63) and the matching 64 up rows
// (lanes 64-127) of a stacked [2f, K] weight; the epilogue emits silu(gate) * up for the 64
// outputs with vLLM's cast points (vllm:model_executor/layers/activation.py:137-141).
__device__ __forceinline__ float silu_gate_bf16(float g_acc, float u_acc) {
  const float g = rbf(g_acc), u = rbf(u_acc);            // the gate_up GEMM output is bf16
  const float sg = rbf(__fdividef(g, 1.0f + __expf(-g))); // F.silu rounds to bf16 (fast exp/div: <= 2 ulp in fp32)
  return sg * u;                                          // caller rounds the product once
}

__device__ __forceinline__ void tma_load_2d_cg2_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1, uint64_t pol) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(dst), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1), "l"(pol) : "memory");
}
__device__ __forceinline__ void tma_load_2d_cg2(uint32_t dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_commit_cg2(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_cg2(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}"
               ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster) : "memory");
}

// PAIR (prefill, TOKN = 256): cta_group::2.  The two CTAs of a cluster form one 256 x 256 MMA tile:
// each holds its own 128 weight rows (A) and HALF of the token tile (B, 128 tokens) in shared memory
// (32 KB per stage instead of 48 KB: less L2->SM traffic and less shared-memory bandwidth), the
// leader CTA issues tcgen05.mma.cta_group::2 for both, accumulators land in each CTA's own TMEM.
// All TMA loads of a stage complete on the LEADER's full barrier; tcgen05.commit multicasts the
// "stage free" and "accumulator ready" arrivals to both CTAs; both epilogues release the accumulator
// on the leader's tmem_empty barrier.
template <int TOKN, int STAGES, bool SPLIT, bool GATED, bool PAIR = false, bool WS = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
               const GemmParams p) {
  static_assert(!(PAIR && SPLIT) && !(WS && (SPLIT || PAIR)), "one split scheme at a time");
  constexpr uint32_t A_BYTES = BLK_M * BLK_K * 2, B_BYTES = (PAIR ? 128 : TOKN) * BLK_K * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = (2 * TOKN < 32) ? 32 : 2 * TOKN;
  constexpr uint32_t IDESC = (1u << 4) | KGE_IDESC_AB | ((uint32_t)(TOKN >> 3) << 17) |
                             ((uint32_t)((PAIR ? 2 * BLK_M : BLK_M) >> 4) << 24);

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  // bars: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  float* xbuf = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES + 2048);   // SPLIT: exchange buffer;
                                                                                // GATED && !SPLIT: up->gate hand-off [2][16][64]
  constexpr int ROWS_OUT = GATED ? 64 : BLK_M;           // output columns produced per tile
  const uint32_t smem_base = smem_addr(smem);
  const uint32_t bar_base = smem_addr(bars);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    trace_stamp(p, 0);
    if (p.trace) { uint32_t sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); p.trace[(size_t)blockIdx.x * 16 + 15] = sm; }
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), PAIR ? 8 : 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_slot)), "n"(TMEM_COLS) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_slot)), "n"(TMEM_COLS) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) trace_stamp(p, 1);
  if (PAIR) { cluster_arrive(); cluster_wait(); }   // peer barriers are initialised before any remote signal
  // TMEM is held: the next kernel may start arriving now (see attn_tc.cu for why not earlier)
  pdl_launch_dependents();
  if (warp != 0) pdl_wait();          // warp 0 (TMA producer) prefetches weights first, then waits

  // work decomposition: persistent round-robin over tiles, or (SPLIT) one tile per cluster with
  // the cluster rank selecting the K range
  const int S = SPLIT ? (int)cluster_nctarank() : 1;
  const int crank = SPLIT ? (int)cluster_ctarank() : 0;
  constexpr bool TWO = PAIR;
  const int mrank = TWO ? (int)cluster_ctarank() : 0;            // which tile of the weight-row pair
  const int total = (TWO ? (p.m_tiles + 1) / 2 : p.m_tiles) * p.n_tiles;
  const int w_first = SPLIT ? (int)blockIdx.x / S : TWO ? (int)blockIdx.x / 2 : (int)blockIdx.x;
  const int w_step = SPLIT ? (1 << 30) : TWO ? (int)gridDim.x / 2 : (int)gridDim.x;
  auto m_tile_of = [&](int w) { return TWO ? 2 * (w / p.n_tiles) + mrank : w / p.n_tiles; };
  // PAIR: every TMA of a stage completes on the leader's (rank 0) full barrier
  auto lead_full = [&](int s) { return PAIR ? mapa_u32(full_bar(s), 0u) : full_bar(s); };
  const int kb0f = crank * p.kb_per_split, kb1f = min(p.kb_total, kb0f + p.kb_per_split);
  const int cols_per = (TOKN + S - 1) / S;               // columns finished by each cluster rank
  // A CTA's work as a sequence of segments (tile w, k-blocks [kb0, kb1)).  Persistent / split variants: its tiles, each
  // with the same k range.  WS (stream-K): its contiguous share [u_begin, u_end) of the tiles * kb_total units.
  const int KB = p.kb_total;
  const int64_t U = (int64_t)total * KB;
  const int64_t u_begin = WS ? U * (int64_t)blockIdx.x / (int64_t)gridDim.x : 0;
  const int64_t u_end = WS ? U * ((int64_t)blockIdx.x + 1) / (int64_t)gridDim.x : 0;
  struct SegIter {
    int64_t cur, end;
    int step, KB, k0, k1;
    __device__ __forceinline__ bool next(int& w, int& kb0, int& kb1) {
      if (cur >= end) return false;
      if (WS) {
        w = (int)(cur / KB); kb0 = (int)(cur % KB);
        const int64_t left = end - cur;
        kb1 = (left < (int64_t)(KB - kb0)) ? kb0 + (int)left : KB;
        cur += kb1 - kb0;
      } else {
        w = (int)cur; kb0 = k0; kb1 = k1; cur += step;
      }
      return true;
    }
  };
  auto segments = [&]() { return WS ? SegIter{u_begin, u_end, 0, KB, 0, 0} : SegIter{(int64_t)w_first, (int64_t)total, w_step, KB, kb0f, kb1f}; };
  if (SPLIT) cluster_arrive_relaxed();                   // phase A: every CTA of the cluster has started (nothing to publish: no release)

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      const bool hint = p.l2_hints != 0;
      const uint64_t pol_w = l2_policy(false), pol_x = l2_policy(true);
      auto load_a = [&](int st, int kb, int m_tile) {
        const uint32_t a_dst = smem_base + st * STAGE_BYTES;
        if (PAIR && hint) {
          if (GATED) {
            tma_load_2d_cg2_hint(a_dst, &tmap_w, lead_full(st), kb * BLK_K, m_tile * 64, pol_w);
            tma_load_2d_cg2_hint(a_dst + 64 * BLK_K * 2, &tmap_w, lead_full(st), kb * BLK_K, p.N + m_tile * 64, pol_w);
          } else {
            tma_load_2d_cg2_hint(a_dst, &tmap_w, lead_full(st), kb * BLK_K, m_tile * BLK_M, pol_w);
          }
        } else if (PAIR) {
          if (GATED) {
            tma_load_2d_cg2(a_dst, &tmap_w, lead_full(st), kb * BLK_K, m_tile * 64);
            tma_load_2d_cg2(a_dst + 64 * BLK_K * 2, &tmap_w, lead_full(st), kb * BLK_K, p.N + m_tile * 64);
          } else {
            tma_load_2d_cg2(a_dst, &tmap_w, lead_full(st), kb * BLK_K, m_tile * BLK_M);
          }
        } else if (hint) {
          if (GATED) {
            tma_load_2d_hint(a_dst, &tmap_w, full_bar(st), kb * BLK_K, m_tile * 64, pol_w);
            tma_load_2d_hint(a_dst + 64 * BLK_K * 2, &tmap_w, full_bar(st), kb * BLK_K, p.N + m_tile * 64, pol_w);
          } else {
            tma_load_2d_hint(a_dst, &tmap_w, full_bar(st), kb * BLK_K, m_tile * BLK_M, pol_w);
          }
        } else if (GATED) {
          tma_load_2d(a_dst, &tmap_w, full_bar(st), kb * BLK_K, m_tile * 64);                          // gate rows
          tma_load_2d(a_dst + 64 * BLK_K * 2, &tmap_w, full_bar(st), kb * BLK_K, p.N + m_tile * 64);   // up rows
        } else {
          tma_load_2d(a_dst, &tmap_w, full_bar(st), kb * BLK_K, m_tile * BLK_M);
        }
      };
      // the barrier that counts a stage's bytes: own (48 KB) or, PAIR, the leader's for both CTAs (2 x 32 KB)
      auto expect = [&](int st) {
        if (!PAIR) mbar_expect_tx(full_bar(st), STAGE_BYTES);
        else if (mrank == 0) mbar_expect_tx(full_bar(st), 2 * STAGE_BYTES);
      };
      // Weights do not depend on the previous kernel: stream the first stages of W while it is
      // still running, then wait for its output (the activations) and add the X tiles.
      int pre = 0;
      auto load_b = [&](int st, int kb, int n_tile) {
        const uint32_t b_dst = smem_base + st * STAGE_BYTES + A_BYTES;
        if (PAIR && hint) tma_load_2d_cg2_hint(b_dst, &tmap_x, lead_full(st), kb * BLK_K, n_tile * TOKN + mrank * 128, pol_x);
        else if (PAIR) tma_load_2d_cg2(b_dst, &tmap_x, lead_full(st), kb * BLK_K, n_tile * TOKN + mrank * 128);
        else if (hint) tma_load_2d_hint(b_dst, &tmap_x, full_bar(st), kb * BLK_K, n_tile * TOKN, pol_x);
        else tma_load_2d(b_dst, &tmap_x, full_bar(st), kb * BLK_K, n_tile * TOKN);
      };
      int w, kb0, kb1;
      trace_stamp(p, 12);
      {
        SegIter peek = segments();
        if (peek.next(w, kb0, kb1)) {            // the first segment's first stages (a later segment's are loaded in the loop)
          const int m_tile = m_tile_of(w);
          pre = min(STAGES, kb1 - kb0);
          for (int i = 0; i < pre; ++i) {
            expect(i);
            load_a(i, kb0 + i, m_tile);
            if (i == 0) trace_stamp(p, 13);
          }
        }
      }
      trace_stamp(p, 2);
      pdl_wait();
      trace_stamp(p, 3);
      SegIter it = segments();
      bool first_seg = true;
      while (it.next(w, kb0, kb1)) {
        const int n_tile = w % p.n_tiles, m_tile = m_tile_of(w);
        for (int kb = kb0; kb < kb1; ++kb) {
          if (first_seg && kb - kb0 < pre) {
            load_b(stage, kb, n_tile);
          } else {
            mbar_wait(empty_bar(stage), phase ^ 1);
            expect(stage);
            load_a(stage, kb, m_tile);
            load_b(stage, kb, n_tile);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        first_seg = false;
      }
      trace_stamp(p, 4);
    }
  } else if (warp == 1) {
    if (lane == 0 && (!PAIR || mrank == 0)) {          // PAIR: the leader issues for both CTAs
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      SegIter it = segments();
      int w, kb0, kb1;
      if (p.trace) { mbar_wait(full_bar(0), 0); trace_stamp(p, 5); }      // (a wait does not consume the phase)
      // This one thread paces a decode-sized GEMM (every instruction between two k-blocks is serial latency), so the
      // loop carries the stage's descriptors along instead of rebuilding them.
      const uint64_t adesc0 = umma_desc_sw128(smem_base), bdesc0 = umma_desc_sw128(smem_base + A_BYTES);
      uint64_t adesc = adesc0, bdesc = bdesc0;
      while (it.next(w, kb0, kb1)) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * TOKN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < BLK_K / UMMA_K; ++k) {
            // +32 B per 16-element k step inside the 128 B swizzle row (start address field is >>4)
            if (PAIR) tc_mma_bf16_cg2(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
            else tc_mma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          if (PAIR) tc_commit_cg2(empty_bar(stage), (uint16_t)3); // frees the stage in both CTAs
          else tc_commit(empty_bar(stage));     // frees the smem slot when these MMAs retire
          adesc += STAGE_BYTES >> 4; bdesc += STAGE_BYTES >> 4;
          if (++stage == STAGES) { stage = 0; phase ^= 1; adesc = adesc0; bdesc = bdesc0; }
        }
        if (PAIR) tc_commit_cg2(tfull_bar(acc), (uint16_t)3);   // both CTAs' epilogues
        else tc_commit(tfull_bar(acc));         // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      trace_stamp(p, 6);
    }
  } else if (warp >= 4) {
    const int q = warp - 4;                      // TMEM lane quarter == warp % 4
    const int et = threadIdx.x - 128;            // 0..127 == weight row within the tile
    int acc = 0; uint32_t acc_phase = 0;
    SegIter it = segments();
    int w, kb0, kb1;
    while (it.next(w, kb0, kb1)) {
      const int n_tile = w % p.n_tiles, m_tile = m_tile_of(w);
      const int n = m_tile * ROWS_OUT + (GATED ? (et & 63) : et);   // output column of this thread's row
      const int t0 = n_tile * TOKN;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      if (et == 0) trace_stamp(p, 7);              // (of the CTA's last segment: later segments overwrite the stamp)
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * TOKN;
      const float bv = (p.bias && n < p.N) ? bf2f(p.bias[n]) : 0.f;
      if (WS && (kb0 != 0 || kb1 != KB)) {
        // ---- stream-K: this segment is only part of tile w's k range.  Its fp32 partial goes to the workspace
        // ([column][128 rows]: a warp stores 128 consecutive bytes per column); slot 0 of a CTA holds the partial that
        // STARTS its share, slot 1 the one that ends it.
        const int slot = ((int64_t)w * KB + kb0 == u_begin) ? 0 : 1;
        float* part = p.ws_part + ((int64_t)(blockIdx.x * 2 + slot) * TOKN) * BLK_M;
#pragma unroll 1
        for (int c = 0; c < TOKN; c += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + c, r);
#pragma unroll
          for (int j = 0; j < 16; ++j) part[(c + j) * BLK_M + et] = __uint_as_float(r[j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(acc));       // the accumulator is free: the next segment's MMAs may start
        __threadfence();                                   // the partial is visible device-wide before this CTA arrives
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // which CTAs hold a part of tile w: the one containing its first unit ... the one containing its last unit
        const int64_t G = gridDim.x, x0 = (int64_t)w * KB, x1 = x0 + KB - 1;
        const int c_first = (int)(((x0 + 1) * G + U - 1) / U - 1), c_last = (int)(((x1 + 1) * G + U - 1) / U - 1);
        int* s_last = reinterpret_cast<int*>(tmem_slot + 1);
        if (et == 0) *s_last = (atomicAdd(p.ws_count + w, 1u) == (unsigned)(c_last - c_first)) ? 1 : 0;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) trace_stamp(p, 8);
        if (*s_last) {
          // the LAST contributor to arrive reduces the tile: partials summed in k order (fixed), then the usual epilogue
          __threadfence();
          constexpr int QUADS = GATED ? 16 : 32;           // 4 consecutive weight rows per thread (GATED: + their up rows)
          constexpr int NGRP = 128 / QUADS;                // column groups
          constexpr int CPG = TOKN / NGRP;                 // columns per group
          constexpr int CH = CPG < 8 ? CPG : 8;            // columns held in registers at a time
          const int quad = et % QUADS, grp = et / QUADS;
          const int n0 = m_tile * ROWS_OUT + quad * 4;
          float bv4[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) bv4[k] = (p.bias && n0 + k < p.N) ? bf2f(p.bias[n0 + k]) : 0.f;
          const bool vec = (p.N & 3) == 0 && n0 + 3 < p.N;
#pragma unroll 1
          for (int cb = grp * CPG; cb < (grp + 1) * CPG; cb += CH) {
            float a[CH][4], u[GATED ? CH : 1][4];
#pragma unroll
            for (int ci = 0; ci < CH; ++ci) {
              a[ci][0] = a[ci][1] = a[ci][2] = a[ci][3] = 0.f;
              if (GATED) { u[ci][0] = u[ci][1] = u[ci][2] = u[ci][3] = 0.f; }
            }
            for (int cc = c_first; cc <= c_last; ++cc) {
              const int sl = (U * cc / G >= x0) ? 0 : 1;     // that CTA's share starts inside this tile: its slot 0
              const float* src = p.ws_part + ((int64_t)(cc * 2 + sl) * TOKN + cb) * BLK_M + quad * 4;
#pragma unroll
              for (int ci = 0; ci < CH; ++ci) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(src + ci * BLK_M));
                a[ci][0] += v.x; a[ci][1] += v.y; a[ci][2] += v.z; a[ci][3] += v.w;
                if (GATED) {
                  const float4 v2 = __ldcg(reinterpret_cast<const float4*>(src + ci * BLK_M + 64));
                  u[ci][0] += v2.x; u[ci][1] += v2.y; u[ci][2] += v2.z; u[ci][3] += v2.w;
                }
              }
            }
#pragma unroll
            for (int ci = 0; ci < CH; ++ci) {
              const int t = t0 + cb + ci;
              if (t < p.T) {
                float o[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  if (GATED) o[k] = silu_gate_bf16(a[ci][k], u[ci][k]);
                  else { o[k] = a[ci][k] + bv4[k]; if (p.act == 1) o[k] = fmaxf(o[k], 0.f); }
                }
                bf16* dst = p.y + (int64_t)t * p.N + n0;
                if (vec) {
                  uint2 w2;
                  w2.x = pack_bf2(o[0], o[1]); w2.y = pack_bf2(o[2], o[3]);
                  *reinterpret_cast<uint2*>(dst) = w2;
                } else {
#pragma unroll
                  for (int k = 0; k < 4; ++k) if (n0 + k < p.N) dst[k] = f2bf(o[k]);
                }
              }
            }
          }
          if (et == 0) p.ws_count[w] = 0;                  // ready for the next launch
        }
      } else if (!SPLIT && !GATED) {
#pragma unroll 1
        for (int c = 0; c < TOKN; c += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + c, r);
          if (n < p.N) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int t = t0 + c + j;
              if (t < p.T) {
                float v = __uint_as_float(r[j]) + bv;
                if (p.act == 1) v = fmaxf(v, 0.f);
                p.y[(int64_t)t * p.N + n] = f2bf(v);
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (PAIR) mbar_arrive_cluster(mapa_u32(tempty_bar(acc), 0u)); else mbar_arrive(tempty_bar(acc)); }
      } else if (!SPLIT) {
        // Gate rows live in warps q=0,1 and their up rows in warps q=2,3 (TMEM lane quarters), so each
        // pair swaps half of every 16-column chunk through shared memory: the gate warp finishes
        // columns 0-7, the up warp columns 8-15.  Double-buffered, one 64-thread named barrier per chunk.
        const int pair = q & 1, r64 = et & 63;
        const int bar_id = 2 + pair;
        const bool is_up = q >= 2;
#pragma unroll 1
        for (int c = 0; c < TOKN; c += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + c, r);
          float* hb = xbuf + ((c >> 4) & 1) * (2 * 8 * 64);     // [0]: up cols 0-7, [1]: gate cols 8-15
          if (is_up) {
#pragma unroll
            for (int j = 0; j < 8; ++j) hb[j * 64 + r64] = __uint_as_float(r[j]);
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) hb[(8 + j) * 64 + r64] = __uint_as_float(r[8 + j]);
          }
          asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
          if (n < p.N) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int col = is_up ? 8 + j : j;
              const int t = t0 + c + col;
              const float gacc = is_up ? hb[(8 + j) * 64 + r64] : __uint_as_float(r[j]);
              const float uacc = is_up ? __uint_as_float(r[8 + j]) : hb[j * 64 + r64];
              if (t < p.T) p.y[(int64_t)t * p.N + n] = f2bf(silu_gate_bf16(gacc, uacc));
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (PAIR) mbar_arrive_cluster(mapa_u32(tempty_bar(acc), 0u)); else mbar_arrive(tempty_bar(acc)); }
      } else {
        cluster_wait();                          // phase A done: peers' shared memory is live
        // Phase 1: my 128 x TOKN fp32 partial goes to the owners' exchange buffers, [src rank][local column][128 rows]:
        // one warp-wide store is 128 contiguous bytes of one remote column.
        const uint32_t xb = smem_addr(xbuf);
        {
          int owner = 0, lc = 0;
#pragma unroll 1
          for (int c = 0; c < TOKN; c += 16) {
            uint32_t r[16];
            tmem_ld16(taddr + c, r);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const uint32_t local = xb + (uint32_t)(((crank * cols_per + lc) * BLK_M + et) * 4);
              st_cluster_f32(mapa_u32(local, (uint32_t)owner), __uint_as_float(r[j]));
              if (++lc == cols_per) { lc = 0; ++owner; }
            }
          }
        }
        tc_fence_before();
        if (et == 0) trace_stamp(p, 8);
        cluster_arrive();                        // phase B: my partials are in the owners' buffers
        cluster_wait();
        if (et == 0) trace_stamp(p, 9);
        // Phase 2: my columns, summed over the ranks in rank order.  One item = 8 consecutive weight rows of one column:
        // 32 contiguous bytes per rank in shared memory, one 16-byte store of the output row.
        const int c_begin = crank * cols_per, c_end = min(TOKN, c_begin + cols_per);
        constexpr int RG = GATED ? 8 : 16;                       // 8-row groups per column
        const int n_items = (c_end - c_begin) * RG;
        const bool vec_ok = (p.N & 7) == 0;
        for (int item = et; item < n_items; item += 128) {
          const int lc = item / RG, rg = item % RG;
          const int t = t0 + c_begin + lc;
          if (t >= p.T) break;
          float a[8], u[GATED ? 8 : 1];
#pragma unroll
          for (int k = 0; k < 8; ++k) { a[k] = 0.f; if (GATED) u[k] = 0.f; }
          for (int s2 = 0; s2 < S; ++s2) {
            const uint32_t src = xb + (uint32_t)(((s2 * cols_per + lc) * BLK_M + rg * 8) * 4);
            const float4 v0 = lds_f4(src), v1 = lds_f4(src + 16);
            a[0] += v0.x; a[1] += v0.y; a[2] += v0.z; a[3] += v0.w;
            a[4] += v1.x; a[5] += v1.y; a[6] += v1.z; a[7] += v1.w;
            if (GATED) {
              const float4 w0 = lds_f4(src + 64 * 4), w1 = lds_f4(src + 64 * 4 + 16);
              u[0] += w0.x; u[1] += w0.y; u[2] += w0.z; u[3] += w0.w;
              u[4] += w1.x; u[5] += w1.y; u[6] += w1.z; u[7] += w1.w;
            }
          }
          const int n0 = m_tile * ROWS_OUT + rg * 8;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            if (GATED) {
              a[k] = silu_gate_bf16(a[k], u[k]);
            } else {
              if (p.bias && n0 + k < p.N) a[k] += bf2f(p.bias[n0 + k]);
              if (p.act == 1) a[k] = fmaxf(a[k], 0.f);
            }
          }
          bf16* dst = p.y + (int64_t)t * p.N + n0;
          if (vec_ok && n0 + 7 < p.N) {
            uint4 o;
            o.x = pack_bf2(a[0], a[1]); o.y = pack_bf2(a[2], a[3]); o.z = pack_bf2(a[4], a[5]); o.w = pack_bf2(a[6], a[7]);
            *reinterpret_cast<uint4*>(dst) = o;
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) if (n0 + k < p.N) dst[k] = f2bf(a[k]);
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (et == 0) trace_stamp(p, 10);
  }
  if (SPLIT && warp < 4) {                       // non-epilogue warps take part in both cluster phases
    cluster_wait();
    cluster_arrive();
    cluster_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) trace_stamp(p, 11);
  if (PAIR) { cluster_arrive(); cluster_wait(); }   // the peer may still signal my barriers until it is done too
  if (warp == 2) {
    tc_fence_after();
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------------------ wide tiles for T <= 64
// A tcgen05.mma whose operands both sit in shared memory costs ~165 cycles whatever its shape (profiles/r02_summary.md
// section 3): with 128 weight rows as M and the tokens as N, a decode-sized GEMM pays four of them for every 16 KB of
// weights and its main loop, not HBM, sets the pace (330-385 ns per k-block).  Here the roles are swapped and the tile
// is wider: the tokens are the M side (64 rows loaded; the MMA is 128 tall and its upper half reads whatever follows
// the token tile in shared memory - rows that are never read back), 256 weight rows are N of ONE MMA, so the same four
// MMAs move 32 KB of weights.  D is [token][weight row] in TMEM: a thread owns a token and 256 consecutive outputs,
// which it stores 16 bytes at a time; the fused SwiGLU epilogue finds gate (columns 0-127) and up (128-255) of an
// output in the same thread.  One tile per CTA; K split over a cluster as above, the exchange buffer laid over the
// pipeline stages once every CTA of the cluster is past its main loop.
constexpr int WIDE_N = 256;                 // weight rows per tile (GATED: 128 gate rows + their 128 up rows)
constexpr int WIDE_M = 64;                  // token rows per tile
constexpr uint32_t WIDE_A_BYTES = WIDE_M * BLK_K * 2, WIDE_B_BYTES = WIDE_N * BLK_K * 2;
constexpr uint32_t WIDE_STAGE = WIDE_A_BYTES + WIDE_B_BYTES;
constexpr int kWideStages = 5;
constexpr int kWideThreads = 384;           // warps 0-3 as in gemm_tc_kernel, two epilogue warpgroups

__device__ __forceinline__ void st_cluster_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <bool SPLIT, bool GATED>
__global__ void __launch_bounds__(kWideThreads, 1)
gemm_wide_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x, const GemmParams p) {
  constexpr int STAGES = kWideStages;
  constexpr uint32_t TMEM_COLS = WIDE_N;
  // p.rows outputs per tile (a multiple of 16, up to 256; GATED: up to 128 gate rows + their up rows): the MMA's N
  const int OUT_PER_TILE = p.rows;
  const int N_MMA = GATED ? 2 * OUT_PER_TILE : OUT_PER_TILE;
  const uint32_t IDESC = (1u << 4) | KGE_IDESC_AB | ((uint32_t)(N_MMA >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  constexpr int UNIT = 16;                                     // outputs per exchange unit (GATED: 16 gate + 16 up columns)
  const int N_UNITS = OUT_PER_TILE / UNIT;
  constexpr int UNIT_F = GATED ? 2 * UNIT : UNIT;              // fp32 values of a unit
  const uint32_t tx_bytes = WIDE_A_BYTES + (uint32_t)N_MMA * BLK_K * 2;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * WIDE_STAGE);    // full[STAGES], empty[STAGES], tmem_full
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 2);
  const uint32_t smem_base = smem_addr(smem);
  const uint32_t bar_base = smem_addr(bars);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  const uint32_t tfull_bar = bar_base + 8u * (2 * STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) trace_stamp(p, 0);
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tfull_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_slot)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) trace_stamp(p, 1);
  pdl_launch_dependents();
  if (warp != 0) pdl_wait();          // warp 0 (TMA producer) prefetches weights first, then waits

  const int S = SPLIT ? (int)cluster_nctarank() : 1;
  const int crank = SPLIT ? (int)cluster_ctarank() : 0;
  const int m_tile = SPLIT ? (int)blockIdx.x / S : (int)blockIdx.x;            // one tile per CTA (cluster)
  const int kb0 = crank * p.kb_per_split, kb1 = min(p.kb_total, kb0 + p.kb_per_split);
  const int n0 = m_tile * OUT_PER_TILE;                                         // first output column of the tile

  if (warp == 0) {
    if (lane == 0) {
      auto load_w = [&](int st, int kb) {
        const uint32_t dst = smem_base + st * WIDE_STAGE + WIDE_A_BYTES;
        if (GATED) {
          tma_load_2d(dst, &tmap_w, full_bar(st), kb * BLK_K, n0);                              // the gate rows
          tma_load_2d(dst + OUT_PER_TILE * BLK_K * 2, &tmap_w, full_bar(st), kb * BLK_K, p.N + n0);   // their up rows
        } else {
          tma_load_2d(dst, &tmap_w, full_bar(st), kb * BLK_K, n0);
        }
      };
      trace_stamp(p, 12);
      const int pre = min(STAGES, kb1 - kb0);
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(full_bar(i), tx_bytes);
        load_w(i, kb0 + i);
        if (i == 0) trace_stamp(p, 13);
      }
      trace_stamp(p, 2);
      pdl_wait();
      trace_stamp(p, 3);
      int stage = 0; uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        if (kb - kb0 >= pre) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_expect_tx(full_bar(stage), tx_bytes);
          load_w(stage, kb);
        }
        tma_load_2d(smem_base + stage * WIDE_STAGE, &tmap_x, full_bar(stage), kb * BLK_K, 0);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      trace_stamp(p, 4);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      if (p.trace) { mbar_wait(full_bar(0), 0); trace_stamp(p, 5); }
      int stage = 0; uint32_t phase = 0;
      const uint64_t adesc0 = umma_desc_sw128(smem_base), bdesc0 = umma_desc_sw128(smem_base + WIDE_A_BYTES);
      uint64_t adesc = adesc0, bdesc = bdesc0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BLK_K / UMMA_K; ++k)
          tc_mma_bf16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
        tc_commit(empty_bar(stage));
        adesc += WIDE_STAGE >> 4; bdesc += WIDE_STAGE >> 4;
        if (++stage == STAGES) { stage = 0; phase ^= 1; adesc = adesc0; bdesc = bdesc0; }
      }
      tc_commit(tfull_bar);
      trace_stamp(p, 6);
    }
  }
  else if (warp >= 4) {
    // Two epilogue warpgroups (warps 4-7, 8-11): a warp reads the TMEM lane quarter warp % 4, the tokens live in lanes
    // 0-63, so warps 4, 5 take the first half of the tile's units for their tokens and warps 8, 9 the second half; all
    // eight warps share the reduction of the split variant.
    const int et = threadIdx.x - 128, q = warp & 3, eg = (warp - 4) >> 2;
    const int tok = q * 32 + lane;                          // my token (TMEM lane) if q < 2
    const int u_half = (N_UNITS + 1) / 2;
    const int u_lo = eg * u_half, u_hi = eg ? N_UNITS : u_half;
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    if (et == 0) trace_stamp(p, 7);
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const bool vec_ok = (p.N & 7) == 0;
    // the 16 outputs of unit u for my token from raw accumulators: bias / ReLU, or silu(gate) * up
    auto finish = [&](int u, const float (&a)[16], const float (&up)[16], float (&o)[16]) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (GATED) {
          o[j] = silu_gate_bf16(a[j], up[j]);
        } else {
          const int n = n0 + u * UNIT + j;
          float v = a[j] + ((p.bias && n < p.N) ? bf2f(p.bias[n]) : 0.f);
          if (p.act == 1) v = fmaxf(v, 0.f);
          o[j] = v;
        }
      }
    };
    auto store16 = [&](int t, int u, const float (&o)[16]) {
      if (t >= p.T) return;
      const int n = n0 + u * UNIT;
      bf16* dst = p.y + (int64_t)t * p.N + n;
      if (vec_ok && n + 15 < p.N) {
        uint4 w0, w1;
        w0.x = pack_bf2(o[0], o[1]); w0.y = pack_bf2(o[2], o[3]); w0.z = pack_bf2(o[4], o[5]); w0.w = pack_bf2(o[6], o[7]);
        w1.x = pack_bf2(o[8], o[9]); w1.y = pack_bf2(o[10], o[11]); w1.z = pack_bf2(o[12], o[13]); w1.w = pack_bf2(o[14], o[15]);
        reinterpret_cast<uint4*>(dst)[0] = w0;
        reinterpret_cast<uint4*>(dst)[1] = w1;
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) if (n + j < p.N) dst[j] = f2bf(o[j]);
      }
    };
    if (!SPLIT) {
      if (q < 2) {
#pragma unroll 1
        for (int u = u_lo; u < u_hi; ++u) {
          uint32_t r[16], r2[16];
          float a[16], up[16], o[16];
          tmem_ld16(taddr + u * UNIT, r);
          if (GATED) tmem_ld16(taddr + OUT_PER_TILE + u * UNIT, r2);
#pragma unroll
          for (int j = 0; j < 16; ++j) { a[j] = __uint_as_float(r[j]); up[j] = GATED ? __uint_as_float(r2[j]) : 0.f; }
          finish(u, a, up, o);
          store16(tok, u, o);
        }
      }
    } else {
      const int units_per = (N_UNITS + S - 1) / S;
      constexpr int NJ = UNIT_F / 4;                        // 16-byte pieces of a unit
      // exchange buffer of a CTA: [source rank][local unit][piece][token][4 fp32] - a warp stores (and later loads) 512
      // contiguous bytes
      const uint32_t xb = smem_base;                        // over the pipeline stages: dead once phase A completes
      cluster_arrive();                                     // phase A: my MMAs have read their last stage
      cluster_wait();
      if (q < 2) {
#pragma unroll 1
        for (int u = u_lo; u < u_hi; ++u) {
          uint32_t r[16], r2[16];
          tmem_ld16(taddr + u * UNIT, r);
          if (GATED) tmem_ld16(taddr + OUT_PER_TILE + u * UNIT, r2);
          const int owner = u / units_per, lu = u - owner * units_per;
          const uint32_t dst = mapa_u32(xb + (uint32_t)(((((crank * units_per + lu) * NJ) * WIDE_M) + tok) * 16), (uint32_t)owner);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            st_cluster_f4(dst + j * (WIDE_M * 16), __uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
          if (GATED) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_cluster_f4(dst + (4 + j) * (WIDE_M * 16), __uint_as_float(r2[4 * j]), __uint_as_float(r2[4 * j + 1]), __uint_as_float(r2[4 * j + 2]), __uint_as_float(r2[4 * j + 3]));
          }
        }
      }
      tc_fence_before();
      if (et == 0) trace_stamp(p, 8);
      cluster_arrive();                                     // phase B: my partials are in the owners' buffers
      cluster_wait();
      if (et == 0) trace_stamp(p, 9);
      // my units, summed over the ranks in rank order; one item = one token's 16 outputs of one unit
      const int u_begin = crank * units_per;
      const int my_units = min(N_UNITS, u_begin + units_per) - u_begin;
      for (int item = et; item < my_units * WIDE_M; item += 256) {
        const int t = item & (WIDE_M - 1), lu = item / WIDE_M;
        float a[16], up[16], o[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) { a[j] = 0.f; up[j] = 0.f; }
        for (int s2 = 0; s2 < S; ++s2) {
          const uint32_t src = xb + (uint32_t)(((((s2 * units_per + lu) * NJ) * WIDE_M) + t) * 16);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 v = lds_f4(src + j * (WIDE_M * 16));
            a[4 * j] += v.x; a[4 * j + 1] += v.y; a[4 * j + 2] += v.z; a[4 * j + 3] += v.w;
          }
          if (GATED) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float4 v = lds_f4(src + (4 + j) * (WIDE_M * 16));
              up[4 * j] += v.x; up[4 * j + 1] += v.y; up[4 * j + 2] += v.z; up[4 * j + 3] += v.w;
            }
          }
        }
        finish(u_begin + lu, a, up, o);
        store16(t, u_begin + lu, o);
      }
    }
    if (et == 0) trace_stamp(p, 10);
  }
  if (SPLIT && warp < 4) {                       // non-epilogue warps take part in both cluster phases
    cluster_arrive();
    cluster_wait();
    cluster_arrive();
    cluster_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) trace_stamp(p, 11);
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(ptr);
  return fn;
}

// 2-D row-major bf16 [rows, cols] with a {64 x box_rows} box and 128B swizzle.
static int encode_tmap(CUtensorMap* m, const void* ptr, int64_t rows, int64_t cols, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  KGE_REQUIRE(enc, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BLK_K, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
#ifdef KGE_HALF
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
#else
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
#endif
  CUresult r = enc(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  KGE_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld", (int)r, (long long)rows, (long long)cols);
  return 0;
}

// A tensor map is a pure function of (address, shape, box): the engine's weights and activation buffers keep their
// addresses, so each map is encoded once and looked up afterwards (two host-side driver calls per GEMM launch otherwise,
// which is what an un-graphed step pays on every one of its ~200 GEMMs).
struct TmapKey {
  const void* p; int64_t rows, cols; int box;
  bool operator==(const TmapKey& o) const { return p == o.p && rows == o.rows && cols == o.cols && box == o.box; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = (uint64_t)(uintptr_t)k.p * 0x9E3779B97F4A7C15ull;
    h ^= ((uint64_t)k.rows * 0xBF58476D1CE4E5B9ull) ^ ((uint64_t)k.cols << 20) ^ (uint64_t)k.box;
    return (size_t)(h ^ (h >> 29));
  }
};
int make_tmap(CUtensorMap* m, const void* ptr, int64_t rows, int64_t cols, int box_rows) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  const TmapKey key{ptr, rows, cols, box_rows};
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *m = it->second; return 0; }
  if (int rc = encode_tmap(m, ptr, rows, cols, box_rows)) return rc;
  if (cache.size() >= 16384) cache.clear();       // callers that cycle through fresh buffers (tests) stay bounded
  cache.emplace(key, *m);
  return 0;
}

constexpr int kGemmSmemBudget = 224 * 1024;

template <int TOKN, bool SPLIT, bool GATED, bool PAIR = false, bool WS = false>
static int launch_gemm(const CUtensorMap& tw, const CUtensorMap& tx, const GemmParams& p, cudaStream_t st) {
  constexpr int stage_bytes = BLK_M * BLK_K * 2 + (PAIR ? 128 : TOKN) * BLK_K * 2;
  // barriers + column map, then the cluster exchange buffer (SPLIT) or the up->gate hand-off (GATED)
  constexpr int extra = 2048 + (SPLIT ? xbuf_bytes(TOKN) : (GATED ? 2 * 16 * 64 * 4 : 0));
  constexpr int budget = kGemmSmemBudget - 1024 - extra;
  constexpr int STAGES = (budget / stage_bytes) > 8 ? 8 : (budget / stage_bytes);
  static_assert(STAGES >= 3, "pipeline too shallow");
  constexpr int smem = STAGES * stage_bytes + 1024 /*align*/ + extra;
  auto kern = gemm_tc_kernel<TOKN, STAGES, SPLIT, GATED, PAIR, WS>;
  KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  prefer_max_shared(kern);
  const int tiles = p.m_tiles * p.n_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_for_this_launch() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (WS) {
    cfg.gridDim = dim3(p.splits);                // stream-K: `splits` CTAs (one per SM), each a contiguous share of all k-blocks
  } else if (SPLIT) {
    cfg.gridDim = dim3(tiles * p.splits);
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = p.splits; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
  } else if (PAIR) {
    const int pairs = ((p.m_tiles + 1) / 2) * p.n_tiles;
    cfg.gridDim = dim3(2 * (pairs < 74 ? pairs : 74));
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
  } else {
    constexpr int slots = 148;
    int grid = tiles < slots ? tiles : slots;
    if (tunable(TUN_GEMM_EVEN_GRID) && tiles > 148) {
      // persistent round-robin: 224 tiles on 148 CTAs leaves 72 CTAs idle for the second half; 112 CTAs with two
      // tiles each keep every CTA streaming to the end (a weight-streaming GEMM needs bytes in flight, not SMs)
      const int rounds = (tiles + 147) / 148;
      grid = (tiles + rounds - 1) / rounds;
    }
    cfg.gridDim = dim3(grid);
  }
  KGE_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tw, tx, p));
  return 0;
}

template <int TOKN>
static int query_max_clusters(int S) {
  constexpr int stage_bytes = BLK_M * BLK_K * 2 + TOKN * BLK_K * 2;
  constexpr int extra = 2048 + xbuf_bytes(TOKN);
  constexpr int budget = kGemmSmemBudget - 1024 - extra;
  constexpr int STAGES = (budget / stage_bytes) > 8 ? 8 : (budget / stage_bytes);
  constexpr int smem = STAGES * stage_bytes + 1024 + extra;
  auto kern = gemm_tc_kernel<TOKN, STAGES, true, false>;   // the gated variant has the same footprint
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(S * 64); cfg.blockDim = dim3(kGemmThreads); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = S; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
  return n;
}

// How many clusters of size S of the split kernel can be co-resident (GPC packing decides: 148 SMs
// hold e.g. 34 clusters of 4 but fewer than 148/3 clusters of 3).  Cached per (token tile, S).
static int max_clusters(int tokn, int S) {
  static int cache[4][9];
  static bool have[4][9];
  const int ti = tokn == 16 ? 0 : tokn == 32 ? 1 : tokn == 64 ? 2 : 3;
  if (!have[ti][S]) {
    cache[ti][S] = ti == 0 ? query_max_clusters<16>(S) : ti == 1 ? query_max_clusters<32>(S)
                 : ti == 2 ? query_max_clusters<64>(S) : query_max_clusters<128>(S);
    have[ti][S] = true;
  }
  return cache[ti][S];
}

struct GemmPlan { int tokn, m_tiles, n_tiles, splits, kb_total, kb_per_split; bool ws; int rows; };

constexpr int64_t kWsCounterBytes = 1024;        // per-tile arrival counters at the head of the workspace (<= 256 tiles)

// ws_ok: the caller can provide a workspace, so split-K may go through it (two CTAs per SM) instead of a cluster
static GemmPlan plan_gemm(int T, int N, int K, int rows_out = BLK_M, bool ws_ok = false) {
  GemmPlan g;
  g.ws = false;
  g.tokn = T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256;
  g.m_tiles = (N + rows_out - 1) / rows_out;
  g.n_tiles = (T + g.tokn - 1) / g.tokn;
  g.kb_total = (K + BLK_K - 1) / BLK_K;
  const int tiles = g.m_tiles * g.n_tiles;
  int splits = 1;
  const int slots = 148;                       // CTAs that can be resident at once
  if (ws_ok && g.tokn <= 128 && tiles <= (int)(kWsCounterBytes / 4) && tiles % 148 != 0 && tunable(TUN_GEMM_WS)) {
    // stream-K: one CTA per SM, every CTA the same number of k-blocks (at least 4, i.e. 256 of K)
    const int64_t units = (int64_t)tiles * g.kb_total;
    int ctas = units / 4 < 148 ? (int)(units / 4) : 148;
    if (ctas >= 2) { g.ws = true; splits = ctas; }
  }
  if (!g.ws && tiles <= slots / 2 && g.tokn <= 128) {
    int smax = slots / tiles;                   // never more CTAs than resident slots
    const int max_by_k = g.kb_total / 4 > 0 ? g.kb_total / 4 : 1;   // >= 4 k-blocks (256 of K) per split
    if (smax > max_by_k) smax = max_by_k;
    if (smax > 8) smax = 8;                     // portable cluster size
    for (int s = smax; s >= 2; --s)             // largest S whose clusters all fit in ONE wave
      if (tiles <= max_clusters(g.tokn, s)) { splits = s; break; }
  }
  if (g.ws) { g.kb_per_split = g.kb_total; g.splits = splits; return g; }      // `splits` carries the grid size
  g.kb_per_split = (g.kb_total + splits - 1) / splits;
  g.splits = (g.kb_total + g.kb_per_split - 1) / g.kb_per_split;   // drop empty splits
  return g;
}

// ---- wide tiles (T <= 64): one tile of 256 weight rows per CTA / cluster
constexpr int kWideSmem = kWideStages * (int)WIDE_STAGE + 1024 /*align*/ + 2048 /*barriers*/;

template <bool SPLIT, bool GATED>
static int launch_wide(const CUtensorMap& tw, const CUtensorMap& tx, const GemmParams& p, cudaStream_t st) {
  auto kern = gemm_wide_kernel<SPLIT, GATED>;
  KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideSmem));
  prefer_max_shared(kern);
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kWideThreads);
  cfg.dynamicSmemBytes = kWideSmem;
  cfg.stream = st;
  cfg.gridDim = dim3(p.m_tiles * p.splits);
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_for_this_launch() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (SPLIT) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = p.splits; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
  }
  KGE_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tw, tx, p));
  return 0;
}

static int max_clusters_wide(int S) {
  static int cache[9];
  static bool have[9];
  if (!have[S]) {
    auto kern = gemm_wide_kernel<true, false>;
    int n = 0;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideSmem) == cudaSuccess) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(S * 64); cfg.blockDim = dim3(kWideThreads); cfg.dynamicSmemBytes = kWideSmem;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = S; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr; cfg.numAttrs = 1;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { (void)cudaGetLastError(); n = 0; }
    }
    cache[S] = n; have[S] = true;
  }
  return cache[S];
}

// Tile width (outputs per tile, a multiple of 16: up to 256, or 128 for the fused gate/up) and K split of the wide
// kernel: full-width tiles with the largest cluster split that fits one round; where no split is possible (more than
// 74 tiles) narrower tiles, down to 3/4 of the width, if that puts more CTAs to work (gate_up of Llama-8B: 128 tiles of
// 112 outputs instead of 112 of 128).  A cost model over all widths and splits (k-block = max(MMA time, bytes at the
// per-SM rate), price per split) picked plans that were faster alone and 0.35 ms per decode step SLOWER in the engine
// (clusters of 3 and 5 next to the neighbouring kernels' CTAs); it was removed (profiles/r02_summary.md section 7).
// False: this problem keeps the 128-row kernels.
static bool plan_wide(int T, int n_out, int K, bool gated, GemmPlan& g) {
  if (!tunable(TUN_GEMM_WIDE) || T > WIDE_M) return false;
  const int kb_total = (K + BLK_K - 1) / BLK_K;
  const int max_rows = gated ? WIDE_N / 2 : WIDE_N;
  const int max_by_k = kb_total / 4 > 0 ? kb_total / 4 : 1;   // >= 4 k-blocks (256 of K) per split
  double best = 1e30;
  int best_rows = 0, best_s = 0;
  auto try_rows = [&](int rows) {               // largest split that fits one round; more CTAs win, then the wider tile
    const int tiles = (n_out + rows - 1) / rows;
    if (tiles > 148) return;
    int sp = 148 / tiles;
    if (sp > max_by_k) sp = max_by_k;
    if (sp > 8) sp = 8;
    while (sp > 1 && tiles > max_clusters_wide(sp)) --sp;
    const double score = -(double)(tiles * sp) - rows * 1e-3;
    if (score < best) { best = score; best_rows = rows; best_s = sp; }
  };
  try_rows(max_rows);
  if (best_s == 1)                              // no split possible: spread the rows over more CTAs instead
    for (int rows = max_rows - 16; rows >= (3 * max_rows) / 4; rows -= 16) try_rows(rows);
  if (best_rows == 0) return false;
  g.ws = false;
  g.tokn = WIDE_M; g.n_tiles = 1;
  g.rows = best_rows;
  g.m_tiles = (n_out + best_rows - 1) / best_rows;
  g.kb_total = kb_total;
  g.kb_per_split = (kb_total + best_s - 1) / best_s;
  g.splits = (kb_total + g.kb_per_split - 1) / g.kb_per_split;
  // only where it keeps at least as many CTAs streaming as the 128-row plan (measured: otherwise it loses)
  const GemmPlan narrow = plan_gemm(T, n_out, K, gated ? 64 : BLK_M, false);
  const int narrow_tiles = narrow.m_tiles * narrow.n_tiles;
  int narrow_ctas = narrow_tiles * narrow.splits;
  if (narrow_tiles > 148) {                                // persistent: the grid launch_gemm evens out
    const int rounds = (narrow_tiles + 147) / 148;
    narrow_ctas = tunable(TUN_GEMM_EVEN_GRID) ? (narrow_tiles + rounds - 1) / rounds : 148;
  }
  return g.m_tiles * g.splits >= narrow_ctas;
}

// prefill-sized problems run as CTA pairs (cta_group::2, 256 x 256 MMA tiles)
static bool gemm_uses_pair(const GemmPlan& g) { return g.tokn == 256 && g.splits == 1 && g.m_tiles >= 2; }

template <bool GATED>
static int dispatch_gemm(const GemmPlan& g, const CUtensorMap& tw, const CUtensorMap& tx, const GemmParams& p, cudaStream_t st) {
  if (g.ws) {
    switch (g.tokn) {
      case 16: return launch_gemm<16, false, GATED, false, true>(tw, tx, p, st);
      case 32: return launch_gemm<32, false, GATED, false, true>(tw, tx, p, st);
      case 64: return launch_gemm<64, false, GATED, false, true>(tw, tx, p, st);
      default: return launch_gemm<128, false, GATED, false, true>(tw, tx, p, st);
    }
  }
  if (g.splits > 1) {
    switch (g.tokn) {
      case 16: return launch_gemm<16, true, GATED>(tw, tx, p, st);
      case 32: return launch_gemm<32, true, GATED>(tw, tx, p, st);
      case 64: return launch_gemm<64, true, GATED>(tw, tx, p, st);
      default: return launch_gemm<128, true, GATED>(tw, tx, p, st);
    }
  }
  switch (g.tokn) {
    case 16: return launch_gemm<16, false, GATED>(tw, tx, p, st);
    case 32: return launch_gemm<32, false, GATED>(tw, tx, p, st);
    case 64: return launch_gemm<64, false, GATED>(tw, tx, p, st);
    case 128: return launch_gemm<128, false, GATED>(tw, tx, p, st);
    default:
      if (gemm_uses_pair(g)) return launch_gemm<256, false, GATED, true>(tw, tx, p, st);
      return launch_gemm<256, false, GATED>(tw, tx, p, st);
  }
}

}  // namespace KGE_ELT_NS
}  // namespace kge

using namespace kge;

#ifndef KGE_HALF
unsigned long long* g_gemm_trace = nullptr;
// Measurement hook of tools/gemm_trace.py: device buffer of 16 timestamps per CTA of the following GEMM launches
// (both element types), or null to stop.  No counterpart in the reference.
extern "C" void kge_gemm_set_trace(void* device_buffer) { g_gemm_trace = (unsigned long long*)device_buffer; }
#else
extern unsigned long long* g_gemm_trace;
#endif

static int64_t ws_bytes_of(const GemmPlan& g) {        // counters + two partial tiles per CTA
  return g.ws ? kWsCounterBytes + (int64_t)g.splits * 2 * g.tokn * BLK_M * 4 : 0;
}

#ifndef KGE_HALF      // the same for both element types: exported once
extern "C" int64_t kge_gemm_workspace_bytes(int T, int N, int K) {
  if (T <= 0 || N <= 0 || K <= 0) return 0;
  // enough for the plain projection and for the fused SiLU-gate one with f = N (64 output rows per tile); 0: this shape
  // does not go through a workspace
  const int64_t a = ws_bytes_of(plan_gemm(T, N, K, BLK_M, true)), b = ws_bytes_of(plan_gemm(T, N, K, 64, true));
  return a > b ? a : b;
}

#endif

#ifdef KGE_HALF
#define KGE_GEMM_NAME kge_gemm_f16
#define KGE_GEMM_GATE_NAME kge_gemm_silu_gate_f16
#define KGE_GEMM_GATE_WS_NAME kge_gemm_silu_gate_ws_f16
#else
#define KGE_GEMM_NAME kge_gemm_bf16
#define KGE_GEMM_GATE_NAME kge_gemm_silu_gate_bf16
#define KGE_GEMM_GATE_WS_NAME kge_gemm_silu_gate_ws_bf16
#endif
extern "C" int KGE_GEMM_NAME(void* y, const void* x, const void* w, const void* bias, int T, int N, int K,
                             int act, void* workspace, int64_t workspace_bytes, void* stream) {
  KGE_REQUIRE(y && x && w, "kge_gemm_bf16: null pointer");
  KGE_REQUIRE(T >= 0 && N > 0 && K > 0 && K % 8 == 0, "kge_gemm_bf16: K=%d must be a positive multiple of 8", K);
  KGE_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)w & 15) == 0, "kge_gemm_bf16: x and w must be 16-byte aligned");
  KGE_REQUIRE(act == 0 || act == 1, "kge_gemm_bf16: act=%d unsupported", act);
  if (T == 0) return 0;
  bool use_ws = workspace && tunable(TUN_GEMM_WS);          // stream-K through the caller's scratch, if it is large enough
  if (use_ws) {
    const GemmPlan gq = plan_gemm(T, N, K, BLK_M, true);
    use_ws = gq.ws && workspace_bytes >= ws_bytes_of(gq) && ((uintptr_t)workspace & 15) == 0;
  }
  if (!use_ws) {
    GemmPlan gw;
    if (plan_wide(T, N, K, false, gw)) {
      GemmParams p;
      p.y = (bf16*)y; p.bias = (const bf16*)bias; p.T = T; p.N = N; p.K = K; p.act = act;
      p.l2_hints = 0; p.trace = g_gemm_trace; p.ws_count = nullptr; p.ws_part = nullptr;
      p.m_tiles = gw.m_tiles; p.n_tiles = 1; p.splits = gw.splits; p.kb_total = gw.kb_total; p.kb_per_split = gw.kb_per_split;
      CUtensorMap tw, tx;
      p.rows = gw.rows;
      if (int rc = make_tmap(&tw, w, N, K, gw.rows)) return rc;
      if (int rc = make_tmap(&tx, x, T, K, WIDE_M)) return rc;
      return gw.splits > 1 ? launch_wide<true, false>(tw, tx, p, (cudaStream_t)stream) : launch_wide<false, false>(tw, tx, p, (cudaStream_t)stream);
    }
  }
  GemmPlan g = plan_gemm(T, N, K, BLK_M, workspace != nullptr);
  if (g.ws && (workspace_bytes < ws_bytes_of(g) || ((uintptr_t)workspace & 15) != 0))
    g = plan_gemm(T, N, K);                        // scratch too small for this shape: split K inside a cluster instead
  GemmParams p;
  p.y = (bf16*)y; p.bias = (const bf16*)bias; p.T = T; p.N = N; p.K = K; p.act = act;
  p.l2_hints = tunable(TUN_GEMM_L2_HINTS);
  p.trace = g_gemm_trace;
  p.ws_count = g.ws ? (unsigned int*)workspace : nullptr;
  p.ws_part = g.ws ? (float*)((uint8_t*)workspace + kWsCounterBytes) : nullptr;
  p.m_tiles = g.m_tiles; p.n_tiles = g.n_tiles; p.splits = g.splits;
  p.kb_total = g.kb_total; p.kb_per_split = g.kb_per_split;
  CUtensorMap tw, tx;
  if (int rc = make_tmap(&tw, w, N, K, BLK_M)) return rc;
  if (int rc = make_tmap(&tx, x, T, K, gemm_uses_pair(g) ? 128 : g.tokn)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  return dispatch_gemm<false>(g, tw, tx, p, st);
}

extern "C" int KGE_GEMM_GATE_WS_NAME(void* y, const void* x, const void* w_gate_up, int T, int f, int K, void* workspace,
                                    int64_t workspace_bytes, void* stream) {
  KGE_REQUIRE(y && x && w_gate_up, "kge_gemm_silu_gate_bf16: null pointer");
  KGE_REQUIRE(T >= 0 && f > 0 && K > 0 && K % 8 == 0, "kge_gemm_silu_gate_bf16: K=%d must be a positive multiple of 8", K);
  KGE_REQUIRE(f % 64 == 0, "kge_gemm_silu_gate_bf16: f=%d must be a multiple of 64", f);
  KGE_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)w_gate_up & 15) == 0, "kge_gemm_silu_gate_bf16: x and w must be 16-byte aligned");
  if (T == 0) return 0;
  bool use_ws = workspace && tunable(TUN_GEMM_WS);
  if (use_ws) {
    const GemmPlan gq = plan_gemm(T, f, K, 64, true);
    use_ws = gq.ws && workspace_bytes >= ws_bytes_of(gq) && ((uintptr_t)workspace & 15) == 0;
  }
  if (!use_ws) {
    GemmPlan gw;
    if (plan_wide(T, f, K, true, gw)) {
      GemmParams p;
      p.y = (bf16*)y; p.bias = nullptr; p.T = T; p.N = f; p.K = K; p.act = 2;
      p.l2_hints = 0; p.trace = g_gemm_trace; p.ws_count = nullptr; p.ws_part = nullptr;
      p.m_tiles = gw.m_tiles; p.n_tiles = 1; p.splits = gw.splits; p.kb_total = gw.kb_total; p.kb_per_split = gw.kb_per_split;
      CUtensorMap tw, tx;
      p.rows = gw.rows;
      if (int rc = make_tmap(&tw, w_gate_up, 2 * (int64_t)f, K, gw.rows)) return rc;
      if (int rc = make_tmap(&tx, x, T, K, WIDE_M)) return rc;
      return gw.splits > 1 ? launch_wide<true, true>(tw, tx, p, (cudaStream_t)stream) : launch_wide<false, true>(tw, tx, p, (cudaStream_t)stream);
    }
  }
  GemmPlan g = plan_gemm(T, f, K, 64, workspace != nullptr);
  if (g.ws && (workspace_bytes < ws_bytes_of(g) || ((uintptr_t)workspace & 15) != 0)) g = plan_gemm(T, f, K, 64);
  GemmParams p;
  p.y = (bf16*)y; p.bias = nullptr; p.T = T; p.N = f; p.K = K; p.act = 2;
  p.l2_hints = tunable(TUN_GEMM_L2_HINTS);
  p.trace = g_gemm_trace;
  p.ws_count = g.ws ? (unsigned int*)workspace : nullptr;
  p.ws_part = g.ws ? (float*)((uint8_t*)workspace + kWsCounterBytes) : nullptr;
  p.m_tiles = g.m_tiles; p.n_tiles = g.n_tiles; p.splits = g.splits;
  p.kb_total = g.kb_total; p.kb_per_split = g.kb_per_split;
  CUtensorMap tw, tx;
  if (int rc = make_tmap(&tw, w_gate_up, 2 * (int64_t)f, K, 64)) return rc;     // 64-row boxes: gate half, up half
  if (int rc = make_tmap(&tx, x, T, K, gemm_uses_pair(g) ? 128 : g.tokn)) return rc;
  return dispatch_gemm<true>(g, tw, tx, p, (cudaStream_t)stream);
}

extern "C" int KGE_GEMM_GATE_NAME(void* y, const void* x, const void* w_gate_up, int T, int f, int K, void* stream) {
  return KGE_GEMM_GATE_WS_NAME(y, x, w_gate_up, T, f, K, nullptr, 0, stream);
}
