// Prefill / mixed-batch paged attention on the 5th-gen tensor cores.
//
// One CTA = 128 query rows ((token, q-head-in-group) pairs of one kv head of one request) x the
// causal prefix of that request, walked in 64-key tiles (two CTAs per SM):
//   S  = Q K^T      tcgen05.mma, A = Q tile, B = K tile (both K-major, 128B swizzle), D in TMEM (2 buffers)
//   P  = softmax    4 warps, ONE THREAD PER ROW: tcgen05.ld its S row, scale + causal mask + running
//                   max/sum in fp32 (no shuffles), bf16 P written straight into the swizzled A-operand
//                   layout in shared memory
//   O_t = P V       tcgen05.mma, A = P, B = V tile as stored ([key][hd] = MN-major), D in TMEM [128,128+hd)
//   O accumulates in TMEM across tiles; it is rescaled (tcgen05.ld / st) only when a row's running maximum has grown
//   by more than 2^8 since its reference maximum was fixed (lazy rescale, see kRescaleThreshold below)
// K/V tiles (TMA = true, the default): a producer thread fetches every tile from the paged cache with
// cp.async.bulk.tensor.2d boxes of {64 columns, block_size rows} (the cache of a layer seen as a 2-D tensor [rows, hd],
// 128-byte swizzle = the layout the MMA descriptors expect), two tiles ahead for K and for V, completion counted by
// mbarriers; tcgen05.commit tells it when a stage is free.  The softmax warps issue no loads.
// TMA = false (knob "attn_tma" 0, block sizes below 8): the round-1 path, tiles gathered by the softmax threads with
// 16-byte cp.async (K two tiles ahead, V one; block-table lookup once per row) and published with fence.proxy.async.
// Tensor-bound: 4 * hd * sum_b q_b (ctx0_b + (q_b+1)/2) flop per q head.
#include <cuda.h>
#include <string.h>

#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

namespace atc {

constexpr int BM = 128, BNK = 64;              // query rows, keys per tile (64: two CTAs fit one SM)
constexpr uint32_t kSpin = 1u << 26;   // ~seconds: a lost arrival traps instead of hanging the GPU
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ uint32_t s_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
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
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void cp16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// shared-memory operand tiles: [64-element column half][rows][128 B], 16-byte chunks XOR-swizzled by
// row % 8 (SWIZZLE_128B); `rows` = 128 for every tile here
template <int ROWS>
__device__ __forceinline__ uint32_t tile_off(int row, int chunk /* 16-byte chunk of the full row */) {
  return (uint32_t)((chunk >> 3) * (ROWS * 128) + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}
// K-major operand (A, or B = K tile): rows of 128 B, 8-row atoms 1024 B apart
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
// MN-major B operand (V tile [key][hd]): 64-wide hd halves BNK*128 B apart (LBO), 8-key atoms 1024 B apart (SBO)
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((BNK * 128) >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// Pipeline (per 128-key tile j): the MMA thread issues S_{j+1} = Q K_{j+1}^T into the other S buffer
// as soon as K_{j+1} has landed, so the tensor core works on it while the softmax warps turn S_j
// into P_j; PV_j accumulates straight into the O columns of TMEM.  O is only rescaled when a row's
// maximum grows by more than 2^8 since its reference maximum was fixed ("lazy rescale": P stays
// <= 256, well inside bf16/fp32 range), so most tiles never touch O from the softmax side.
constexpr float kRescaleThreshold = 8.0f;
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int HD, bool TMA>
__global__ void __launch_bounds__(TMA ? 192 : 160, 2)
attn_tc_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
               bf16* __restrict__ out, const bf16* __restrict__ q, int q_stride, const bf16* __restrict__ k_cache,
               const bf16* __restrict__ v_cache, const int32_t* __restrict__ block_table, int max_blocks,
               const int32_t* __restrict__ seq_lens, const int32_t* __restrict__ qsl, int H, int KV, int bs_shift,
               float scale_log2) {
  constexpr int CH = HD / 8;                       // 16-byte chunks per row
  constexpr uint32_t Q_BYTES = 128 * HD * 2;
  constexpr uint32_t TILE_BYTES = BNK * HD * 2;    // K or V tile
  constexpr uint32_t P_BYTES = 128 * BNK * 2;
  constexpr uint32_t TMEM_COLS = 256;              // S0 [0,64), S1 [64,128), O [128,128+HD)
  constexpr uint32_t O_COL = 128;
  // S: M=128, N=128, both K-major.  PV: M=128, N=HD, A K-major, B MN-major (bit 16).
  constexpr uint32_t IDESC_S = (1u << 4) | KGE_IDESC_AB | ((uint32_t)(BNK >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
  constexpr uint32_t IDESC_O = (1u << 4) | KGE_IDESC_AB | (1u << 16) | ((uint32_t)(HD >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

  const int b = blockIdx.z, kvh = blockIdx.y;
  const int G = H / KV;
  const int TT = BM / G;
  const int q0 = qsl[b], q_len = qsl[b + 1] - q0;
  // heaviest (latest) query tiles first: their CTAs walk the longest causal prefix
  const int tok0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * TT;
  if (tok0 >= q_len) return;
  const int seq_len = seq_lens[b];
  const int ctx0 = seq_len - q_len;
  const int ntok = min(TT, q_len - tok0);
  const int kv_len = ctx0 + tok0 + ntok;
  const int ntiles = (kv_len + BNK - 1) / BNK;
  const int32_t* btab = block_table + (int64_t)b * max_blocks;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;                        // no static shared memory: the dynamic window starts 1024-aligned
  if ((s_addr(smem) & 1023u) != 0) __trap();
  const uint32_t sQ = s_addr(smem);
  const uint32_t sK = sQ + Q_BYTES;                // 2 stages
  const uint32_t sV = sK + 2 * TILE_BYTES;         // 2 stages
  const uint32_t sP = sV + 2 * TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Q_BYTES + 4 * TILE_BYTES + P_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
  const uint32_t bar0 = s_addr(bars);
  // k_full[2], v_full[2], s_full[2], p_full, o_full, q_full, v_empty[2]
  auto k_full = [&](int s) { return bar0 + 8u * s; };
  auto v_full = [&](int s) { return bar0 + 16u + 8u * s; };
  auto s_full = [&](int s) { return bar0 + 32u + 8u * s; };
  const uint32_t p_full = bar0 + 48, o_full = bar0 + 56, q_full = bar0 + 64;
  auto v_empty = [&](int s) { return bar0 + 72u + 8u * s; };

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (warp == 4) {
    if (lane == 0) {
      // TMA: the producer's expect_tx is the one arrival of a K/V barrier; cp.async: every loading thread arrives
      mbar_init(k_full(0), TMA ? 1 : 128); mbar_init(k_full(1), TMA ? 1 : 128);
      mbar_init(v_full(0), TMA ? 1 : 128); mbar_init(v_full(1), TMA ? 1 : 128);
      mbar_init(s_full(0), 1); mbar_init(s_full(1), 1);
      mbar_init(p_full, 128); mbar_init(o_full, 1);
      mbar_init(q_full, 128); mbar_init(v_empty(0), 1); mbar_init(v_empty(1), 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      if (TMA) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
      }
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_addr(tmem_slot)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem = *tmem_slot;
  // TMEM is held: only now may the next kernel's CTAs arrive on this SM (they would otherwise
  // compete for TMEM columns with a kernel they have to wait for)
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 4) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      auto issue_s = [&](int j) {
        const int st = j & 1;
        mbar_wait(k_full(st), (uint32_t)((j >> 1) & 1));
        tc_after();
        const uint32_t kt = sK + st * TILE_BYTES;
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) {
          const uint32_t qoff = (uint32_t)((k >> 2) * (128 * 128) + (k & 3) * 32);
          const uint32_t koff = (uint32_t)((k >> 2) * (BNK * 128) + (k & 3) * 32);
          tc_mma(tmem + st * BNK, desc_kmajor(sQ + qoff), desc_kmajor(kt + koff), IDESC_S, k > 0 ? 1u : 0u);
        }
        tc_commit(s_full(st));
      };
      if (TMA) { mbar_wait(q_full, 0u); tc_after(); }      // the Q tile (cp.async by the softmax threads) has landed
      issue_s(0);
      for (int j = 0; j < ntiles; ++j) {
        // S_{j+1} may overwrite S buffer (j+1)&1 = the buffer of S_{j-1}: the softmax warps finished
        // reading it before they arrived on p_full(j-1), which this thread has already waited for
        if (j + 1 < ntiles) issue_s(j + 1);
        mbar_wait(p_full, (uint32_t)(j & 1));
        mbar_wait(v_full(j & 1), (uint32_t)((j >> 1) & 1));
        tc_after();
        const uint32_t vt = sV + (j & 1) * TILE_BYTES;
#pragma unroll
        for (int k = 0; k < BNK / 16; ++k) {
          const uint32_t aoff = (uint32_t)((k >> 2) * (128 * 128) + (k & 3) * 32);   // P: keys are the K dimension (one 64-key half)
          const uint32_t boff = (uint32_t)(k * 16 * 128);                             // V: 16 keys = 16 rows of 128 B
          tc_mma(tmem + O_COL, desc_kmajor(sP + aoff), desc_mnmajor(vt + boff), IDESC_O, (j > 0 || k > 0) ? 1u : 0u);
        }
        tc_commit(o_full);
        if (TMA) tc_commit(v_empty(j & 1));      // PV_j has read V stage j&1: the producer may refill it
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------------ TMA producer (TMA only: 192 threads)
    // One thread: per tile it reads the tile's block ids and issues one box per (block, 64-column half).  (Spreading the
    // issue over several lanes with the block ids handed out by shuffle measured SLOWER here: 364 vs 471 TFLOP/s at
    // 4 x 2048, profiles/r02_vs_libs_*.json; a prefill tile lasts long enough for one thread to stay ahead.)
    if (lane == 0) {
      const int bs = 1 << bs_shift;
      const int box_rows = bs < BNK ? bs : BNK;
      const int boxes = BNK / box_rows;
      auto fetch = [&](int t, bool is_k) {
        const int st = t & 1;
        const uint32_t dst = (is_k ? sK : sV) + st * TILE_BYTES;
        const uint32_t bar = is_k ? k_full(st) : v_full(st);
        mbar_expect_tx(bar, TILE_BYTES);
        for (int i = 0; i < boxes; ++i) {
          const int key = t * BNK + i * box_rows;
          const int bi = key >> bs_shift;
          const int blk = bi < max_blocks ? btab[bi] : 0;      // past the table: block 0 (those keys are masked)
          const int row = (blk * KV + kvh) * bs + (key & (bs - 1));
#pragma unroll
          for (int h = 0; h < HD / 64; ++h)
            tma_load_2d(dst + (uint32_t)(h * (BNK * 128) + i * box_rows * 128), is_k ? &tm_k : &tm_v, bar, h * 64, row);
        }
      };
      fetch(0, true);
      if (ntiles > 1) fetch(1, true);
      fetch(0, false);
      if (ntiles > 1) fetch(1, false);
      for (int j = 0; j + 2 < ntiles; ++j) {
        mbar_wait(s_full(j & 1), (uint32_t)((j >> 1) & 1));      // S_j done: K stage j&1 is free
        fetch(j + 2, true);
        mbar_wait(v_empty(j & 1), (uint32_t)((j >> 1) & 1));     // PV_j done: V stage j&1 is free
        fetch(j + 2, false);
      }
    }
  } else {
    // ------------------------------------------------------------------ loaders + softmax (one row per thread)
    const int r = tid;
    const int tk = r / G, hh = r % G;
    const bool row_ok = tk < ntok;
    const int lim = row_ok ? ctx0 + tok0 + tk : -1;          // last visible key of this row
    constexpr int RP = 128 / CH, NP = BNK / RP, NPQ = 128 / RP;
    const int lrow0 = tid / CH, lc = tid % CH;
    const int64_t head_off = ((int64_t)kvh << bs_shift) * HD;
    const int64_t blk_stride = ((int64_t)KV << bs_shift) * HD;
    const int bs_mask = (1 << bs_shift) - 1;
    // cp.async groups are committed in a fixed rhythm (K two tiles ahead, V one tile ahead), so
    // "at most 2 groups pending" always means the tile about to be published has landed
    auto load_tile = [&](int t, bool is_k) {
      if (t < ntiles) {
        const uint32_t dt = (is_k ? sK : sV) + (t & 1) * TILE_BYTES;
        const bf16* cache = is_k ? k_cache : v_cache;
#pragma unroll
        for (int i = 0; i < NP; ++i) {
          const int row = lrow0 + i * RP;
          const int key = t * BNK + row;
          const bool ok = key < kv_len;
          int64_t off = 0;
          if (ok) off = (int64_t)btab[key >> bs_shift] * blk_stride + head_off + (key & bs_mask) * HD + lc * 8;
          cp16(dt + tile_off<BNK>(row, lc), cache + off, ok);
        }
      }
      cp_commit();
    };
#pragma unroll
    for (int i = 0; i < NPQ; ++i) {
      const int row = lrow0 + i * RP;
      const int t2 = row / G, h2 = row % G;
      const bool ok = t2 < ntok;
      const bf16* src = ok ? q + (int64_t)(q0 + tok0 + t2) * q_stride + (kvh * G + h2) * HD + lc * 8 : q;
      cp16(sQ + tile_off<128>(row, lc), src, ok);
    }
    if (TMA) {
      cp_commit();
      cp_wait<0>();              // this thread's part of the Q tile
      fence_async_smem();
      mbar_arrive(q_full);
    } else {
      // commit order K0 | K1 | V0 | K2 | V1 | K3 | V2 ...: whenever a tile is published, exactly two
      // younger groups may still be in flight
      load_tile(0, true);        // group 0: Q + K0 (what S_0 needs)
      load_tile(1, true);        // group 1: K1
      load_tile(0, false);       // group 2: V0
      cp_wait<2>();              // group 0 landed
      fence_async_smem();
      mbar_arrive(k_full(0));
    }

    float m_ref = -INFINITY, l = 0.f;
    const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);

    for (int j = 0; j < ntiles; ++j) {
      mbar_wait(s_full(j & 1), (uint32_t)((j >> 1) & 1));
      tc_after();
      if (!TMA) {
        // S_j is complete, so K stage j&1 is free: fetch K_{j+2}; then publish K_{j+1} (two groups back)
        load_tile(j + 2, true);
        if (j + 1 < ntiles) { cp_wait<2>(); fence_async_smem(); mbar_arrive(k_full((j + 1) & 1)); }
      }
      const int kbase = j * BNK;
      uint32_t sv[BNK];
#pragma unroll
      for (int c = 0; c < BNK; c += 32) tmem_ld32(t_row + (j & 1) * BNK + c, *reinterpret_cast<uint32_t(*)[32]>(&sv[c]));
      // The softmax warps are issue-bound: keep the per-score work to max, one FMA, one MUFU, one add.  Scores stay
      // unscaled (scale > 0 keeps the order); the scale is folded into the exponent's FMA.  Only tiles that reach
      // past some row's last visible key (the diagonal) pay for masking.
      // four independent running maxima (and, below, sums): with two softmax warps per scheduler there is little
      // else to hide the latency of a 64-long dependent chain behind
      float mxa[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (__all_sync(0xffffffffu, kbase + BNK - 1 <= lim)) {
#pragma unroll
        for (int i = 0; i < BNK; ++i) mxa[i & 3] = fmaxf(mxa[i & 3], __uint_as_float(sv[i]));
      } else {
#pragma unroll
        for (int i = 0; i < BNK; ++i) {
          const float x = (kbase + i <= lim) ? __uint_as_float(sv[i]) : -INFINITY;
          sv[i] = __float_as_uint(x);
          mxa[i & 3] = fmaxf(mxa[i & 3], x);
        }
      }
      const float mx_raw = fmaxf(fmaxf(mxa[0], mxa[1]), fmaxf(mxa[2], mxa[3]));
      const float mx = mx_raw * scale_log2;                     // -inf stays -inf
      // PV_{j-1} must be done before O is rescaled, sP is overwritten or V stage (j+1)&1 is refilled
      if (j > 0) { mbar_wait(o_full, (uint32_t)((j - 1) & 1)); tc_after(); }
      if (!TMA) load_tile(j + 1, false);
      const bool grow = mx > m_ref + kRescaleThreshold;        // includes the first finite maximum
      if (__any_sync(0xffffffffu, grow)) {
        float alpha = 1.f;
        if (grow) {
          alpha = (m_ref == -INFINITY) ? 0.f : exp2f(m_ref - mx);
          m_ref = mx;
          l *= alpha;
        }
        if (j > 0) {
#pragma unroll 1
          for (int c = 0; c < HD; c += 32) {
            uint32_t v[32];
            tmem_ld32(t_row + O_COL + c, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st32(t_row + O_COL + c, v);
          }
        }
      }
      const float base = (m_ref == -INFINITY) ? 0.f : m_ref;
      float rsa[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c8 = 0; c8 < BNK / 8; ++c8) {
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float p0 = ex2_ftz(fmaf(__uint_as_float(sv[c8 * 8 + i * 2]), scale_log2, -base));       // masked: -inf -> 0
          const float p1 = ex2_ftz(fmaf(__uint_as_float(sv[c8 * 8 + i * 2 + 1]), scale_log2, -base));
          rsa[i] += p0 + p1;
          pk[i] = pack_bf2(p0, p1);
        }
        const uint32_t dst = sP + tile_off<128>(r, c8);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
      }
      l += (rsa[0] + rsa[1]) + (rsa[2] + rsa[3]);
      if (!TMA) cp_wait<2>();             // V_j (committed one tile ago) has landed
      fence_async_smem();                 // P (and, cp.async path, V_j) bytes -> tensor-core proxy
      tc_before();                        // my TMEM reads/writes are ordered before the MMAs that follow
      if (!TMA) mbar_arrive(v_full(j & 1));
      mbar_arrive(p_full);
    }
    if (!TMA) cp_wait<0>();
    mbar_wait(o_full, (uint32_t)((ntiles - 1) & 1));
    tc_after();
    {
      // tcgen05.ld is warp-collective (.sync.aligned): every lane loads, only rows with a token store
      const float inv = l > 0.f ? 1.f / l : 0.f;
      bf16* dst = out + (int64_t)(q0 + tok0 + (row_ok ? tk : 0)) * H * HD + (kvh * G + hh) * HD;
#pragma unroll 1
      for (int c = 0; c < HD; c += 32) {
        uint32_t v[32];
        tmem_ld32(t_row + O_COL + c, v);
        if (row_ok) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 w;
            w.x = pack_bf2(__uint_as_float(v[i]) * inv, __uint_as_float(v[i + 1]) * inv);
            w.y = pack_bf2(__uint_as_float(v[i + 2]) * inv, __uint_as_float(v[i + 3]) * inv);
            w.z = pack_bf2(__uint_as_float(v[i + 4]) * inv, __uint_as_float(v[i + 5]) * inv);
            w.w = pack_bf2(__uint_as_float(v[i + 6]) * inv, __uint_as_float(v[i + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + c + i) = w;
          }
        }
      }
    }
  }
  tc_before();
  __syncthreads();
  if (warp == 4) {
    tc_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
  }
}

template <int HD> constexpr int tc_smem() { return 128 * HD * 2 + 4 * BNK * HD * 2 + 128 * BNK * 2 + 256; }

}  // namespace atc

int make_tmap(CUtensorMap* m, const void* ptr, int64_t rows, int64_t cols, int box_rows);   // gemm_tc.cu (cached)

template <int HD>
static int launch_tc(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt,
                     int max_blocks, const int32_t* sl, const int32_t* qsl, int B, int max_q_len, int H, int KV, int bs,
                     float scale, cudaStream_t st) {
  int bs_shift = 0;
  while ((1 << bs_shift) < bs) ++bs_shift;
  const int G = H / KV;
  const int TT = atc::BM / G;
  CUtensorMap tk, tv;
  const bool tma = tunable(TUN_ATTN_TMA) != 0 && bs >= 8;
  if (tma) {
    const int box_rows = bs < atc::BNK ? bs : atc::BNK;
    const int64_t rows = (int64_t)1 << 30;      // the pool's extent is not known here; it only bounds out-of-range handling, never relied on
    if (int rc = make_tmap(&tk, kc, rows, HD, box_rows)) return rc;
    if (int rc = make_tmap(&tv, vc, rows, HD, box_rows)) return rc;
    auto kern = atc::attn_tc_kernel<HD, true>;
    KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::tc_smem<HD>()));
    KGE_CUDA_OK(launch_pdl(kern, dim3((max_q_len + TT - 1) / TT, KV, B), dim3(192), atc::tc_smem<HD>(), st, tk, tv, out, q, q_stride,
                           kc, vc, bt, max_blocks, sl, qsl, H, KV, bs_shift, scale * atc::kLog2e));
    return 0;
  }
  memset(&tk, 0, sizeof tk); memset(&tv, 0, sizeof tv);
  auto kern = atc::attn_tc_kernel<HD, false>;
  KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::tc_smem<HD>()));
  KGE_CUDA_OK(launch_pdl(kern, dim3((max_q_len + TT - 1) / TT, KV, B), dim3(160), atc::tc_smem<HD>(), st, tk, tv, out, q, q_stride, kc,
                         vc, bt, max_blocks, sl, qsl, H, KV, bs_shift, scale * atc::kLog2e));
  return 0;
}

// Called by kge_paged_attn (attn.cu) for batches with query chunks longer than one token.
int paged_attn_prefill_tc(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt,
                          int max_blocks, const int32_t* sl, const int32_t* qsl, int B, int max_q_len, int H, int KV,
                          int hd, int bs, float scale, cudaStream_t st) {
  KGE_REQUIRE(H / KV <= 128, "kge_paged_attn: GQA group %d > 128 unsupported", H / KV);
  if (hd == 128) return launch_tc<128>(out, q, q_stride, kc, vc, bt, max_blocks, sl, qsl, B, max_q_len, H, KV, bs, scale, st);
  return launch_tc<64>(out, q, q_stride, kc, vc, bt, max_blocks, sl, qsl, B, max_q_len, H, KV, bs, scale, st);
}

}  // namespace KGE_ELT_NS
}  // namespace kge
