// Prefill / mixed-batch paged attention on the 5th-gen tensor cores.
//
// One CTA = 128 query rows ((token, q-head-in-group) pairs of one kv head of one request) x the
// causal prefix of that request, walked in 128-key tiles:
//   S  = Q K^T      tcgen05.mma, A = Q tile, B = K tile (both K-major, 128B swizzle), D in TMEM [0,128)
//   P  = softmax    4 warps, ONE THREAD PER ROW: tcgen05.ld its S row, scale + causal mask + running
//                   max/sum in fp32 (no shuffles), bf16 P written straight into the swizzled A-operand
//                   layout in shared memory
//   O_t = P V       tcgen05.mma, A = P, B = V tile as stored ([key][hd] = MN-major), D in TMEM [128,128+hd)
//   O  = O*alpha + O_t in registers (exact online softmax, no lazy rescale)
// K/V tiles are gathered from the paged cache with cp.async (double-buffered, 16-byte chunks placed at
// their 128B-swizzled positions, block-table lookup once per row), made visible to the tensor-core
// (async) proxy with fence.proxy.async, and handed over with mbarriers; tcgen05.commit signals the
// softmax warps.  Tensor-bound: 4 * hd * sum_b q_b (ctx0_b + (q_b+1)/2) flop per q head.
#include "common.cuh"
#include "kge.h"

namespace kge {

namespace atc {

constexpr int BM = 128, BNK = 128;             // query rows, keys per tile
constexpr uint32_t kSpin = 1u << 30;
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
__device__ __forceinline__ uint32_t tile_off(int row, int chunk /* 16-byte chunk of the full row */) {
  return (uint32_t)((chunk >> 3) * (128 * 128) + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}
// K-major operand (A, or B = K tile): rows of 128 B, 8-row atoms 1024 B apart
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
// MN-major B operand (V tile [key][hd]): 64-wide hd halves 16 KB apart (LBO), 8-key atoms 1024 B apart (SBO)
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((128 * 128) >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}

template <int HD>
__global__ void __launch_bounds__(160, 1)
attn_tc_kernel(bf16* __restrict__ out, const bf16* __restrict__ q, int q_stride, const bf16* __restrict__ k_cache,
               const bf16* __restrict__ v_cache, const int32_t* __restrict__ block_table, int max_blocks,
               const int32_t* __restrict__ seq_lens, const int32_t* __restrict__ qsl, int H, int KV, int bs_shift,
               float scale_log2) {
  constexpr int CH = HD / 8;                       // 16-byte chunks per row
  constexpr uint32_t TILE_BYTES = 128 * HD * 2;    // Q, K or V tile
  constexpr uint32_t P_BYTES = 128 * BNK * 2;
  constexpr uint32_t TMEM_COLS = 256;              // S [0,128) + O_t [128,128+HD)
  // S: M=128, N=128, both K-major.  PV: M=128, N=HD, A K-major, B MN-major (bit 16).
  constexpr uint32_t IDESC_S = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BNK >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
  constexpr uint32_t IDESC_O = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(HD >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

  const int b = blockIdx.z, kvh = blockIdx.y;
  const int G = H / KV;
  const int TT = BM / G;
  const int q0 = qsl[b], q_len = qsl[b + 1] - q0;
  const int tok0 = blockIdx.x * TT;
  if (tok0 >= q_len) return;
  const int seq_len = seq_lens[b];
  const int ctx0 = seq_len - q_len;
  const int ntok = min(TT, q_len - tok0);
  const int kv_len = ctx0 + tok0 + ntok;
  const int ntiles = (kv_len + BNK - 1) / BNK;
  const int32_t* btab = block_table + (int64_t)b * max_blocks;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sQ = s_addr(smem);
  const uint32_t sK = sQ + TILE_BYTES;             // 2 stages
  const uint32_t sV = sK + 2 * TILE_BYTES;         // 2 stages
  const uint32_t sP = sV + 2 * TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 5 * TILE_BYTES + P_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  const uint32_t bar0 = s_addr(bars);
  const uint32_t kv_full0 = bar0, kv_full1 = bar0 + 8, s_full = bar0 + 16, p_full = bar0 + 24, o_full = bar0 + 32;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (warp == 4) {
    if (lane == 0) {
      mbar_init(kv_full0, 128); mbar_init(kv_full1, 128);
      mbar_init(s_full, 1); mbar_init(p_full, 128); mbar_init(o_full, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_addr(tmem_slot)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 4) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      for (int j = 0; j < ntiles; ++j) {
        const int st = j & 1;
        mbar_wait(st ? kv_full1 : kv_full0, (uint32_t)((j >> 1) & 1));
        tc_after();
        const uint32_t kt = sK + st * TILE_BYTES, vt = sV + st * TILE_BYTES;
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) {
          const uint32_t off = (uint32_t)((k >> 2) * (128 * 128) + (k & 3) * 32);
          tc_mma(tmem, desc_kmajor(sQ + off), desc_kmajor(kt + off), IDESC_S, k > 0 ? 1u : 0u);
        }
        tc_commit(s_full);
        mbar_wait(p_full, (uint32_t)(j & 1));
        tc_after();
#pragma unroll
        for (int k = 0; k < BNK / 16; ++k) {
          const uint32_t aoff = (uint32_t)((k >> 2) * (128 * 128) + (k & 3) * 32);   // P: keys are the K dimension
          const uint32_t boff = (uint32_t)(k * 16 * 128);                             // V: 16 keys = 16 rows of 128 B
          tc_mma(tmem + 128, desc_kmajor(sP + aoff), desc_mnmajor(vt + boff), IDESC_O, k > 0 ? 1u : 0u);
        }
        tc_commit(o_full);
      }
    }
  } else {
    // ------------------------------------------------------------------ loaders + softmax (one row per thread)
    const int r = tid;
    const int tk = r / G, hh = r % G;
    const bool row_ok = tk < ntok;
    const int lim = row_ok ? ctx0 + tok0 + tk : -1;          // last visible key of this row
    // loader geometry: thread t owns chunk column t % CH of rows t / CH + i * (128 / CH)
    constexpr int RP = 128 / CH, NP = 128 / RP;
    const int lrow0 = tid / CH, lc = tid % CH;
    const int64_t head_off = ((int64_t)kvh << bs_shift) * HD;
    const int64_t blk_stride = ((int64_t)KV << bs_shift) * HD;
    const int bs_mask = (1 << bs_shift) - 1;
    auto load_kv = [&](int j) {
      const uint32_t kt = sK + (j & 1) * TILE_BYTES, vt = sV + (j & 1) * TILE_BYTES;
#pragma unroll
      for (int i = 0; i < NP; ++i) {
        const int row = lrow0 + i * RP;
        const int key = j * BNK + row;
        const bool ok = key < kv_len;
        int64_t off = 0;
        if (ok) off = (int64_t)btab[key >> bs_shift] * blk_stride + head_off + (key & bs_mask) * HD + lc * 8;
        const uint32_t d = tile_off(row, lc);
        cp16(kt + d, k_cache + off, ok);
        cp16(vt + d, v_cache + off, ok);
      }
    };
    // Q tile
#pragma unroll
    for (int i = 0; i < NP; ++i) {
      const int row = lrow0 + i * RP;
      const int t2 = row / G, h2 = row % G;
      const bool ok = t2 < ntok;
      const bf16* src = ok ? q + (int64_t)(q0 + tok0 + t2) * q_stride + (kvh * G + h2) * HD + lc * 8 : q;
      cp16(sQ + tile_off(row, lc), src, ok);
    }
    load_kv(0);
    cp_commit();

    float o[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);

    for (int j = 0; j < ntiles; ++j) {
      if (j + 1 < ntiles) { load_kv(j + 1); cp_commit(); cp_wait<1>(); }
      else cp_wait<0>();
      fence_async_smem();                 // my cp.async bytes -> visible to the tensor-core proxy
      tc_before();                        // my TMEM reads of the previous tile are ordered before the next MMA
      mbar_arrive((j & 1) ? kv_full1 : kv_full0);

      mbar_wait(s_full, (uint32_t)(j & 1));
      tc_after();
      const int kbase = j * BNK;
      // pass 1: row maximum of the masked, scaled scores
      float mx = m;
#pragma unroll 1
      for (int c = 0; c < BNK; c += 32) {
        uint32_t v[32];
        tmem_ld32(t_row + c, v);
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (kbase + c + i <= lim) mx = fmaxf(mx, __uint_as_float(v[i]) * scale_log2);
      }
      const float base = (mx == -INFINITY) ? 0.f : mx;
      const float alpha = exp2f(m - base);
      m = mx;
      // pass 2: P = exp2(s - max) as bf16 into the A-operand tile, row sum
      float rs = 0.f;
#pragma unroll 1
      for (int c = 0; c < BNK; c += 32) {
        uint32_t v[32];
        tmem_ld32(t_row + c, v);
#pragma unroll
        for (int g8 = 0; g8 < 4; ++g8) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int c0 = c + g8 * 8 + i * 2;
            const float p0 = (kbase + c0 <= lim) ? exp2f(__uint_as_float(v[g8 * 8 + i * 2]) * scale_log2 - base) : 0.f;
            const float p1 = (kbase + c0 + 1 <= lim) ? exp2f(__uint_as_float(v[g8 * 8 + i * 2 + 1]) * scale_log2 - base) : 0.f;
            rs += p0 + p1;
            pk[i] = pack_bf2(p0, p1);
          }
          const uint32_t dst = sP + tile_off(r, (c >> 3) + g8);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
        }
      }
      l = l * alpha + rs;
      fence_async_smem();                 // P bytes -> tensor-core proxy
      tc_before();
      mbar_arrive(p_full);

      mbar_wait(o_full, (uint32_t)(j & 1));
      tc_after();
#pragma unroll
      for (int c = 0; c < HD; c += 32) {
        uint32_t v[32];
        tmem_ld32(t_row + 128 + c, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c + i] = o[c + i] * alpha + __uint_as_float(v[i]);
      }
    }
    if (row_ok) {
      const float inv = l > 0.f ? 1.f / l : 0.f;
      bf16* dst = out + (int64_t)(q0 + tok0 + tk) * H * HD + (kvh * G + hh) * HD;
#pragma unroll
      for (int c = 0; c < HD; c += 8) {
        uint4 w;
        w.x = pack_bf2(o[c] * inv, o[c + 1] * inv); w.y = pack_bf2(o[c + 2] * inv, o[c + 3] * inv);
        w.z = pack_bf2(o[c + 4] * inv, o[c + 5] * inv); w.w = pack_bf2(o[c + 6] * inv, o[c + 7] * inv);
        *reinterpret_cast<uint4*>(dst + c) = w;
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

template <int HD> constexpr int tc_smem() { return 5 * 128 * HD * 2 + 128 * BNK * 2 + 256 + 1024; }

}  // namespace atc

template <int HD>
static int launch_tc(bf16* out, const bf16* q, int q_stride, const bf16* kc, const bf16* vc, const int32_t* bt,
                     int max_blocks, const int32_t* sl, const int32_t* qsl, int B, int max_q_len, int H, int KV, int bs,
                     float scale, cudaStream_t st) {
  int bs_shift = 0;
  while ((1 << bs_shift) < bs) ++bs_shift;
  const int G = H / KV;
  const int TT = atc::BM / G;
  auto kern = atc::attn_tc_kernel<HD>;
  KGE_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::tc_smem<HD>()));
  kern<<<dim3((max_q_len + TT - 1) / TT, KV, B), 160, atc::tc_smem<HD>(), st>>>(out, q, q_stride, kc, vc, bt, max_blocks, sl,
                                                                              qsl, H, KV, bs_shift, scale * atc::kLog2e);
  KGE_LAUNCH_OK();
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

}  // namespace kge
