// Tensor-parallel exchange over peer memory (NVLink / NVSwitch): the row-parallel projections' all-reduce fused with
// the residual add and the norm that always follows it, and the all-gather of vocab-sharded logits.
//
// Replaces, per o_proj / down_proj (fc2 / out_proj): ncclAllReduce + fused_add_rms_norm
//   vllm:model_executor/layers/linear.py:1559 (all-reduce of the bf16 partials) followed by
//   vllm:ir/ops/layernorm.py:40-58 (llama.py:323-331,430) or nn.LayerNorm after the residual add (opt.py:181-196);
// and vllm:model_executor/layers/logits_processor.py:83 (all-gather of the vocab shards).
//
// Every rank owns one WINDOW of device memory (cudaMalloc'ed, exported with cudaIpcGetMemHandle or, for ranks that are
// threads of one process, used directly after cudaDeviceEnablePeerAccess).  All ranks map all windows.  A window holds a
// small control block (epoch, arrival counter, per-peer flags) and the caller's exchange buffers at offsets that are
// the same on every rank.
//
// reduce + add + norm, two-shot and partitioned by TOKEN rows (rank r owns rows [r*ceil(T/N), ...)):
//   1. START barrier: "my partial [T,d] is complete in my window" (st.release.sys of the epoch into every peer's flag
//      slot, each CTA spins on its own window's slots);
//   2. each owned row: 16-byte loads of that row from every rank's partial (ld.relaxed.sys - never a stale L1 line),
//      fp32 sum in rank order, rounded once to bf16 (the all-reduce's bf16 result), residual add and norm with the
//      arithmetic of norm_kernel (elementwise.cu), residual kept for OWNED rows only (nobody else reads it), the normed
//      row stored into EVERY rank's output buffer (plain 16-byte stores to peer memory);
//   3. DONE barrier: the last CTA to finish (device-scope arrival counter after fence.sys) releases the epoch to every
//      peer and waits for theirs - when the kernel ends, every rank's rows have landed here and nobody still reads this
//      rank's partial.
// So a rank reads (N-1)/N of T*d*2 bytes and writes the same amount, each row is normalised once in the whole group
// (not once per rank), and one launch replaces an NCCL kernel plus a norm kernel.  The epoch lives in the window, not in
// the launch arguments, so captured CUDA graphs replay.  Spins are bounded: a peer that never arrives traps the kernel
// (an error on the stream) instead of hanging the GPU.
//
// No TMA and no multicast anywhere in this file: peer memory is only touched by plain and .sys-scoped loads / stores.
// (TMA multicast loads in this library read only this GPU's memory, since on B200 a TMA multicast load from another
// GPU's memory hung the GPU.)  The GEMMs that consume the output read it through tensor maps over the LOCAL window.
#include <string.h>
#include <unistd.h>

#include "common.cuh"
#include "kge.h"

namespace kge {
inline namespace KGE_ELT_NS {

constexpr int kTpMaxRanks = 16;
constexpr int64_t kTpCtrlBytes = 4096;            // control block at the start of every window
constexpr int kFlagStride = 32;                   // one 32-byte sector per flag
constexpr long long kSpinLimitCycles = 40000000000ll;   // ~20 s at 2 GHz: a missing peer becomes a trap, not a hang

struct TpParams {
  uint8_t* win[kTpMaxRanks];
  int rank, world;
};

__device__ __forceinline__ uint32_t* tp_flag(uint8_t* win, int phase, int src) {
  return reinterpret_cast<uint32_t*>(win + 256 + (phase * kTpMaxRanks + src) * kFlagStride);
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void tp_spin(const uint32_t* f, uint32_t epoch) {
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
    if (clock64() - t0 > kSpinLimitCycles) __trap();
    __nanosleep(20);
  }
}

// Phase 0 of every collective here: tell every peer that this rank's input is complete, wait until theirs is.
__device__ __forceinline__ uint32_t tp_start_barrier(const TpParams& p) {
  uint8_t* me = p.win[p.rank];
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(me) + 1;
  const int tid = threadIdx.x;
  if (blockIdx.x == 0 && tid < p.world) st_release_sys(tp_flag(p.win[tid], 0, p.rank), epoch);
  if (tid < p.world) tp_spin(tp_flag(me, 0, tid), epoch);
  __syncthreads();
  return epoch;
}
// Last phase: when every CTA of this rank has stored its share, tell every peer and wait for all of them.
__device__ __forceinline__ void tp_done_barrier(const TpParams& p, uint32_t epoch, int* s_last) {
  uint8_t* me = p.win[p.rank];
  uint32_t* done = reinterpret_cast<uint32_t*>(me + 4);
  const int tid = threadIdx.x;
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();                                   // this CTA's peer stores are performed before it arrives
    *s_last = (atomicAdd(done, 1u) == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (*s_last) {
    if (tid == 0) {
      *done = 0;
      *reinterpret_cast<volatile uint32_t*>(me) = epoch;      // the next collective's epoch starts from here
      __threadfence_system();
    }
    __syncthreads();
    if (tid < p.world) {
      st_release_sys(tp_flag(p.win[tid], 1, p.rank), epoch);
      tp_spin(tp_flag(me, 1, tid), epoch);
    }
  }
}

constexpr int kTpMaxVec = 4;      // 16-byte vectors per thread per row: d <= 32 * blockDim

// MODE 1: h = bf16(sum of partials); s = h + res (fp32); res <- bf16(s); y = rmsnorm(s) * w
// MODE 3: h = bf16(sum of partials (+ bias, rank 0's partial carries it)); o = bf16(res + h); res <- o; y = layernorm(o)
// MODE 0: y = bf16(sum of partials)   (plain all-reduce, every rank gets every row)
template <int MODE>
__global__ void __launch_bounds__(512)
tp_reduce_norm_kernel(const TpParams p, int64_t x_off, int64_t y_off, bf16* __restrict__ residual,
                      const bf16* __restrict__ w, const bf16* __restrict__ b, int T, int d, float eps) {
  __shared__ float red[33];
  __shared__ int s_last;
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = tp_start_barrier(p);
  const int rows_per = (T + p.world - 1) / p.world;
  const int r0 = p.rank * rows_per, r1 = min(T, r0 + rows_per);
  const int nvec = d / 8;
  for (int row = r0 + (int)blockIdx.x; row < r1; row += (int)gridDim.x) {
    const int64_t base = (int64_t)row * d;
    float v[kTpMaxVec][8];
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < kTpMaxVec; ++j) {
      const int i = threadIdx.x + j * blockDim.x;
      if (i < nvec) {
        float* f = v[j];
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = 0.f;
        // all ranks' loads of this vector are issued before the first one is consumed (one NVLink round trip, not
        // one per rank); the sum runs in rank order, so it is the same wherever it is computed
        for (int q0 = 0; q0 < p.world; q0 += 8) {
          uint4 a[8];
#pragma unroll
          for (int u = 0; u < 8; ++u)
            if (q0 + u < p.world) a[u] = ld_relaxed_sys_v4(p.win[q0 + u] + x_off + (base + i * 8) * 2);
#pragma unroll
          for (int u = 0; u < 8; ++u)
            if (q0 + u < p.world) {
              f[0] += bf_lo(a[u].x); f[1] += bf_hi(a[u].x); f[2] += bf_lo(a[u].y); f[3] += bf_hi(a[u].y);
              f[4] += bf_lo(a[u].z); f[5] += bf_hi(a[u].z); f[6] += bf_lo(a[u].w); f[7] += bf_hi(a[u].w);
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = rbf(f[k]);         // the all-reduce's result is a bf16 tensor
        if (MODE == 1 || MODE == 3) {
          const uint4 r = *reinterpret_cast<const uint4*>(residual + base + i * 8);
          f[0] += bf_lo(r.x); f[1] += bf_hi(r.x); f[2] += bf_lo(r.y); f[3] += bf_hi(r.y);
          f[4] += bf_lo(r.z); f[5] += bf_hi(r.z); f[6] += bf_lo(r.w); f[7] += bf_hi(r.w);
          uint4 o;
          o.x = pack_bf2(f[0], f[1]); o.y = pack_bf2(f[2], f[3]);
          o.z = pack_bf2(f[4], f[5]); o.w = pack_bf2(f[6], f[7]);
          *reinterpret_cast<uint4*>(residual + base + i * 8) = o;
          if (MODE == 3) {                                    // OPT normalises the ROUNDED sum (opt.py:181-186)
            f[0] = bf_lo(o.x); f[1] = bf_hi(o.x); f[2] = bf_lo(o.y); f[3] = bf_hi(o.y);
            f[4] = bf_lo(o.z); f[5] = bf_hi(o.z); f[6] = bf_lo(o.w); f[7] = bf_hi(o.w);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) acc += (MODE == 1) ? f[k] * f[k] : f[k];
      }
    }
    float mean = 0.f, rstd = 1.f;
    if (MODE == 1) {
      rstd = 1.0f / sqrtf(block_sum(acc, red) / (float)d + eps);
    } else if (MODE == 3) {
      mean = block_sum(acc, red) / (float)d;
      float a2 = 0.f;
#pragma unroll
      for (int j = 0; j < kTpMaxVec; ++j) {
        const int i = threadIdx.x + j * blockDim.x;
        if (i < nvec) {
#pragma unroll
          for (int k = 0; k < 8; ++k) { const float c = v[j][k] - mean; a2 += c * c; }
        }
      }
      rstd = 1.0f / sqrtf(block_sum(a2, red) / (float)d + eps);
    }
#pragma unroll
    for (int j = 0; j < kTpMaxVec; ++j) {
      const int i = threadIdx.x + j * blockDim.x;
      if (i < nvec) {
        float o[8];
        if (MODE == 0) {
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] = v[j][k];
        } else {
          const uint4 ww = *reinterpret_cast<const uint4*>(w + i * 8);
          const float wf[8] = {bf_lo(ww.x), bf_hi(ww.x), bf_lo(ww.y), bf_hi(ww.y), bf_lo(ww.z), bf_hi(ww.z), bf_lo(ww.w), bf_hi(ww.w)};
          if (MODE == 1) {
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = rbf(v[j][k] * rstd) * wf[k];     // weight multiply in bf16
          } else {
            const uint4 bb = *reinterpret_cast<const uint4*>(b + i * 8);
            const float bfv[8] = {bf_lo(bb.x), bf_hi(bb.x), bf_lo(bb.y), bf_hi(bb.y), bf_lo(bb.z), bf_hi(bb.z), bf_lo(bb.w), bf_hi(bb.w)};
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = (v[j][k] - mean) * rstd * wf[k] + bfv[k];
          }
        }
        uint4 ov;
        ov.x = pack_bf2(o[0], o[1]); ov.y = pack_bf2(o[2], o[3]);
        ov.z = pack_bf2(o[4], o[5]); ov.w = pack_bf2(o[6], o[7]);
        for (int q = 0; q < p.world; ++q)                     // the finished row goes to every rank (own window included)
          *reinterpret_cast<uint4*>(p.win[q] + y_off + (base + i * 8) * 2) = ov;
      }
    }
  }
  tp_done_barrier(p, epoch, &s_last);
}

// dst[r, col0 + c] of EVERY rank's window <- src[r, c] (this rank's column shard), r < R, c < cols.
template <bool VEC>
__global__ void __launch_bounds__(256)
tp_all_gather_cols_kernel(const TpParams p, int64_t dst_off, const bf16* __restrict__ src, int R, int cols, int ld_dst, int col0) {
  __shared__ int s_last;
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t epoch = tp_start_barrier(p);       // nobody is still reading last step's gathered rows
  if (VEC) {
    const int cv = cols / 8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (int64_t)R * cv; i += (int64_t)gridDim.x * blockDim.x) {
      const int r = (int)(i / cv), c = (int)(i % cv) * 8;
      const uint4 a = *reinterpret_cast<const uint4*>(src + (int64_t)r * cols + c);
      for (int q = 0; q < p.world; ++q)
        *reinterpret_cast<uint4*>(p.win[q] + dst_off + ((int64_t)r * ld_dst + col0 + c) * 2) = a;
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (int64_t)R * cols; i += (int64_t)gridDim.x * blockDim.x) {
      const int r = (int)(i / cols), c = (int)(i % cols);
      const bf16 a = src[(int64_t)r * cols + c];
      for (int q = 0; q < p.world; ++q)
        *reinterpret_cast<bf16*>(p.win[q] + dst_off + ((int64_t)r * ld_dst + col0 + c) * 2) = a;
    }
  }
  tp_done_barrier(p, epoch, &s_last);
}

static inline int tp_row_threads(int d) {
  const int v = d / 8;
  int t = 32;
  while (t < v && t < 512) t <<= 1;
  return t;
}

}  // namespace KGE_ELT_NS
}  // namespace kge

using namespace kge;

struct kge_tp_comm {
  int rank = 0, world = 1, device = 0;
  int64_t bytes = 0;                     // whole window, control block included
  uint8_t* window = nullptr;             // own window (cudaMalloc)
  uint8_t* peer[kTpMaxRanks] = {};       // every rank's window as mapped here (peer[rank] == window)
  bool opened[kTpMaxRanks] = {};         // mapped with cudaIpcOpenMemHandle (to be closed)
  bool connected = false;
  int sm_count = 148;
};

#ifndef KGE_HALF      // window management and the byte-wise gather do not depend on the element type
extern "C" int kge_tp_comm_create(kge_tp_comm** out, int rank, int world, int device, int64_t data_bytes) {
  KGE_REQUIRE(out, "kge_tp_comm_create: null argument");
  *out = nullptr;
  KGE_REQUIRE(world >= 1 && world <= kTpMaxRanks && rank >= 0 && rank < world, "kge_tp_comm_create: rank %d of %d (at most %d ranks)", rank, world, kTpMaxRanks);
  KGE_REQUIRE(data_bytes >= 0, "kge_tp_comm_create: negative size");
  KGE_CUDA_OK(cudaSetDevice(device));
  kge_tp_comm* c = new kge_tp_comm();
  c->rank = rank; c->world = world; c->device = device;
  c->bytes = kTpCtrlBytes + ((data_bytes + 255) / 256) * 256;
  if (cudaMalloc((void**)&c->window, (size_t)c->bytes) != cudaSuccess) {
    kge::set_error("kge_tp_comm_create: cudaMalloc of a %lld-byte window failed: %s", (long long)c->bytes, cudaGetErrorString(cudaGetLastError()));
    delete c;
    return -2;
  }
  cudaMemset(c->window, 0, (size_t)kTpCtrlBytes);      // epoch 0, no flags: must be in place before any peer maps it
  cudaDeviceSynchronize();
  c->peer[rank] = c->window;
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (world == 1) c->connected = true;
  *out = c;
  return 0;
}

extern "C" void kge_tp_comm_destroy(kge_tp_comm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < c->world; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->peer[r]);
  if (c->window) cudaFree(c->window);
  delete c;
}

extern "C" int kge_tp_comm_info(kge_tp_comm* c, kge_tp_peer_info* out) {
  KGE_REQUIRE(c && out, "kge_tp_comm_info: null argument");
  memset(out, 0, sizeof(*out));
  KGE_CUDA_OK(cudaSetDevice(c->device));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  cudaIpcMemHandle_t h;
  KGE_CUDA_OK(cudaIpcGetMemHandle(&h, c->window));
  memcpy(out->ipc_handle, &h, 64);
  out->ptr = (uint64_t)(uintptr_t)c->window;
  out->pid = (int64_t)getpid();
  out->bytes = c->bytes;
  out->device = c->device;
  out->rank = c->rank;
  return 0;
}

extern "C" int kge_tp_comm_connect(kge_tp_comm* c, const kge_tp_peer_info* all) {
  KGE_REQUIRE(c && all, "kge_tp_comm_connect: null argument");
  KGE_CUDA_OK(cudaSetDevice(c->device));
  for (int r = 0; r < c->world; ++r) {
    KGE_REQUIRE(all[r].rank == r, "kge_tp_comm_connect: entry %d describes rank %d", r, all[r].rank);
    KGE_REQUIRE(all[r].bytes == c->bytes, "kge_tp_comm_connect: rank %d has a %lld-byte window, this rank %lld (different engine configurations?)",
                r, (long long)all[r].bytes, (long long)c->bytes);
    if (r == c->rank) continue;
    if (all[r].pid == (int64_t)getpid()) {
      // ranks as threads of one process (the server): the peer's pointer is valid here once peer access is on
      int can = 0;
      KGE_CUDA_OK(cudaDeviceCanAccessPeer(&can, c->device, all[r].device));
      KGE_REQUIRE(can, "kge_tp_comm_connect: device %d cannot access device %d's memory", c->device, all[r].device);
      cudaError_t e = cudaDeviceEnablePeerAccess(all[r].device, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
      else KGE_CUDA_OK(e);
      c->peer[r] = (uint8_t*)(uintptr_t)all[r].ptr;
    } else {
      cudaIpcMemHandle_t h;
      memcpy(&h, all[r].ipc_handle, 64);
      void* ptr = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        kge::set_error("kge_tp_comm_connect: cudaIpcOpenMemHandle for rank %d failed: %s", r, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return -2;
      }
      c->peer[r] = (uint8_t*)ptr;
      c->opened[r] = true;
    }
  }
  c->connected = true;
  return 0;
}

extern "C" void* kge_tp_comm_data(kge_tp_comm* c) { return c ? c->window + kTpCtrlBytes : nullptr; }
extern "C" int kge_tp_comm_connected(const kge_tp_comm* c) { return c && c->connected ? 1 : 0; }

#endif

static int tp_params(const kge_tp_comm* c, TpParams* p, const char* who) {
  KGE_REQUIRE(c, "%s: null communicator", who);
  KGE_REQUIRE(c->connected, "%s: kge_tp_comm_connect has not been called", who);
  for (int r = 0; r < kTpMaxRanks; ++r) p->win[r] = r < c->world ? c->peer[r] : nullptr;
  p->rank = c->rank; p->world = c->world;
  return 0;
}

extern "C" int KGE_API(kge_tp_reduce_add_norm)(kge_tp_comm* c, int64_t y_off, void* residual, int64_t x_off, const void* w,
                                      const void* b, int T, int d, float eps, int mode, void* stream) {
  TpParams p;
  if (int rc = tp_params(c, &p, "kge_tp_reduce_add_norm")) return rc;
  KGE_REQUIRE(mode == 0 || mode == 1 || mode == 3, "kge_tp_reduce_add_norm: mode %d (0 sum, 1 add + RMSNorm, 3 add + LayerNorm)", mode);
  KGE_REQUIRE(T >= 0 && d > 0 && d % 8 == 0 && d <= 8 * kTpMaxVec * 512, "kge_tp_reduce_add_norm: d=%d must be a multiple of 8, at most %d", d, 8 * kTpMaxVec * 512);
  KGE_REQUIRE(x_off >= 0 && y_off >= 0 && x_off % 16 == 0 && y_off % 16 == 0, "kge_tp_reduce_add_norm: offsets must be 16-byte aligned");
  const int64_t need = (int64_t)T * d * 2, cap = c->bytes - kTpCtrlBytes;
  KGE_REQUIRE(x_off + need <= cap && y_off + need <= cap, "kge_tp_reduce_add_norm: [T=%d, d=%d] does not fit the %lld-byte window", T, d, (long long)cap);
  KGE_REQUIRE(mode == 0 || (residual && w), "kge_tp_reduce_add_norm: residual and weight are required");
  KGE_REQUIRE(mode != 3 || b, "kge_tp_reduce_add_norm: LayerNorm needs a bias");
  if (T == 0) return 0;
  const int rows_per = (T + c->world - 1) / c->world;
  int threads = tp_row_threads(d);
  while ((d / 8 + threads - 1) / threads > kTpMaxVec) threads <<= 1;
  // one CTA per owned row up to two waves of CTAs; at least one CTA even when this rank owns no row (barriers)
  int grid = rows_per < 1 ? 1 : rows_per;
  const int cap_grid = c->sm_count * (threads <= 256 ? 4 : 2);
  if (grid > cap_grid) grid = cap_grid;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t xo = kTpCtrlBytes + x_off, yo = kTpCtrlBytes + y_off;
  if (mode == 1)
    KGE_CUDA_OK(launch_pdl(tp_reduce_norm_kernel<1>, dim3(grid), dim3(threads), 0, st, p, xo, yo, (bf16*)residual, (const bf16*)w, (const bf16*)b, T, d, eps));
  else if (mode == 3)
    KGE_CUDA_OK(launch_pdl(tp_reduce_norm_kernel<3>, dim3(grid), dim3(threads), 0, st, p, xo, yo, (bf16*)residual, (const bf16*)w, (const bf16*)b, T, d, eps));
  else
    KGE_CUDA_OK(launch_pdl(tp_reduce_norm_kernel<0>, dim3(grid), dim3(threads), 0, st, p, xo, yo, (bf16*)residual, (const bf16*)w, (const bf16*)b, T, d, eps));
  return 0;
}

#ifndef KGE_HALF
extern "C" int kge_tp_all_gather_cols(kge_tp_comm* c, int64_t dst_off, const void* src, int R, int cols, int ld_dst, int col0,
                                      void* stream) {
  TpParams p;
  if (int rc = tp_params(c, &p, "kge_tp_all_gather_cols")) return rc;
  KGE_REQUIRE(src && R >= 0 && cols >= 0 && ld_dst > 0 && col0 >= 0 && col0 + cols <= ld_dst, "kge_tp_all_gather_cols: bad shape (cols %d at %d of %d)", cols, col0, ld_dst);
  KGE_REQUIRE(dst_off >= 0 && dst_off % 16 == 0 && dst_off + (int64_t)R * ld_dst * 2 <= c->bytes - kTpCtrlBytes,
              "kge_tp_all_gather_cols: [R=%d, %d] does not fit the window", R, ld_dst);
  if (R == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = cols % 8 == 0 && col0 % 8 == 0 && ld_dst % 8 == 0 && ((uintptr_t)src & 15) == 0;
  const int64_t work = vec ? (int64_t)R * (cols / 8) : (int64_t)R * cols;
  int grid = (int)((work + 255) / 256);
  if (grid < 1) grid = 1;
  if (grid > c->sm_count * 4) grid = c->sm_count * 4;
  const int64_t off = kTpCtrlBytes + dst_off;
  if (vec) KGE_CUDA_OK(launch_pdl(tp_all_gather_cols_kernel<true>, dim3(grid), dim3(256), 0, st, p, off, (const bf16*)src, R, cols, ld_dst, col0));
  else KGE_CUDA_OK(launch_pdl(tp_all_gather_cols_kernel<false>, dim3(grid), dim3(256), 0, st, p, off, (const bf16*)src, R, cols, ld_dst, col0));
  return 0;
}
#endif
