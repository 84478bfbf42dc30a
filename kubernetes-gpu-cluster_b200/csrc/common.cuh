// Shared device/host helpers for the kge CUDA sources (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

namespace kge {

// Thread-local error text behind kge_last_error() (api.cc).
void set_error(const char* fmt, ...);
// Process-wide switch for programmatic dependent launch (api.cc).  On by default; turned off by
// KGE_DISABLE_PDL and by tensor-parallel engines (see DESIGN.md "known issues").
bool pdl_enabled();
void pdl_set_enabled(bool on);
// pdl_break_next(): the next launch from this thread is an ordinary stream-ordered launch (it neither starts early nor
// relies on griddepcontrol.wait).  Used right after a kernel that is not ours (NCCL).  pdl_for_this_launch() consumes it.
void pdl_break_next();
bool pdl_for_this_launch();
// Launch-shape knobs behind kge_set_tunable() (api.cc): heuristics that only change HOW work is
// partitioned, never results.  Defaults are the measured-best settings (profiles/).
enum Tunable { TUN_GEMM_EVEN_GRID = 0, TUN_ATTN_DECODE_SPLITS = 1, TUN_ATTN_SPLIT_BALANCE = 2, TUN_ATTN_CLUSTER_MERGE = 3, TUN_ATTN_TMA = 4, TUN_GEMM_WS = 5, TUN_SMEM_CARVEOUT = 6, TUN_GEMM_L2_HINTS = 7, TUN_ATTN_PERSISTENT = 8, TUN_GEMM_WIDE = 9, TUN_COUNT = 10 };
int tunable(Tunable t);

#define KGE_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      kge::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return -2;                                                                       \
    }                                                                                  \
  } while (0)

#define KGE_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      kge::set_error(__VA_ARGS__);        \
      return -1;                          \
    }                                     \
  } while (0)

#define KGE_LAUNCH_OK() KGE_CUDA_OK(cudaGetLastError())

}  // namespace kge

// ---------------------------------------------------------------------------------------------------------------------
// Element type.  The kernel sources are compiled TWICE: as they are (bfloat16) and with -DKGE_HALF (float16, what
// `--dtype float16` of ref:values-01-minimal-example8.yaml:30-31 asks for).  Everything that depends on the 16-bit format
// goes through the handful of helpers below; the two builds live in two inline namespaces (so their kernels and host
// helpers are distinct symbols) and export their C entry points under KGE_API(name): `name` and `name_f16`.
// In the float16 build `bf16` simply names the 16-bit storage element.
#ifdef KGE_HALF
#include <cuda_fp16.h>
#define KGE_ELT_NS h16
#define KGE_API(name) name##_f16
#define KGE_MMA_ELT "f16"
#define KGE_IDESC_AB 0u                              /* tcgen05 instruction descriptor: A and B are F16 */
#else
#define KGE_ELT_NS b16
#define KGE_API(name) name
#define KGE_MMA_ELT "bf16"
#define KGE_IDESC_AB ((1u << 7) | (1u << 10))        /* tcgen05 instruction descriptor: A and B are BF16 */
#endif

namespace kge {
inline namespace KGE_ELT_NS {

#ifdef KGE_HALF
typedef __half bf16;
#else
typedef __nv_bfloat16 bf16;
#endif

#ifdef __CUDACC__
#ifdef KGE_HALF
__device__ __forceinline__ float bf2f(bf16 v) { return __half2float(v); }
__device__ __forceinline__ bf16 f2bf(float v) { return __float2half_rn(v); }
__device__ __forceinline__ float rbf(float v) { return __half2float(__float2half_rn(v)); }
__device__ __forceinline__ float bf_lo(uint32_t u) { return __half2float(__ushort_as_half((unsigned short)(u & 0xffffu))); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __half2float(__ushort_as_half((unsigned short)(u >> 16))); }
__device__ __forceinline__ uint32_t pack_bf2(float lo, float hi) {
  __half2 p = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}
#else
__device__ __forceinline__ float bf2f(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ bf16 f2bf(float v) { return __float2bfloat16_rn(v); }
// round-trip through the element type: the "every op rounds in the model dtype" steps of the reference
__device__ __forceinline__ float rbf(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}
#endif
#endif

}  // namespace KGE_ELT_NS
}  // namespace kge

namespace kge {
#ifdef __CUDACC__
// Programmatic dependent launch: every kernel of the forward pass lets its successor start early
// (launch latency, prologue and - for the GEMM - the first weight tiles overlap this kernel's
// execution) and waits for its predecessor's results right before it first touches them.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Every kernel of a step asks for the same L1 / shared-memory split (all shared memory): the GEMM and attention kernels
// need ~200 KB per SM anyway, and an SM that has to change its split between two kernels must drain first, which
// serialises exactly the kernel boundaries that programmatic dependent launch is there to overlap.  (Knob
// "smem_carveout" 0 leaves the small kernels at the driver's default split.)
template <typename K>
inline void prefer_max_shared(K kern) {
  if (tunable(TUN_SMEM_CARVEOUT)) (void)cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  prefer_max_shared(kern);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_for_this_launch() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024; `red` is >= 33 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : 0.f;
  t = warp_sum(t);
  return t;
}
#endif

}  // namespace kge
