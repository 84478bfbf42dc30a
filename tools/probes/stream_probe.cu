// Micro-benchmark (measurement only, not part of the library): how fast can N CTAs stream weight tiles from HBM into shared
// memory through a ring of TMA stages, as a function of the tile's shape in memory?
//   mode 0  2-D tensor map over a row-major [rows, 4096] bf16 matrix, box = 64 columns x R rows: R separate 128-byte runs
//           8 KB apart (how the GEMM reads a K-major weight today)
//   mode 1  cp.async.bulk of one contiguous chunk of R x 128 bytes (a weight repacked tile by tile)
//   mode 2  2-D tensor map over a [rows, 64] matrix (row pitch 128 B), box = 64 x R: the same contiguous bytes through the
//           tiled path (128-byte swizzle)
// No math: a consumer thread waits for each stage and hands it back.  Build + run: tools/probes/run_stream_probe.sh
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t ph) {
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(b), "r"(ph) : "memory");
}

struct P { const uint8_t* src; long long base; int mode, rows, kblocks, stages, xrows; unsigned long long* out; };

__global__ void __launch_bounds__(64, 1) stream_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tx, const P p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int S = p.stages;
  const uint32_t stage_bytes = (uint32_t)p.rows * 128u;
  const uint32_t x_bytes = (uint32_t)p.xrows * 128u;      // the activation tile of the k-block (L2 resident), 0 = none
  uint64_t* bars = (uint64_t*)(smem + (size_t)S * 24576);
  const uint32_t bb = saddr(bars);
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2 * S; ++s) mbar_init(bb + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned long long t0 = 0;
  if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0) :: "memory");
  // CTA c streams k-blocks 0..kblocks-1 of row panel c (modes 0/2) or chunks c*kblocks.. (mode 1)
  if (threadIdx.x == 0) {
    int st = 0; uint32_t ph = 0;
    for (int kb = 0; kb < p.kblocks; ++kb) {
      if (kb >= S) mbar_wait(bb + 8 * (S + st), ph ^ 1);
      mbar_expect(bb + 8 * st, stage_bytes + x_bytes);
      const uint32_t dst = saddr(smem + (size_t)st * 24576);
      if (p.mode == 1) {
        const uint8_t* src = p.src + p.base + ((long long)blockIdx.x * p.kblocks + kb) * (long long)stage_bytes;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(stage_bytes), "r"(bb + 8 * st) : "memory");
      } else if (p.mode == 0) {
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(dst), "l"(&tm), "r"(bb + 8 * st), "r"((kb & 63) * 64), "r"((int)(p.base / 8192) + ((int)blockIdx.x * (p.kblocks / 64) + (kb >> 6)) * p.rows) : "memory");
      } else {
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(dst), "l"(&tm), "r"(bb + 8 * st), "r"(0), "r"((int)(p.base / 128) + ((int)blockIdx.x * p.kblocks + kb) * p.rows) : "memory");
      }
      if (x_bytes) {
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(dst + 16384), "l"(&tx), "r"(bb + 8 * st), "r"((kb & 63) * 64), "r"(0) : "memory");
      }
      if (++st == S) { st = 0; ph ^= 1; }
    }
  } else if (threadIdx.x == 32) {
    int st = 0; uint32_t ph = 0;
    for (int kb = 0; kb < p.kblocks; ++kb) {
      mbar_wait(bb + 8 * st, ph);
      mbar_arrive(bb + 8 * (S + st));
      if (++st == S) { st = 0; ph ^= 1; }
    }
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1) :: "memory");
    p.out[blockIdx.x * 2 + 1] = t1;
  }
  if (threadIdx.x == 0) p.out[blockIdx.x * 2] = t0;
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncodeFn enc = (EncodeFn)fp;
  const size_t bytes = (size_t)3 << 30;
  uint8_t* buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 1, bytes));
  uint8_t* xbuf; CK(cudaMalloc(&xbuf, 128 * 8192)); CK(cudaMemset(xbuf, 1, 128 * 8192));
  unsigned long long* out; CK(cudaMalloc(&out, 4096 * 16));
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  const int kblocks = argc > 1 ? atoi(argv[1]) : 256;    // per CTA: 256 = four row panels of K = 4096
  const int grids[] = {48, 96, 128, 148};
  struct Cfg { int mode, rows, stages, xrows; };
  const Cfg cfgs[] = {{0, 128, 4, 0}, {0, 128, 8, 0}, {0, 128, 13, 0}, {0, 64, 8, 0}, {0, 64, 13, 0}, {0, 64, 26, 0},
                      {1, 128, 4, 0}, {1, 128, 8, 0}, {1, 128, 13, 0}, {1, 64, 8, 0}, {1, 64, 13, 0},
                      {2, 128, 8, 0}, {2, 128, 13, 0},
                      {0, 128, 8, 64}, {0, 128, 9, 64}, {0, 128, 8, 16}, {0, 128, 8, 32}, {0, 64, 9, 64}, {1, 128, 8, 64}, {1, 128, 9, 64}};
  CUtensorMap tx;
  {
    cuuint64_t dims[2] = {4096, 128}; cuuint64_t str[1] = {8192}; cuuint32_t es[2] = {1, 1};
    cuuint32_t box[2] = {64, 64};
    if (enc(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, xbuf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) { fprintf(stderr, "encode x\n"); return 1; }
  }
  // read everything once: no dirty lines of the memset left in L2
  {
    CUtensorMap tm;
    cuuint64_t dims[2] = {64, (cuuint64_t)(bytes / 128)}; cuuint64_t str[1] = {128};
    cuuint32_t box[2] = {64, 128}; cuuint32_t es[2] = {1, 1};
    enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    P p{buf, 0, 1, 128, (int)(bytes / 16384 / 148), 8, 0, out};
    stream_kernel<<<148, 64, 8 * 24576 + 2048>>>(tm, tx, p);
    CK(cudaDeviceSynchronize());
  }
  for (const Cfg& c : cfgs) for (int grid : grids) {
    const int mode = c.mode, rows = c.rows, stages = c.stages;
    if (c.xrows && rows < 128 && false) continue;
    CUtensorMap tm;
    cuuint32_t es[2] = {1, 1};
    cuuint32_t box[2] = {64, (cuuint32_t)rows};
    if (mode == 0) {                              // [rows, 4096] bf16
      cuuint64_t dims[2] = {4096, (cuuint64_t)(bytes / 8192)}; cuuint64_t str[1] = {8192};
      if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) { fprintf(stderr, "encode\n"); return 1; }
    } else {                                      // [many, 64] bf16
      cuuint64_t dims[2] = {64, (cuuint64_t)(bytes / 128)}; cuuint64_t str[1] = {128};
      if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) { fprintf(stderr, "encode\n"); return 1; }
    }
    const size_t per_launch = (size_t)grid * kblocks * rows * 128;
    const size_t smem = (size_t)stages * 24576 + 1024 + 2 * stages * 8 + 64;
    if (smem > 227 * 1024) continue;
    double best_us = 1e30, sum_us = 0; int reps = 4;
    for (int r = 0; r < reps; ++r) {
      // every launch reads its own region of the 3 GiB buffer (>= 5 regions): nothing it reads is in L2, nothing in L2 is dirty
      const size_t region = (per_launch + (1 << 20) - 1) & ~(size_t)((1 << 20) - 1);
      static size_t cursor = 0;
      if (cursor + region > bytes) cursor = 0;
      P p{buf, (long long)cursor, mode, rows, kblocks, stages, c.xrows, out};
      cursor += region;
      stream_kernel<<<grid, 64, smem>>>(tm, tx, p);
      CK(cudaDeviceSynchronize());
      std::vector<unsigned long long> h(grid * 2);
      CK(cudaMemcpy(h.data(), out, grid * 16, cudaMemcpyDeviceToHost));
      unsigned long long t0 = ~0ull, t1 = 0;
      for (int cc = 0; cc < grid; ++cc) { if (h[2 * cc] < t0) t0 = h[2 * cc]; if (h[2 * cc + 1] > t1) t1 = h[2 * cc + 1]; }
      const double us = (t1 - t0) / 1e3;
      if (us < best_us) best_us = us;
      sum_us += us;
    }
    const double mb = (double)per_launch / 1e6;
    printf("{\"mode\": %d, \"rows\": %d, \"stages\": %d, \"x_rows\": %d, \"ctas\": %d, \"MB\": %.1f, \"us_best\": %.2f, \"us_mean\": %.2f, \"TB/s_best\": %.2f, \"GB/s_per_cta\": %.1f, \"ns_per_kblock\": %.0f, \"w_in_flight_KB_per_cta\": %d}\n",
           mode, rows, stages, c.xrows, grid, mb, best_us, sum_us / reps, mb / best_us, mb * 1e3 / best_us / grid, best_us * 1e3 / kblocks, stages * rows * 128 / 1024);
    fflush(stdout);
  }
  return 0;
}
