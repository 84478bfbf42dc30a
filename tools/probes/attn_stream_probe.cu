// Micro-benchmark (measurement only): the decode-attention K/V stream without any math.  296 CTAs (two per SM), three
// 32 KB stages each; a stage is one 64-token K block and one V block of one kv head, 16 KB contiguous each, fetched either
//   mode 0: as the kernel does today - the cache as a [rows, 128] matrix (256-byte rows), boxes of 64 columns x 64 rows:
//           two half-row boxes per block (SWIZZLE_128B cannot take 256-byte rows)
//   mode 1: the cache as a [2 * rows, 64] matrix (128-byte rows), ONE box of 64 x 128 per block: 16 KB contiguous
//   mode 2: cp.async.bulk of the 16 KB
// Build + run: nvcc -arch=sm_100a -o /tmp/asp tools/probes/attn_stream_probe.cu -lcuda && /tmp/asp
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
__device__ __forceinline__ void tma2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
constexpr int S = 3;
struct P { const uint8_t* k; const uint8_t* v; int mode, tiles; const int* perm; unsigned long long* out; };

__global__ void __launch_bounds__(64, 2) kern(const __grid_constant__ CUtensorMap tk, const __grid_constant__ CUtensorMap tv, const P p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + S * 32768);
  const uint32_t bb = saddr(bars);
  if (threadIdx.x == 0) { for (int s = 0; s < 2 * S; ++s) mbar_init(bb + 8 * s, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  unsigned long long t0 = 0;
  if (threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0) :: "memory");
  if (threadIdx.x == 0) {
    int st = 0; uint32_t ph = 0;
    for (int i = 0; i < p.tiles; ++i) {
      const int blk = p.perm[blockIdx.x * p.tiles + i];           // a random 16 KB block, like a paged cache
      if (i >= S) mbar_wait(bb + 8 * (S + st), ph ^ 1);
      mbar_expect(bb + 8 * st, 32768);
      const uint32_t dk = saddr(smem + st * 32768), dv = dk + 16384;
      if (p.mode == 0) {
        tma2d(dk, &tk, bb + 8 * st, 0, blk * 64); tma2d(dv, &tv, bb + 8 * st, 0, blk * 64);
        tma2d(dk + 8192, &tk, bb + 8 * st, 64, blk * 64); tma2d(dv + 8192, &tv, bb + 8 * st, 64, blk * 64);
      } else if (p.mode == 1) {
        tma2d(dk, &tk, bb + 8 * st, 0, blk * 128); tma2d(dv, &tv, bb + 8 * st, 0, blk * 128);
      } else {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dk), "l"(p.k + (size_t)blk * 16384), "r"(16384), "r"(bb + 8 * st) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dv), "l"(p.v + (size_t)blk * 16384), "r"(16384), "r"(bb + 8 * st) : "memory");
      }
      if (++st == S) { st = 0; ph ^= 1; }
    }
  } else if (threadIdx.x == 32) {
    int st = 0; uint32_t ph = 0;
    for (int i = 0; i < p.tiles; ++i) { mbar_wait(bb + 8 * st, ph); mbar_arrive(bb + 8 * (S + st)); if (++st == S) { st = 0; ph ^= 1; } }
    unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1) :: "memory");
    p.out[blockIdx.x * 2 + 1] = t1;
  }
  if (threadIdx.x == 0) p.out[blockIdx.x * 2] = t0;
}
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncodeFn enc = (EncodeFn)fp;
  const int grid = 296, tiles = 59, nblk = 1 << 16;               // 64 Ki blocks x 16 KB = 1 GiB per pool
  uint8_t *k, *v; CK(cudaMalloc(&k, (size_t)nblk * 16384)); CK(cudaMalloc(&v, (size_t)nblk * 16384));
  unsigned long long* out; CK(cudaMalloc(&out, 4096 * 16));
  std::vector<int> perm(nblk); for (int i = 0; i < nblk; ++i) perm[i] = i;
  srand(1); for (int i = nblk - 1; i > 0; --i) { int j = rand() % (i + 1); std::swap(perm[i], perm[j]); }
  int* dperm; CK(cudaMalloc(&dperm, nblk * 4)); CK(cudaMemcpy(dperm, perm.data(), nblk * 4, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  for (int mode = 0; mode < 3; ++mode) {
    CUtensorMap tk, tv;
    cuuint32_t es[2] = {1, 1};
    for (int w = 0; w < 2; ++w) {
      void* base = w ? (void*)v : (void*)k;
      CUtensorMap* tm = w ? &tv : &tk;
      if (mode == 0) {
        cuuint64_t dims[2] = {128, (cuuint64_t)nblk * 64}; cuuint64_t str[1] = {256}; cuuint32_t box[2] = {64, 64};
        if (enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) return 1;
      } else {
        cuuint64_t dims[2] = {64, (cuuint64_t)nblk * 128}; cuuint64_t str[1] = {128}; cuuint32_t box[2] = {64, 128};
        if (enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) return 1;
      }
    }
    for (int rep = 0; rep < 4; ++rep) {
      // every repetition reads other blocks (offset into the permutation): nothing it reads is in L2
      P p{k, v, mode, tiles, dperm + ((rep * grid * tiles) % (nblk - grid * tiles)), out};
      kern<<<grid, 64, S * 32768 + 1024 + 128>>>(tk, tv, p);
      CK(cudaDeviceSynchronize());
      std::vector<unsigned long long> h(grid * 2);
      CK(cudaMemcpy(h.data(), out, grid * 16, cudaMemcpyDeviceToHost));
      unsigned long long t0 = ~0ull, t1 = 0;
      for (int c = 0; c < grid; ++c) { if (h[2 * c] < t0) t0 = h[2 * c]; if (h[2 * c + 1] > t1) t1 = h[2 * c + 1]; }
      const double us = (t1 - t0) / 1e3, mb = (double)grid * tiles * 32768 / 1e6;
      printf("{\"mode\": %d, \"rep\": %d, \"MB\": %.1f, \"us\": %.2f, \"TB/s\": %.2f}\n", mode, rep, mb, us, mb / us);
    }
  }
  return 0;
}
