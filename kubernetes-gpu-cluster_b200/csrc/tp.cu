// Tensor-parallel plumbing: NCCL communicator per engine (one rank per GPU / process) and the
// bf16 sum all-reduce after the row-parallel projections (o_proj / down_proj, fc2 / out_proj):
// vllm:model_executor/layers/linear.py:1559; the reference's TP values files force the NCCL path
// with --disable-custom-all-reduce (ref:values-01-minimal-example8.yaml:32).
// NCCL is loaded with dlopen at TP initialisation so a single-GPU engine (and a Python process
// that already carries torch's bundled NCCL) has no link-time dependency on it.
#include <dlfcn.h>
#include <string.h>

#include "engine.h"

namespace kge {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  if (api.lib) return &api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) { set_error("cannot load libnccl.so.2: %s", dlerror()); return nullptr; }
  api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.lib, "ncclGetUniqueId");
  api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.lib, "ncclCommInitRank");
  api.AllReduce = (decltype(api.AllReduce))dlsym(api.lib, "ncclAllReduce");
  api.AllGather = (decltype(api.AllGather))dlsym(api.lib, "ncclAllGather");
  api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.lib, "ncclCommDestroy");
  api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.lib, "ncclGetErrorString");
  if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.AllGather || !api.CommDestroy || !api.GetErrorString) {
    set_error("libnccl is missing a required symbol");
    api.lib = nullptr;
    return nullptr;
  }
  return &api;
}

int tp_all_reduce(kge_engine* e, bf16* buf, int64_t n) {
  KGE_REQUIRE(e->comm, "tensor parallel size %d but neither kge_engine_tp_connect nor kge_engine_init_tp was called", e->tp_size);
  NcclApi* api = nccl_api();
  if (!api) return -1;
  ncclResult_t r = api->AllReduce(buf, buf, (size_t)n, e->f16 ? ncclFloat16 : ncclBfloat16, ncclSum, e->comm, e->stream);
  KGE_REQUIRE(r == ncclSuccess, "ncclAllReduce: %s", api->GetErrorString(r));
  pdl_break_next();        // whatever follows the NCCL kernel is launched in plain stream order
  return 0;
}

// What ncclAllGather delivers - rank k's block at [k][R * shard], holding its [R, cols_k] logits densely (cols_k = the
// columns rank k really has: the last rank's shard may be narrower) - laid out as logits [R][vocab].
__global__ void reorder_logits_kernel(bf16* __restrict__ dst, const bf16* __restrict__ src, int R, int shard, int vocab, int tp) {
  const int64_t n = (int64_t)R * vocab;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int col = (int)(i % vocab), r = (int)(i / vocab);
    const int k = col / shard, c = col - k * shard;
    const int cols_k = min(shard, vocab - k * shard);
    dst[i] = src[(int64_t)k * R * shard + (int64_t)r * cols_k + c];
  }
}

// NCCL path of the vocab-sharded LM head: all ranks' [R, vocab_shard] blocks gathered, then laid out as [R, vocab].
int tp_all_gather_logits(kge_engine* e, int R) {
  KGE_REQUIRE(e->comm, "tensor parallel size %d but neither kge_engine_tp_connect nor kge_engine_init_tp was called", e->tp_size);
  NcclApi* api = nccl_api();
  if (!api) return -1;
  if (!e->logits_gather) {
    KGE_CUDA_OK(cudaMalloc((void**)&e->logits_gather, (size_t)e->tp_size * e->cfg.max_num_seqs * e->vocab_shard * 2));
    e->allocs.push_back(e->logits_gather);
  }
  ncclResult_t r = api->AllGather(e->logits_loc, e->logits_gather, (size_t)R * e->vocab_shard, ncclBfloat16, e->comm, e->stream);
  KGE_REQUIRE(r == ncclSuccess, "ncclAllGather: %s", api->GetErrorString(r));
  reorder_logits_kernel<<<148 * 2, 256, 0, e->stream>>>(e->logits, e->logits_gather, R, e->vocab_shard, e->spec.vocab, e->tp_size);
  KGE_LAUNCH_OK();
  pdl_break_next();
  return 0;
}

}  // namespace kge

using namespace kge;

extern "C" int kge_engine_tp_info(kge_engine* e, kge_tp_peer_info* out) {
  KGE_REQUIRE(e && out, "kge_engine_tp_info: null argument");
  KGE_REQUIRE(e->tpc, "kge_engine_tp_info: the engine was created with tensor_parallel_size 1");
  return kge_tp_comm_info(e->tpc, out);
}

extern "C" int kge_engine_tp_connect(kge_engine* e, const kge_tp_peer_info* all) {
  KGE_REQUIRE(e && all, "kge_engine_tp_connect: null argument");
  KGE_REQUIRE(e->tpc, "kge_engine_tp_connect: the engine was created with tensor_parallel_size 1");
  if (int rc = kge_tp_comm_connect(e->tpc, all)) return rc;
  // the exchange kernels are the engine's own: programmatic dependent launch and captured decode steps are safe again
  if (!getenv("KGE_DISABLE_PDL")) kge::pdl_set_enabled(true);
  e->use_graphs = getenv("KGE_DISABLE_GRAPHS") == nullptr;
  return 0;
}

extern "C" int kge_nccl_unique_id(char* out128) {
  KGE_REQUIRE(out128, "kge_nccl_unique_id: null argument");
  NcclApi* api = nccl_api();
  if (!api) return -1;
  ncclUniqueId id;
  ncclResult_t r = api->GetUniqueId(&id);
  KGE_REQUIRE(r == ncclSuccess, "ncclGetUniqueId: %s", api->GetErrorString(r));
  static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
  memcpy(out128, &id, 128);
  return 0;
}

extern "C" int kge_engine_init_tp(kge_engine* e, const char* id128) {
  KGE_REQUIRE(e && id128, "kge_engine_init_tp: null argument");
  if (e->tp_size == 1) return 0;
  NcclApi* api = nccl_api();
  if (!api) return -1;
  KGE_CUDA_OK(cudaSetDevice(e->cfg.device));
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  ncclResult_t r = api->CommInitRank(&e->comm, e->tp_size, id, e->tp_rank);
  KGE_REQUIRE(r == ncclSuccess, "ncclCommInitRank: %s", api->GetErrorString(r));
  return 0;
}
