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
  api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.lib, "ncclCommDestroy");
  api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.lib, "ncclGetErrorString");
  if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy || !api.GetErrorString) {
    set_error("libnccl is missing a required symbol");
    api.lib = nullptr;
    return nullptr;
  }
  return &api;
}

int tp_all_reduce(kge_engine* e, bf16* buf, int64_t n) {
  KGE_REQUIRE(e->comm, "tensor parallel size %d but kge_engine_init_tp was not called", e->tp_size);
  NcclApi* api = nccl_api();
  if (!api) return -1;
  ncclResult_t r = api->AllReduce(buf, buf, (size_t)n, ncclBfloat16, ncclSum, e->comm, e->stream);
  KGE_REQUIRE(r == ncclSuccess, "ncclAllReduce: %s", api->GetErrorString(r));
  pdl_break_next();        // whatever follows the NCCL kernel is launched in plain stream order
  return 0;
}

}  // namespace kge

using namespace kge;

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
