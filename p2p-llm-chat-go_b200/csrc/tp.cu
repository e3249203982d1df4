// tp.cu — tensor parallelism plumbing: NCCL over NVLink / NVSwitch for the 70B configuration.
//
// Column-parallel qkv / gate / up and the vocab-parallel lm_head need no communication; the row-parallel
// o and down projections produce partial sums of the [d_model] residual update that are all-reduced
// (2 per layer; 16 KiB at batch 1 for d_model 8192 in fp32, so latency- not bandwidth-bound), and the
// greedy token is the max over ranks of a packed 64-bit (logit, ~id) key.
// libnccl is opened at run time (dlopen) so that single-GPU users need no NCCL at all; whichever
// libnccl.so.2 the process already has (e.g. PyTorch's) is the one that gets used.
#include <dlfcn.h>
#include <nccl.h>

#include "engine.h"

namespace b200 {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  const char* (*GetErrorString)(ncclResult_t);
};

static NcclApi* nccl() {
  static NcclApi api;
  static bool tried = false, ok = false;
  if (!tried) {
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (h) {
      api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
      api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
      api.AllGather = (decltype(api.AllGather))dlsym(h, "ncclAllGather");
      api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
      api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
      ok = api.GetUniqueId && api.CommInitRank && api.AllReduce && api.AllGather && api.CommDestroy && api.GetErrorString;
    }
  }
  return ok ? &api : nullptr;
}

static int nccl_fail(ncclResult_t r, const char* what) {
  NcclApi* a = nccl();
  set_error("NCCL error %d (%s) at %s", (int)r, a ? a->GetErrorString(r) : "?", what);
  return B200_ECUDA;
}

int tp_unique_id(void* out128) {
  NcclApi* a = nccl();
  if (!a) { set_error("libnccl.so.2 not found"); return B200_ENODEV; }
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  ncclResult_t r = a->GetUniqueId(&id);
  if (r != ncclSuccess) return nccl_fail(r, "ncclGetUniqueId");
  memcpy(out128, &id, 128);
  return 0;
}

int Engine::tp_init(const void* unique_id) {
  if (tp_size <= 1) return 0;
  NcclApi* a = nccl();
  if (!a) { set_error("libnccl.so.2 not found"); return B200_ENODEV; }
  ncclUniqueId id;
  memcpy(&id, unique_id, 128);
  ncclComm_t comm;
  ncclResult_t r = a->CommInitRank(&comm, tp_size, id, tp_rank);
  if (r != ncclSuccess) return nccl_fail(r, "ncclCommInitRank");
  nccl_comm = comm;
  return 0;
}

void tp_destroy(void* comm) {
  NcclApi* a = nccl();
  if (a && comm) a->CommDestroy((ncclComm_t)comm);
}

int Engine::tp_allreduce_sum(float* buf, size_t n, cudaStream_t s) {
  ncclResult_t r = nccl()->AllReduce(buf, buf, n, ncclFloat32, ncclSum, (ncclComm_t)nccl_comm, s);
  return r == ncclSuccess ? 0 : nccl_fail(r, "ncclAllReduce(sum)");
}
int Engine::tp_allreduce_max_u64(unsigned long long* buf, cudaStream_t s) {
  ncclResult_t r = nccl()->AllReduce(buf, buf, 1, ncclUint64, ncclMax, (ncclComm_t)nccl_comm, s);
  return r == ncclSuccess ? 0 : nccl_fail(r, "ncclAllReduce(max)");
}
int Engine::tp_allgather(const float* send, float* recv, size_t n_per_rank, cudaStream_t s) {
  ncclResult_t r = nccl()->AllGather(send, recv, n_per_rank, ncclFloat32, (ncclComm_t)nccl_comm, s);
  return r == ncclSuccess ? 0 : nccl_fail(r, "ncclAllGather");
}

}  // namespace b200
