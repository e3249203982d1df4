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
#include <unistd.h>

#include <cstring>
#include <vector>

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
int Engine::tp_allreduce_max_u64_n(unsigned long long* buf, size_t n, cudaStream_t s) {
  ncclResult_t r = nccl()->AllReduce(buf, buf, n, ncclUint64, ncclMax, (ncclComm_t)nccl_comm, s);
  return r == ncclSuccess ? 0 : nccl_fail(r, "ncclAllReduce(max)");
}
int Engine::tp_allgather(const float* send, float* recv, size_t n_per_rank, cudaStream_t s) {
  ncclResult_t r = nccl()->AllGather(send, recv, n_per_rank, ncclFloat32, (ncclComm_t)nccl_comm, s);
  return r == ncclSuccess ? 0 : nccl_fail(r, "ncclAllGather");
}

}  // namespace b200

// ---- fused all-reduce over peer memory -------------------------------------------------------------------
// Every rank owns two receive buffers ([tp][d_model] fp32, alternating between consecutive all-reduces) and a
// flag word per (buffer, source rank).  Their addresses are exchanged once with an NCCL all-gather: ranks that
// live in the same process use the raw pointers after cudaDeviceEnablePeerAccess, ranks in other processes
// (torchrun) open CUDA IPC handles.  From then on the decode step needs NCCL only for the 8-byte argmax.
namespace b200 {

struct TpXchg {
  long long pid;
  int device, pad;
  void *recv0, *recv1, *flags;
  cudaIpcMemHandle_t h0, h1, hf;
  void *brecv0, *brecv1;          // batched-decode receive buffers (flags live in the same small block as `flags`)
  cudaIpcMemHandle_t hb0, hb1;
};

int Engine::tp_setup_link() {
  const int d = cfg.d_model, n = tp_size;
  if (n > kMaxTp) { set_error("tensor parallel degree %d > %d", n, kMaxTp); return B200_EINVAL; }
  TpLink& L = tp_link_host;
  memset(&L, 0, sizeof L);
  L.rank = tp_rank; L.size = n; L.d = d; L.per_step = 2 * cfg.n_layers;
  unsigned int* small = nullptr;  // flags[2][kMaxTp] | epoch | ticket | error
  // every float travels with its exchange's sequence number beside it (kernels.h: TpLink), hence 8 bytes per element
  B200_CUDA(cudaMalloc((void**)&L.recv[0], (size_t)n * d * 8));
  B200_CUDA(cudaMalloc((void**)&L.recv[1], (size_t)n * d * 8));
  B200_CUDA(cudaMalloc((void**)&small, 256));
  allocs.push_back(L.recv[0]); allocs.push_back(L.recv[1]); allocs.push_back(small);
  B200_CUDA(cudaMemset(L.recv[0], 0, (size_t)n * d * 8));
  B200_CUDA(cudaMemset(L.recv[1], 0, (size_t)n * d * 8));
  B200_CUDA(cudaMemset(small, 0, 256));
  if (d % 64) { set_error("tensor parallelism needs d_model %% 64 == 0 (got %d)", d); return B200_EINVAL; }
  for (int b = 0; b < 2; ++b) {
    B200_CUDA(cudaMalloc((void**)&L.xsum[b], (size_t)d * 8));
    allocs.push_back(L.xsum[b]);
    B200_CUDA(cudaMemset(L.xsum[b], 0, (size_t)d * 8));
  }
  L.flags = small; L.epoch = small + 2 * kMaxTp; L.ticket = small + 2 * kMaxTp + 1; L.error = (int*)(small + 2 * kMaxTp + 2);
  // batched decode: [tp][kTpBatchRows * d] per buffer; flags [2][kMaxTp] and a ticket further into `small`
  TpBatch& Bt = tp_batch;
  memset(&Bt, 0, sizeof Bt);
  Bt.rank = tp_rank; Bt.size = n; Bt.slot_floats = (long long)kTpBatchRows * d;
  const size_t bbytes = (size_t)n * kTpBatchRows * d * 4;
  B200_CUDA(cudaMalloc((void**)&Bt.recv[0], bbytes));
  B200_CUDA(cudaMalloc((void**)&Bt.recv[1], bbytes));
  allocs.push_back(Bt.recv[0]); allocs.push_back(Bt.recv[1]);
  B200_CUDA(cudaMemset(Bt.recv[0], 0, bbytes));
  B200_CUDA(cudaMemset(Bt.recv[1], 0, bbytes));
  constexpr int kBFlagsOff = 2 * kMaxTp + 8;  // words into `small`
  static_assert((kBFlagsOff + 2 * kMaxTp + 1) * 4 <= 256, "flag block too small");
  Bt.flags = small + kBFlagsOff; Bt.ticket = small + kBFlagsOff + 2 * kMaxTp; Bt.error = L.error;
  // measured on 8 x B200 (70B, 32 sequences): 17.1 ms per step with the push against 15.1 ms with NCCL, whose 1 MB
  // all-reduce is reduced inside the NVSwitch (NVLS) in ~28 us while the push moves 8 x 1 MB per rank: opt-in
  tp_batch_push = getenv("B200_TP_PUSH") != nullptr && getenv("B200_TP_NCCL") == nullptr;
  int rc;
  if ((rc = alloc((void**)&x_alt, (size_t)d * 4)) || (rc = alloc((void**)&tp_link, sizeof(TpLink)))) return rc;

  TpXchg mine = {};
  mine.pid = (long long)getpid(); mine.device = cfg.device;
  mine.recv0 = L.recv[0]; mine.recv1 = L.recv[1]; mine.flags = L.flags;
  B200_CUDA(cudaIpcGetMemHandle(&mine.h0, L.recv[0]));
  B200_CUDA(cudaIpcGetMemHandle(&mine.h1, L.recv[1]));
  B200_CUDA(cudaIpcGetMemHandle(&mine.hf, small));
  mine.brecv0 = Bt.recv[0]; mine.brecv1 = Bt.recv[1];
  B200_CUDA(cudaIpcGetMemHandle(&mine.hb0, Bt.recv[0]));
  B200_CUDA(cudaIpcGetMemHandle(&mine.hb1, Bt.recv[1]));
  TpXchg* dev_all = nullptr;
  B200_CUDA(cudaMalloc((void**)&dev_all, sizeof(TpXchg) * n));
  B200_CUDA(cudaMemcpy(dev_all + tp_rank, &mine, sizeof mine, cudaMemcpyHostToDevice));
  ncclResult_t r = nccl()->AllGather(dev_all + tp_rank, dev_all, sizeof(TpXchg), ncclChar, (ncclComm_t)nccl_comm, stream);
  if (r != ncclSuccess) { cudaFree(dev_all); return nccl_fail(r, "ncclAllGather(link)"); }
  std::vector<TpXchg> all(n);
  cudaError_t e = cudaMemcpyAsync(all.data(), dev_all, sizeof(TpXchg) * n, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  cudaFree(dev_all);
  if (e != cudaSuccess) return cuda_fail(e, "link exchange");
  for (int p = 0; p < n; ++p) {
    if (p == tp_rank) {
      L.peer_recv[0][p] = L.recv[0]; L.peer_recv[1][p] = L.recv[1]; L.peer_flags[p] = L.flags;
      Bt.peer_recv[0][p] = Bt.recv[0]; Bt.peer_recv[1][p] = Bt.recv[1]; Bt.peer_flags[p] = Bt.flags;
    } else if (all[p].pid == mine.pid) {  // a thread of this process drives that GPU: map it directly
      e = cudaDeviceEnablePeerAccess(all[p].device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
      (void)cudaGetLastError();
      L.peer_recv[0][p] = (float*)all[p].recv0; L.peer_recv[1][p] = (float*)all[p].recv1; L.peer_flags[p] = (unsigned int*)all[p].flags;
      Bt.peer_recv[0][p] = (float*)all[p].brecv0; Bt.peer_recv[1][p] = (float*)all[p].brecv1;
      Bt.peer_flags[p] = (unsigned int*)all[p].flags + kBFlagsOff;
    } else {
      void *a = nullptr, *b = nullptr, *f = nullptr;
      B200_CUDA(cudaIpcOpenMemHandle(&a, all[p].h0, cudaIpcMemLazyEnablePeerAccess));
      B200_CUDA(cudaIpcOpenMemHandle(&b, all[p].h1, cudaIpcMemLazyEnablePeerAccess));
      B200_CUDA(cudaIpcOpenMemHandle(&f, all[p].hf, cudaIpcMemLazyEnablePeerAccess));
      tp_ipc_opened.push_back(a); tp_ipc_opened.push_back(b); tp_ipc_opened.push_back(f);
      L.peer_recv[0][p] = (float*)a; L.peer_recv[1][p] = (float*)b; L.peer_flags[p] = (unsigned int*)f;
      void *ba = nullptr, *bb = nullptr;
      B200_CUDA(cudaIpcOpenMemHandle(&ba, all[p].hb0, cudaIpcMemLazyEnablePeerAccess));
      B200_CUDA(cudaIpcOpenMemHandle(&bb, all[p].hb1, cudaIpcMemLazyEnablePeerAccess));
      tp_ipc_opened.push_back(ba); tp_ipc_opened.push_back(bb);
      Bt.peer_recv[0][p] = (float*)ba; Bt.peer_recv[1][p] = (float*)bb; Bt.peer_flags[p] = (unsigned int*)f + kBFlagsOff;
    }
  }
  B200_CUDA(cudaMemcpy(tp_link, &L, sizeof L, cudaMemcpyHostToDevice));
  // nobody may push before every rank has mapped everything: one more collective as a barrier
  float* one = tp_partial;
  if ((rc = tp_allreduce_sum(one, 1, stream))) return rc;
  B200_CUDA(cudaStreamSynchronize(stream));
  tp_fused = true;
  return 0;
}

int Engine::tp_check_error() {
  if (!tp_fused) return 0;
  int err = 0;
  B200_CUDA(cudaMemcpy(&err, tp_link_host.error, 4, cudaMemcpyDeviceToHost));
  if (err) { set_error("tensor-parallel peer did not deliver its partial sums in time"); return B200_ETIMEOUT; }
  return 0;
}

}  // namespace b200
