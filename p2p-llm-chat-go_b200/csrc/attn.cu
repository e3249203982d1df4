// attn.cu — paged GQA decode attention for one query position (batch-1 decode), sm_100a.
//
// grid = (n_kv_heads, n_splits), 128 threads.  A CTA owns one KV head (its G = n_heads/n_kv_heads query
// heads share every K/V row it reads) and the 32-token tiles  split, split+n_splits, ...  of the context,
// so the grid is fixed (CUDA-graph friendly) while the context grows.  Per tile: K and V rows are gathered
// through the block table into shared memory (bf16), scores and the online-softmax update are computed
// one head per warp (lane = token), P.V one output dim per thread.  Each CTA leaves an un-normalised
// partial (acc, m, l) in scratch; the last CTA of a KV head to finish (atomic ticket) merges the splits and
// writes the attention output, so no second launch is needed.  The ticket counters reset themselves.
//
// HBM traffic: ctx * 2 * 256 B per KV head (bf16 K and V rows, read once) — 4 KiB per context token per
// layer for Llama-3.1 (8 KV heads); the kernel is latency-, not bandwidth-bound at ctx <= 640.
#include <mutex>

#include "attn_decode.cuh"

namespace b200 {
namespace {

template <int G>
__global__ void __launch_bounds__(128) attn_decode_kernel(const AttnArgs a) {
  __shared__ AttnSmem<G> sm;
  pdl_wait();
  pdl_launch_dependents();
  // blockIdx.z: query position within the launch (0 for decode)
  attn_decode_item<G>(a, sm, blockIdx.x, blockIdx.y, gridDim.y, blockIdx.z, gridDim.x, threadIdx.x, 0);
}

}  // namespace

size_t attn_partial_floats(int n_heads, int n_kv_heads, int n_splits) {
  return (size_t)n_kv_heads * n_splits * (n_heads / n_kv_heads) * kPart;
}

cudaError_t launch_attn_decode(const AttnArgs& a, cudaStream_t s, bool pdl) {
  if (a.n_kv_heads <= 0 || a.n_heads % a.n_kv_heads || a.n_splits <= 0 || a.n_splits > kMaxSplits) return cudaErrorInvalidValue;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a.n_kv_heads, a.n_splits, a.n_queries > 1 ? a.n_queries : 1);
  cfg.blockDim = dim3(128);
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  switch (a.n_heads / a.n_kv_heads) {
    case 1: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<1>, a);
    case 2: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<2>, a);
    case 4: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<4>, a);
    case 8: return cudaLaunchKernelEx(&cfg, attn_decode_kernel<8>, a);
  }
  return cudaErrorInvalidValue;
}


// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant (b200_attn_decode_noalloc).
//
// Some hosts of this library have already handed every byte of device memory to an allocator of their own
// (a serving process that sized its KV pool to fill HBM, a PyTorch caching allocator) or run the call on a
// path where a device-wide stall is not acceptable.  For them a cudaMalloc inside the library can fail for
// lack of memory, and it synchronises the whole device even when it succeeds.  This variant therefore takes
// every input and output as a device pointer from the caller and keeps the kernel's scratch (split
// partials + ticket counters) in a fixed array inside the library image.  The array is paid for once by
// every GPU that loads libb200llm.so (kNoAllocScratchBytes, ~0.64 MB) and lives until the process exits.
//
// `g_noalloc_scratch` below is a *host-side handle* for that array: host code must never pass its name to
// a kernel.  The device address is looked up once per device, with that device current, through
// cudaGetSymbolAddress, and cached.  That look-up happens in the first call made on a device; make that
// first call before any stream capture is started.
//
// Every call on a device shares the one array, which still holds the previous call's partials.  Only the
// ticket counters have to start at zero and they are cleared on the caller's stream before each launch.
// Two calls that overlap on the same GPU (different streams, or different threads) race on the scratch:
// keep calls to this entry point on one GPU in stream order.
constexpr int kNoAllocMaxHeads = 64;   // == B200_NOALLOC_MAX_HEADS (Llama-3.1-70B has 64 query heads)
constexpr int kNoAllocSplits = 20;
constexpr size_t kNoAllocPartialFloats = (size_t)kNoAllocMaxHeads * kNoAllocSplits * kPart;
constexpr size_t kNoAllocScratchBytes = kNoAllocPartialFloats * sizeof(float) + kNoAllocMaxHeads * sizeof(unsigned int);
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_noalloc_ptr[kMaxDevices];  // device address of g_noalloc_scratch, per device ordinal
static std::mutex g_noalloc_mu;

// Returns the device address of the scratch array on the CURRENT device (set-up on first use).
static cudaError_t noalloc_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return cudaSuccess;
}

int attn_noalloc_max_heads() { return kNoAllocMaxHeads; }

// All pointers are device pointers owned by the caller; nothing is allocated or freed here.
cudaError_t launch_attn_decode_noalloc(const float* q, const uint16_t* k_cache, const uint16_t* v_cache,
                                       const int32_t* block_table, int n_ctx, int n_heads, int n_kv_heads, float* out,
                                       cudaStream_t s) {
  if (n_heads < 1 || n_heads > kNoAllocMaxHeads || n_kv_heads < 1 || n_heads % n_kv_heads || n_ctx < 1)
    return cudaErrorInvalidValue;
  void* scratch = nullptr;
  cudaError_t e = noalloc_scratch(&scratch);
  if (e != cudaSuccess) return e;
  float* partial = reinterpret_cast<float*>(scratch);
  unsigned int* counters = reinterpret_cast<unsigned int*>(partial + kNoAllocPartialFloats);
  e = cudaMemsetAsync(counters, 0, (size_t)n_kv_heads * sizeof(unsigned int), s);  // must start at zero
  if (e != cudaSuccess) return e;
  AttnArgs a = {};
  a.q = q; a.k_cache = k_cache; a.v_cache = v_cache; a.block_table = block_table; a.st = nullptr;
  a.n_heads = n_heads; a.n_kv_heads = n_kv_heads; a.out = out; a.partial = partial; a.counters = counters;
  a.n_splits = kNoAllocSplits; a.fixed_ctx = n_ctx;
  return launch_attn_decode(a, s, false);
}

}  // namespace b200
