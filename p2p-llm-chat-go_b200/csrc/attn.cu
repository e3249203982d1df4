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

#include "common.cuh"
#include "kernels.h"

namespace b200 {
namespace {

constexpr int kTile = 32;
constexpr int kKStride = 65;  // words per K row in smem (64 data + 1 pad): lane=token reads are conflict-free
constexpr int kPart = kHeadDim + 2;  // acc[128], m, l
constexpr int kMaxSplits = 32;

template <int G>
__global__ void __launch_bounds__(128) attn_decode_kernel(const AttnArgs a) {
  __shared__ uint32_t Ks[kTile * kKStride];
  __shared__ __align__(16) uint16_t Vs[kTile * kHeadDim];
  __shared__ __align__(16) float qs[G * kHeadDim];
  __shared__ float P[G][kTile];
  __shared__ float m_run[G], l_run[G], corr_s[G];
  __shared__ int s_last;

  const int kvh = blockIdx.x, split = blockIdx.y, S = gridDim.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_wait();
  pdl_launch_dependents();
  const int qi = blockIdx.z;  // query position within the launch (0 for decode)
  const int ctx = a.q_ctx ? a.q_ctx[qi] : (a.fixed_ctx > 0 ? a.fixed_ctx + qi : a.st->pos + 1);
  const int32_t* bt = a.q_slot ? a.block_table + (size_t)a.q_slot[qi] * a.pages_per_seq : a.block_table;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  const float* q_in = a.q + (size_t)qi * a.q_stride;
  float* out = a.out + (size_t)qi * a.q_stride;
  float* partial = a.partial + (size_t)qi * gridDim.x * S * G * kPart;
  unsigned int* counters = a.counters + (size_t)qi * gridDim.x;

  for (int i = tid; i < G * kHeadDim; i += 128) qs[i] = q_in[(size_t)kvh * G * kHeadDim + i];
  if (tid < G) {
    m_run[tid] = -INFINITY;
    l_run[tid] = 0.f;
  }
  float acc[G];
#pragma unroll
  for (int h = 0; h < G; ++h) acc[h] = 0.f;
  __syncthreads();

  for (int tile = split; tile * kTile < ctx; tile += S) {
    const int t0 = tile * kTile;
    // gather K and V rows of this tile (zero-fill past the context end)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int item = tid + 128 * j, t = item >> 4, seg = item & 15;
      const int tok = t0 + t;
      uint4 kv = make_uint4(0, 0, 0, 0), vv = make_uint4(0, 0, 0, 0);
      if (tok < ctx) {
        const int page = bt[tok / kPageTokens];
        const size_t row = (((size_t)page * a.n_kv_heads + kvh) * kPageTokens + (tok % kPageTokens)) * kHeadDim;
        kv = *reinterpret_cast<const uint4*>(a.k_cache + row + seg * 8);
        vv = *reinterpret_cast<const uint4*>(a.v_cache + row + seg * 8);
      }
      uint32_t* kd = Ks + t * kKStride + seg * 4;
      kd[0] = kv.x; kd[1] = kv.y; kd[2] = kv.z; kd[3] = kv.w;
      *reinterpret_cast<uint4*>(Vs + t * kHeadDim + seg * 8) = vv;
    }
    __syncthreads();
    // scores + online softmax: one head per warp, lane = token
    for (int h = warp; h < G; h += 4) {
      const uint32_t* kr = Ks + lane * kKStride;
      const float* qh = qs + h * kHeadDim;
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;  // four chains: the dot product is latency-, not issue-bound
#pragma unroll 8
      for (int w = 0; w < kHeadDim / 2; w += 2) {
        const uint32_t ka = kr[w], kb = kr[w + 1];
        const float4 q4 = *reinterpret_cast<const float4*>(qh + 2 * w);
        s0 = fmaf(q4.x, bf16lo(ka), s0);
        s1 = fmaf(q4.y, bf16hi(ka), s1);
        s2 = fmaf(q4.z, bf16lo(kb), s2);
        s3 = fmaf(q4.w, bf16hi(kb), s3);
      }
      float s = (s0 + s1) + (s2 + s3);
      const bool valid = (t0 + lane) < ctx;
      s = valid ? s * scale : -INFINITY;
      const float m_old = m_run[h];
      const float m_new = fmaxf(m_old, warp_max(s));
      const float p = valid ? expf(s - m_new) : 0.f;
      const float l_tile = warp_sum(p);
      P[h][lane] = p;
      if (lane == 0) {
        const float corr = (m_old == -INFINITY) ? 0.f : expf(m_old - m_new);
        corr_s[h] = corr;
        l_run[h] = l_run[h] * corr + l_tile;
        m_run[h] = m_new;
      }
    }
    __syncthreads();
    // P.V: thread = output dim
#pragma unroll
    for (int h = 0; h < G; ++h) acc[h] *= corr_s[h];
#pragma unroll 8
    for (int t = 0; t < kTile; ++t) {
      const float v = __uint_as_float((uint32_t)Vs[t * kHeadDim + tid] << 16);
#pragma unroll
      for (int h = 0; h < G; ++h) acc[h] = fmaf(P[h][t], v, acc[h]);
    }
    __syncthreads();
  }

  float* part = partial + ((size_t)kvh * S + split) * G * kPart;
#pragma unroll
  for (int h = 0; h < G; ++h) part[h * kPart + tid] = acc[h];
  if (tid < G) {
    part[tid * kPart + kHeadDim] = m_run[tid];
    part[tid * kPart + kHeadDim + 1] = l_run[tid];
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(counters + kvh, 1u) == (unsigned)(S - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // merge the splits: (m, l) of every (head, split) into shared memory with one parallel round of loads,
  // softmax weights per split, then one pass over the partial accumulators with independent loads
  const float* base = partial + (size_t)kvh * S * G * kPart;
  float* w_s = reinterpret_cast<float*>(Ks);  // [G][kMaxSplits] softmax weight of every split (reuses the K tile)
  for (int h = warp; h < G; h += 4) {       // one head per warp, lane = split: a single round of loads
    const bool on = lane < S;
    const float* ps = base + ((size_t)(on ? lane : 0) * G + h) * kPart;
    const float ms = on ? __ldcg(ps + kHeadDim) : -INFINITY;
    const float ls = on ? __ldcg(ps + kHeadDim + 1) : 0.f;
    const float M = warp_max(ms);
    const float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
    const float den = warp_sum(ls * w);
    w_s[h * kMaxSplits + lane] = w;
    if (lane == 0) corr_s[h] = 1.0f / den;
  }
  __syncthreads();
  float num[G];
#pragma unroll
  for (int h = 0; h < G; ++h) num[h] = 0.f;
  for (int sp = 0; sp < S; sp += 4) {  // 4 x G independent loads in flight per thread
    float p[4][G];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int h = 0; h < G; ++h) p[u][h] = (sp + u < S) ? __ldcg(base + ((size_t)(sp + u) * G + h) * kPart + tid) : 0.f;
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int h = 0; h < G; ++h) num[h] = fmaf(p[u][h], (sp + u < S) ? w_s[h * kMaxSplits + sp + u] : 0.f, num[h]);
  }
#pragma unroll
  for (int h = 0; h < G; ++h) out[((size_t)kvh * G + h) * kHeadDim + tid] = num[h] * corr_s[h];
  if (tid == 0) counters[kvh] = 0u;
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
