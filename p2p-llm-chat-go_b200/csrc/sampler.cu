// sampler.cu — token sampling on the device (one CTA, 1024 threads, logits stay in L2).
//
// Order of operations = llama.cpp's default sampler chain, which is what Ollama runs when the client sends
// no `options` (the reference never does, web/streamlit_app.py:91-95):
//   repeat penalty over the last `repeat_last_n` tokens -> top-k -> top-p (on the softmax of the kept
//   logits) -> temperature -> inverse-CDF draw with a counter-based RNG (splitmix64(seed + step*phi)).
// Ties are broken towards the lower token id; every candidate gets a unique 64-bit key
// (order-preserving logit bits << 32 | ~id), so "the k largest" is well defined.
//
// Selection: three histogram passes (11+11+10 bits, shared-memory atomics) find the k-th largest logit
// exactly, a fourth pass gathers the candidates, a rank sort (warp shuffles are not needed at k <= 256)
// orders them, and one thread does the short sequential tail (<= 256 exps and adds) in the same order as
// the CPU oracle so that the drawn token id is identical.
#include "common.cuh"
#include "kernels.h"

namespace b200 {
namespace {

constexpr int kThreads = 1024;
constexpr int kMaxK = 256;      // == B200_MAX_TOP_K
constexpr int kMaxEq = 1024;    // candidates tied with the k-th value (beyond this they cannot all be listed)

DEVINL uint32_t ordered_bits(float v) {
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
DEVINL uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

DEVINL float from_ordered_bits(uint32_t o) { return __uint_as_float((o & 0x80000000u) ? (o ^ 0x80000000u) : ~o); }

struct Smem {
  uint32_t hist[2048];
  uint32_t seen[(1 << 18) / 32];  // token bitmap for the repeat penalty (vocab <= 262144)
  unsigned long long cand[kMaxK];
  unsigned long long sorted[kMaxK];
  uint32_t eq_idx[kMaxEq];
  uint32_t prefix, bin, above, n_cand, n_eq;
};

DEVINL float penalised(const float* logits, const uint32_t* seen, int i, float penalty) {
  float v = logits[i];
  if (seen[i >> 5] & (1u << (i & 31))) v = v > 0.f ? v / penalty : v * penalty;
  return v;
}

// Where the candidates come from: every logit of the vocabulary (one-kernel path) ...
struct SrcLogits {
  const float* logits;
  const uint32_t* seen;
  float penalty;
  DEVINL bool get(int i, uint32_t& o, uint32_t& id) const {
    o = ordered_bits(penalised(logits, seen, i, penalty));
    id = (uint32_t)i;
    return true;
  }
};
// ... or the per-slice top-k keys left by sample_part_kernel (two-kernel path): slice c, rank r at part[c*kMaxK + r],
// key = ordered penalised logit << 32 | ~id, 0 = empty slot
struct SrcKeys {
  const unsigned long long* part;
  int k;
  DEVINL bool get(int i, uint32_t& o, uint32_t& id) const {
    const unsigned long long key = part[(size_t)(i / k) * kMaxK + (i % k)];
    o = (uint32_t)(key >> 32);
    id = 0xffffffffu - (uint32_t)(key & 0xffffffffull);
    return key != 0ull;
  }
};

// Finds in hist[0..nbins) the highest bin b such that count(bins > b) < want <= count(bins >= b), all threads
// together: thread t owns bins 2t, 2t+1 (nbins <= 2 * kThreads); a block-wide suffix sum gives every thread the
// count above its bins, and the one thread whose bin straddles `want` publishes it.  (One thread walking the
// 2048 bins from the top cost ~8 us per pass, three passes per token.)
DEVINL void pick_bin(Smem& s, int nbins, uint32_t want) {
  __shared__ uint32_t warp_tot[kThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t h0 = 2 * tid < nbins ? s.hist[2 * tid] : 0u, h1 = 2 * tid + 1 < nbins ? s.hist[2 * tid + 1] : 0u;
  // suffix sum over threads: sfx = sum of (h0 + h1) of all threads with a higher index
  uint32_t v = h0 + h1, incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_down_sync(0xffffffffu, incl, o);
    if (lane + o < 32) incl += up;
  }
  if (lane == 0) warp_tot[warp] = incl;  // total of this warp
  __syncthreads();
  uint32_t higher_warps = 0;
  for (int w = warp + 1; w < kThreads / 32; ++w) higher_warps += warp_tot[w];
  const uint32_t above1 = higher_warps + (incl - v);  // everything above bin 2t+1
  const uint32_t above0 = above1 + h1;                // everything above bin 2t
  // bin 0 takes whatever is left (the serial version stopped at b == 0 unconditionally)
  if (2 * tid + 1 < nbins && above1 < want && want <= above1 + h1) { s.bin = (uint32_t)(2 * tid + 1); s.above += above1; }
  else if (2 * tid < nbins && above0 < want && (want <= above0 + h0 || tid == 0)) { s.bin = (uint32_t)(2 * tid); s.above += above0; }
  __syncthreads();
}

// Exact top-k of the source's candidates (then, in sample_core, top-p / temperature / draw).  Needs s.prefix =
// s.above = s.n_cand = s.n_eq = 0 on entry (one CTA of kThreads threads).
// s.sorted[0..k) <- the k largest keys of the source, descending (k <= number of valid candidates)
template <class Src>
DEVINL void select_topk(Smem& s, const Src& src, int n, int k) {
  const int tid = threadIdx.x;
  // ---- exact k-th largest ordered value: 11 + 11 + 10 bits
  const int shifts[3] = {21, 10, 0}, widths[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int nb = 1 << widths[pass];
    for (int i = tid; i < nb; i += kThreads) s.hist[i] = 0;
    __syncthreads();
    const uint32_t pre = s.prefix;
    const int hi_shift = shifts[pass] + widths[pass];
    for (int i = tid; i < n; i += kThreads) {
      uint32_t o, id;
      if (!src.get(i, o, id)) continue;
      if (pass == 0 || (o >> hi_shift) == (pre >> hi_shift)) atomicAdd(&s.hist[(o >> shifts[pass]) & (nb - 1)], 1u);
    }
    __syncthreads();
    pick_bin(s, nb, (uint32_t)k - s.above);
    __syncthreads();
    if (tid == 0) s.prefix |= s.bin << shifts[pass];
    __syncthreads();
  }
  const uint32_t kth = s.prefix;       // ordered bits of the k-th largest logit
  const uint32_t above = s.above;      // how many are strictly larger (< k)
  // ---- gather: everything above the threshold, plus the lowest-id ties at the threshold
  for (int i = tid; i < n; i += kThreads) {
    uint32_t o, id;
    if (!src.get(i, o, id)) continue;
    if (o > kth) {
      const uint32_t slot = atomicAdd(&s.n_cand, 1u);
      s.cand[slot] = ((unsigned long long)o << 32) | (0xffffffffu - id);
    } else if (o == kth) {
      const uint32_t slot = atomicAdd(&s.n_eq, 1u);
      if (slot < kMaxEq) s.eq_idx[slot] = id;
    }
  }
  __syncthreads();
  const int need = k - (int)above, n_eq = (int)min(s.n_eq, (uint32_t)kMaxEq);
  for (int e = tid; e < n_eq; e += kThreads) {  // rank the ties by id; the `need` smallest ids join
    const uint32_t id = s.eq_idx[e];
    int rank = 0;
    for (int j = 0; j < n_eq; ++j) rank += s.eq_idx[j] < id;
    if (rank < need) s.cand[above + rank] = ((unsigned long long)kth << 32) | (0xffffffffu - id);
  }
  __syncthreads();
  for (int e = tid; e < k; e += kThreads) {  // rank sort, descending key
    const unsigned long long key = s.cand[e];
    int rank = 0;
    for (int j = 0; j < k; ++j) rank += s.cand[j] > key;
    s.sorted[rank] = key;
  }
  __syncthreads();
}

template <class Src>
DEVINL void sample_core(Smem& s, const Src& src, int n, int k, const SampleParams& p, uint64_t step, int32_t* out) {
  select_topk(s, src, n, k);
  if (threadIdx.x != 0) return;
  // ---- sequential tail, same operation order as the oracle (orc_sample)
  // the key holds the penalised logit exactly (ordered_bits is a bijection on the float bits)
  auto logit_of = [&](int i) { return from_ordered_bits((uint32_t)(s.sorted[i] >> 32)); };
  if (p.temperature <= 0.f) {  // greedy over the penalised logits (k == 1)
    *out = (int32_t)(0xffffffffu - (uint32_t)(s.sorted[0] & 0xffffffffull));
    return;
  }
  const float mx = logit_of(0);
  float sum = 0.f;
  for (int i = 0; i < k; ++i) sum += expf(logit_of(i) - mx);
  int keep = k;
  if (p.top_p < 1.0f) {
    float cum = 0.f;
    for (int i = 0; i < k; ++i) {
      cum += expf(logit_of(i) - mx) / sum;
      if (cum >= p.top_p) { keep = i + 1; break; }
    }
  }
  float sum2 = 0.f;
  for (int i = 0; i < keep; ++i) sum2 += expf((logit_of(i) - mx) / p.temperature);
  const uint64_t r = splitmix64(p.seed + step * 0x9E3779B97F4A7C15ull);
  const float uu = (float)(r >> 40) * 5.9604644775390625e-8f * sum2;
  float cum = 0.f;
  int pick = keep - 1;
  for (int i = 0; i < keep; ++i) {
    cum += expf((logit_of(i) - mx) / p.temperature);
    if (uu < cum) { pick = i; break; }
  }
  *out = (int32_t)(0xffffffffu - (uint32_t)(s.sorted[pick] & 0xffffffffull));
}

__global__ void __launch_bounds__(kThreads) sample_kernel(const SampleArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  pdl_wait();
  const StepState* st = a.st;
  if (st != nullptr && (st->greedy || st->pos + 1 < st->n_prompt)) return;
  const int tid = threadIdx.x, V = a.vocab;
  const SampleParams p = *a.params;
  const int n_hist = a.n_hist >= 0 ? a.n_hist : st->pos + 1;
  const uint64_t step = a.n_hist >= 0 ? a.step : (uint64_t)st->n_out;

  for (int i = tid; i < (V + 31) / 32; i += kThreads) s.seen[i] = 0;
  if (tid == 0) { s.prefix = 0; s.above = 0; s.n_cand = 0; s.n_eq = 0; }
  __syncthreads();
  if (p.repeat_penalty != 1.0f && p.repeat_last_n != 0) {
    const int from = p.repeat_last_n < 0 ? 0 : max(0, n_hist - p.repeat_last_n);
    for (int i = from + tid; i < n_hist; i += kThreads) {
      const int t = a.history[i];
      if (t >= 0 && t < V) atomicOr(&s.seen[t >> 5], 1u << (t & 31));
    }
  }
  __syncthreads();

  if (p.temperature <= 0.f) {  // greedy over the penalised logits
    unsigned long long best = 0;
    for (int i = tid; i < V; i += kThreads) {
      const unsigned long long k = ((unsigned long long)ordered_bits(penalised(a.logits, s.seen, i, p.repeat_penalty)) << 32) | (0xffffffffu - (uint32_t)i);
      best = k > best ? k : best;
    }
    unsigned long long* red = s.cand;
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
      best = other > best ? other : best;
    }
    if ((tid & 31) == 0) red[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < kThreads / 32; ++w) best = red[w] > best ? red[w] : best;
      *a.out = (int32_t)(0xffffffffu - (uint32_t)(best & 0xffffffffull));
    }
    return;
  }

  int k = p.top_k <= 0 ? kMaxK : min(p.top_k, kMaxK);
  k = min(k, V);
  const SrcLogits src = {a.logits, s.seen, p.repeat_penalty};
  sample_core(s, src, V, k, p, step, a.out);
}

// ---- two-kernel path: the single CTA above walks the 128 K logits four times (~170 us per sampled token).  Here
// every 1024-logit slice of the vocabulary is ranked by its own CTA (the global top-k is a subset of the union of the
// slices' top-k), and one CTA then selects among <= slices * k keys.  Same keys, same order, same tail: same token.
constexpr int kSlice = 1024;

// a slice's candidates: the 1024 keys of the slice in shared memory (0 = past the end of the vocabulary)
struct SrcSlice {
  const unsigned long long* keys;
  DEVINL bool get(int i, uint32_t& o, uint32_t& id) const {
    const unsigned long long key = keys[i];
    o = (uint32_t)(key >> 32);
    id = 0xffffffffu - (uint32_t)(key & 0xffffffffull);
    return key != 0ull;
  }
};
static_assert(kSlice == kThreads, "one logit per thread");

__global__ void __launch_bounds__(kSlice) sample_part_kernel(const SampleArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(s.seen);  // the token bitmap's space: [kSlice] keys + 32 words
  uint32_t* seen = s.seen + 2 * kSlice;
  pdl_wait();
  pdl_launch_dependents();
  const StepState* st = a.st;
  if (st != nullptr && (st->greedy || st->pos + 1 < st->n_prompt)) return;
  const int tid = threadIdx.x, V = a.vocab, base = blockIdx.x * kSlice;
  const SampleParams p = *a.params;
  const int n_hist = a.n_hist >= 0 ? a.n_hist : st->pos + 1;
  int k = p.temperature <= 0.f ? 1 : (p.top_k <= 0 ? kMaxK : min(p.top_k, kMaxK));
  k = min(k, min(V - base, kSlice));  // a short last slice has fewer candidates than k
  if (tid < kSlice / 32) seen[tid] = 0;
  if (tid < kMaxK) a.part[(size_t)blockIdx.x * kMaxK + tid] = 0ull;  // empty slots
  if (tid == 0) { s.prefix = 0; s.above = 0; s.n_cand = 0; s.n_eq = 0; }
  __syncthreads();
  if (p.repeat_penalty != 1.0f && p.repeat_last_n != 0) {
    const int from = p.repeat_last_n < 0 ? 0 : max(0, n_hist - p.repeat_last_n);
    for (int i = from + tid; i < n_hist; i += kSlice) {
      const int t = a.history[i] - base;
      if (t >= 0 && t < kSlice) atomicOr(&seen[t >> 5], 1u << (t & 31));
    }
  }
  __syncthreads();
  const int id = base + tid;
  unsigned long long key = 0ull;
  if (id < V) {
    float v = a.logits[id];
    if (seen[tid >> 5] & (1u << (tid & 31))) v = v > 0.f ? v / p.repeat_penalty : v * p.repeat_penalty;
    key = ((unsigned long long)ordered_bits(v) << 32) | (0xffffffffu - (uint32_t)id);
  }
  keys[tid] = key;
  __syncthreads();
  const SrcSlice src = {keys};
  select_topk(s, src, kSlice, k);
  if (tid < k) a.part[(size_t)blockIdx.x * kMaxK + tid] = s.sorted[tid];
}

__global__ void __launch_bounds__(kThreads) sample_final_kernel(const SampleArgs a) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  pdl_wait();
  pdl_launch_dependents();
  const StepState* st = a.st;
  if (st != nullptr && (st->greedy || st->pos + 1 < st->n_prompt)) return;
  const SampleParams p = *a.params;
  const uint64_t step = a.n_hist >= 0 ? a.step : (uint64_t)st->n_out;
  int k = p.temperature <= 0.f ? 1 : (p.top_k <= 0 ? kMaxK : min(p.top_k, kMaxK));
  k = min(k, a.vocab);
  if (threadIdx.x == 0) { s.prefix = 0; s.above = 0; s.n_cand = 0; s.n_eq = 0; }
  __syncthreads();
  const int slices = (a.vocab + kSlice - 1) / kSlice;
  const SrcKeys src = {a.part, k};
  sample_core(s, src, slices * k, k, p, step, a.out);
}

}  // namespace

size_t sample_part_bytes(int vocab) { return (size_t)((vocab + kSlice - 1) / kSlice) * kMaxK * sizeof(unsigned long long); }

cudaError_t launch_sample(const SampleArgs& a, cudaStream_t s, bool pdl) {
  if (a.vocab <= 0 || a.vocab > (1 << 18)) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = sizeof(Smem);
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  if (a.part != nullptr && a.vocab >= 4 * kSlice) {  // two kernels: per-slice ranking, then the selection and the draw
    cudaLaunchConfig_t c1 = cfg;
    c1.gridDim = dim3((a.vocab + kSlice - 1) / kSlice);
    c1.blockDim = dim3(kSlice);
    e = cudaFuncSetAttribute(sample_part_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return e;
    count_launch();
    e = cudaLaunchKernelEx(&c1, sample_part_kernel, a);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(sample_final_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return e;
    return count_launch(), cudaLaunchKernelEx(&cfg, sample_final_kernel, a);
  }
  return count_launch(), cudaLaunchKernelEx(&cfg, sample_kernel, a);
}

}  // namespace b200
