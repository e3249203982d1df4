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

// Finds in hist[0..nbins) the highest bin b such that count(bins > b) < want <= count(bins >= b).
DEVINL void pick_bin(Smem& s, int nbins, uint32_t want) {
  if (threadIdx.x == 0) {
    uint32_t above = 0;
    int b = nbins - 1;
    for (; b > 0; --b) {
      if (above + s.hist[b] >= want) break;
      above += s.hist[b];
    }
    s.bin = (uint32_t)b;
    s.above += above;
  }
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
  // ---- exact k-th largest ordered value: 11 + 11 + 10 bits
  const int shifts[3] = {21, 10, 0}, widths[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int nb = 1 << widths[pass];
    for (int i = tid; i < nb; i += kThreads) s.hist[i] = 0;
    __syncthreads();
    const uint32_t pre = s.prefix;
    const int hi_shift = shifts[pass] + widths[pass];
    for (int i = tid; i < V; i += kThreads) {
      const uint32_t o = ordered_bits(penalised(a.logits, s.seen, i, p.repeat_penalty));
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
  for (int i = tid; i < V; i += kThreads) {
    const uint32_t o = ordered_bits(penalised(a.logits, s.seen, i, p.repeat_penalty));
    if (o > kth) {
      const uint32_t slot = atomicAdd(&s.n_cand, 1u);
      s.cand[slot] = ((unsigned long long)o << 32) | (0xffffffffu - (uint32_t)i);
    } else if (o == kth) {
      const uint32_t slot = atomicAdd(&s.n_eq, 1u);
      if (slot < kMaxEq) s.eq_idx[slot] = (uint32_t)i;
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
  if (tid != 0) return;
  // ---- sequential tail, same operation order as the oracle (orc_sample)
  auto logit_of = [&](int i) { return penalised(a.logits, s.seen, (int)(0xffffffffu - (uint32_t)(s.sorted[i] & 0xffffffffull)), p.repeat_penalty); };
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
  *a.out = (int32_t)(0xffffffffu - (uint32_t)(s.sorted[pick] & 0xffffffffull));
}

}  // namespace

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
  return cudaLaunchKernelEx(&cfg, sample_kernel, a);
}

}  // namespace b200
