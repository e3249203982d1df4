// gemv.cu — batch-1 decode mat-vec kernels for sm_100a: y = W.x with W streamed once from HBM.
//
// One kernel template, three weight formats (bf16, Q4_K, Q6_K), fused prologues (RMSNorm, int16 activation
// quantisation) and fused epilogues (residual add, SwiGLU, RoPE + KV-cache append, logits + argmax).
//
// Data movement.  Every warp owns a private ring of shared-memory stages that the TMA unit fills with
// 1-D bulk copies (cp.async.bulk ... mbarrier::complete_tx, SASS UBLKCP) tagged L2 evict_first: weights are
// touched once per token and must not displace activations / KV in the 126 MB L2.  A stage holds R
// consecutive rows x CHUNK columns; the warp multiplies it against the activation vector that the
// prologue left in shared memory, then lane 0 refills the stage.  No block-level barrier is on the
// streaming path.  Work units (R rows, full K) are dealt round-robin to all warps of a grid of one CTA per
// SM, so consecutive warps stream consecutive rows.
//
// PDL.  Weights do not depend on the previous kernel, so a CTA fills its rings first and only then
// executes griddepcontrol.wait; under programmatic dependent launch that prefetch overlaps the previous
// kernel's tail and the launch gap.
//
// Numerics (DESIGN.md "Numerics contract", mirrored by oracle/oracle.cpp mode 0): fp32 activations and
// accumulation; bf16 weights used exactly; Q4_K/Q6_K weights multiply activations quantised to int16 per
// 32-element block (dp2a: int16 x int8) with exact integer sub-block sums.
#include <algorithm>
#include <cstdlib>

#include "gemv_formats.cuh"

namespace b200 {

namespace {

// ------------------------------------------------------------------ the kernel
// cta / n_ctas: this CTA's index among the CTAs working on matrix `a` (the whole grid, or one side of a dual launch)
template <class F, int PRO, int EPI>
DEVINL void gemv_body(const GemvArgs& a, int cta, int n_ctas) {
  extern __shared__ __align__(128) uint8_t smem[];
  // warp index through a shuffle so the compiler knows it is warp-uniform: the TMA issue below then takes
  // its addresses from uniform registers instead of a per-lane "waterfall" loop around every UBLKCP
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const size_t xb = (F::x_bytes(a.K) + 127) & ~(size_t)127;
  uint8_t* xs = smem;
  uint8_t* ring = smem + xb + (size_t)warp * F::STAGES * F::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + xb + (size_t)F::NW * F::STAGES * F::STAGE_BYTES) + warp * F::STAGES;
  float* red = reinterpret_cast<float*>(smem + xb + (size_t)F::NW * F::STAGES * F::STAGE_BYTES + F::NW * F::STAGES * 8);

  // unit u (R rows, full K) -> global warp u % twarps; consecutive units land on consecutive CTAs, so a
  // matrix with fewer units than warps still spreads over every SM
  const int n_units = a.N / F::R;
  const int chunks = a.K / F::CHUNK;
  const int gwarp = warp * n_ctas + cta, twarps = n_ctas * F::NW;
  const int my_units = gwarp < n_units ? (n_units - gwarp + twarps - 1) / twarps : 0;
  const int n_items = my_units * chunks;
  const uint32_t ring_s = smem_u32(ring), bars_s = smem_u32(bars);
  const uint64_t pol = l2_policy_evict_first();

  const bool tracer = a.trace != nullptr && cta == 0 && threadIdx.x == 0;
  if (tracer) a.trace[0] = global_timer_ns();
  if (EPI == EPI_LOGITS) {  // no prefetch before the dependency: the kernel may be skipped entirely
    pdl_wait();
    if (a.st != nullptr && a.st->pos + 1 < a.st->n_prompt) return;
  }
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < F::STAGES; ++s) mbar_init(bars_s + 8 * s, 1);
    mbar_fence_init();
  }
  __syncwarp();

  // producer cursor: kept by every lane (warp-uniform), only the elected lane touches the TMA unit
  int p_item = 0, p_unit = gwarp, p_chunk = 0, p_stage = 0;
  auto produce = [&]() {
    if (lane == 0) {
      const uint32_t bar = bars_s + 8 * p_stage;
      mbar_expect_tx(bar, F::STAGE_BYTES);
      F::issue(ring_s + p_stage * F::STAGE_BYTES, bar, a, p_unit, p_chunk, pol);
    }
    ++p_item;
    if (++p_chunk == chunks) {
      p_chunk = 0;
      p_unit += twarps;
    }
    if (++p_stage == F::STAGES) p_stage = 0;
  };
  // Before the dependency: the whole ring.  (Tried, round 2: one stage before the wait and the rest right after the
  // prologue has issued its activation loads, in case those loads queue behind the ring at the SM's memory port —
  // no gain for Q4_K, and bf16 ffn_down's prologue went from 1.1 to 2.5 us.)
#pragma unroll
  for (int s = 0; s < F::STAGES; ++s)
    if (p_item < n_items) produce();
  auto fill_rest = [] {};
  // RMS-norm weights do not depend on the previous kernel either: this thread's share (the element -> thread mapping of
  // the single-fetch prologues, K <= 8192) goes to registers now, so the prologue waits for the activations only
  constexpr int kPU = ProU<F::NW>::value;
  float4 nw[kPU];
  const uint64_t keep = l2_policy_evict_last();
#pragma unroll
  for (int u = 0; u < kPU; ++u) {
    const int k = 4 * ((int)threadIdx.x + u * F::NW * 32);
    nw[u] = (PRO == PRO_RMSNORM && k < a.K && a.K <= F::NW * 32 * 4 * kPU) ? ldg_keep(reinterpret_cast<const float4*>(a.norm_w + k), keep)
                                                                          : make_float4(1.f, 1.f, 1.f, 1.f);
  }
  if (EPI != EPI_LOGITS) {
    pdl_launch_dependents();
    pdl_wait();
  }
  if (tracer) a.trace[1] = global_timer_ns();
  F::template prologue<PRO>(xs, a, red, warp, lane, nw, fill_rest);
  if (tracer) a.trace[2] = global_timer_ns();
  if (EPI == EPI_LOGITS) pdl_launch_dependents();

  int c_stage = 0;
  uint32_t c_parity = 0;
  float bestv = -INFINITY;
  uint32_t besti = 0;
  for (int u = 0; u < my_units; ++u) {
    float acc[F::R], res[F::R];
#pragma unroll
    for (int r = 0; r < F::R; ++r) acc[r] = 0.f;
    if (EPI == EPI_RESID && lane == 0) {  // the residual these rows are added to: fetched now, not after the last chunk
#pragma unroll
      for (int r = 0; r < F::R; ++r) res[r] = __ldcg(a.y + (size_t)(gwarp + u * twarps) * F::R + r);
    }
    for (int c = 0; c < chunks; ++c) {
      mbar_wait(bars_s + 8 * c_stage, c_parity);
      F::consume(ring + (size_t)c_stage * F::STAGE_BYTES, xs, c, a, lane, acc);
      __syncwarp();
      if (p_item < n_items) {
        if (lane == 0) fence_proxy_async();
        produce();
      }
      if (++c_stage == F::STAGES) {
        c_stage = 0;
        c_parity ^= 1;
      }
    }
#pragma unroll
    for (int r = 0; r < F::R; ++r) acc[r] = warp_sum(acc[r]);
    if (EPI == EPI_RESID) {
#pragma unroll
      for (int r = 0; r < F::R; ++r) acc[r] = res[r] + acc[r];
    }
    epilogue<EPI, F::R>(a, gwarp + u * twarps, lane, acc, bestv, besti);
  }
  if (EPI == EPI_LOGITS && lane == 0 && my_units > 0) atomicMax(a.argmax_key, argmax_key(bestv, besti));
  if (a.trace != nullptr && threadIdx.x == 0) atomicMax(a.trace + 3, global_timer_ns());
  if (a.next_bytes != 0 && lane == 0) {  // this warp's slice of the next kernel's weights -> L2
    const unsigned long long per = ((a.next_bytes / (unsigned long long)twarps) + 127ull) & ~127ull;
    unsigned long long off = per * (unsigned long long)gwarp;
    const unsigned long long end = off + per < a.next_bytes ? off + per : a.next_bytes;
    for (; off < end; off += 8192ull) l2_prefetch(a.next_W + off, (uint32_t)((end - off < 8192ull ? end - off : 8192ull) & ~15ull));
  }
}

template <class F, int PRO, int EPI>
__global__ void __launch_bounds__(F::NW * 32, F::OCC) gemv_kernel(const GemvArgs a) {
  gemv_body<F, PRO, EPI>(a, (int)blockIdx.x, (int)gridDim.x);
}

// Two matrices that read the same input vector in ONE launch: the first a_ctas CTAs stream `a` in format FA, the
// rest stream `b` in format FB (each CTA runs one body, so the two activation layouts never meet).  Used for the
// q,k | v projections of Q4_K_M layers whose V is stored in Q6_K: one launch less per such layer (~7 us each).
template <class FA, class FB, int PRO, int EPI>
__global__ void __launch_bounds__(FA::NW * 32, FA::OCC) gemv_dual_kernel(const GemvArgs a, const GemvArgs b, int a_ctas) {
  if ((int)blockIdx.x < a_ctas) gemv_body<FA, PRO, EPI>(a, (int)blockIdx.x, a_ctas);
  else gemv_body<FB, PRO, EPI>(b, (int)blockIdx.x - a_ctas, (int)gridDim.x - a_ctas);
}

template <class F>
size_t smem_bytes(int K) {
  const size_t xb = (F::x_bytes(K) + 127) & ~(size_t)127;
  return xb + (size_t)F::NW * F::STAGES * F::STAGE_BYTES + F::NW * F::STAGES * 8 + F::NW * 4 + 32;
}

// Dynamic shared memory a CTA of this shape may use: the sm_100 opt-in maximum, or half an SM (minus the 1 KB the
// system keeps per CTA) when two CTAs have to fit.
template <class F>
constexpr int max_smem() { return F::OCC >= 2 ? (228 * 1024) / 2 - 1024 : 227 * 1024; }

template <class F>
bool fits(int K) { return smem_bytes<F>(K) <= (size_t)max_smem<F>(); }

template <class F, int PRO, int EPI>
cudaError_t launch_t(const GemvArgs& a, cudaStream_t s, bool pdl, int sm_count) {
  auto kern = gemv_kernel<F, PRO, EPI>;
  const size_t smem = smem_bytes<F>(a.K);
  // Per-device attribute, cheap to set: no caching, so one host thread may drive several GPUs.  Always the
  // maximum for this shape rather than this launch's size: the same kernel serves matrices with different K,
  // and a tool that re-launches a captured node (ncu replay) reads the attribute as it stands at that time.
  if (smem > (size_t)max_smem<F>()) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem<F>());
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  const int n_units = a.N / F::R;
  int grid = n_units < sm_count ? n_units : sm_count;  // units are dealt CTA-first (see gemv_kernel)
  if (grid < 1) grid = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(F::NW * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return count_launch(), cudaLaunchKernelEx(&cfg, kern, a);
}

template <class F>
cudaError_t launch_f(int pro, int epi, const GemvArgs& a, cudaStream_t s, bool pdl, int sm) {
#define CASE(P, E) \
  if (pro == P && epi == E) return launch_t<F, P, E>(a, s, pdl, sm);
  CASE(PRO_COPY, EPI_STORE)
  CASE(PRO_RMSNORM, EPI_STORE)
  CASE(PRO_COPY, EPI_RESID)
  CASE(PRO_RMSNORM, EPI_SWIGLU)
  CASE(PRO_RMSNORM, EPI_QKV)
  CASE(PRO_RMSNORM, EPI_LOGITS)
  if constexpr (F::OCC == 1) {  // tensor parallelism over peer memory: the fat shape only
    CASE(PRO_COPY, EPI_TP_PUSH)
    CASE(PRO_TP_RMSNORM, EPI_SWIGLU)
    CASE(PRO_TP_RMSNORM, EPI_QKV)
    CASE(PRO_TP_RMSNORM, EPI_LOGITS)
  }
#undef CASE
  return cudaErrorInvalidValue;
}

template <class FA, class FB, int PRO, int EPI>
cudaError_t launch_dual_t(const GemvArgs& a, const GemvArgs& b, cudaStream_t s, bool pdl, int sm_count) {
  auto kern = gemv_dual_kernel<FA, FB, PRO, EPI>;
  static_assert(FA::NW == FB::NW && FA::OCC == FB::OCC, "both sides of a dual launch share the CTA shape");
  const size_t smem = std::max(smem_bytes<FA>(a.K), smem_bytes<FB>(b.K));
  if (smem > (size_t)max_smem<FA>() || sm_count < 2) return cudaErrorInvalidValue;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem<FA>());
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return e;
  // CTAs in proportion to the bytes each side streams
  const double ba = (double)a.N * (double)FA::row_bytes(a.K), bb = (double)b.N * (double)FB::row_bytes(b.K);
  int a_ctas = (int)(sm_count * ba / (ba + bb) + 0.5);
  a_ctas = std::min(std::max(a_ctas, 1), sm_count - 1);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sm_count);
  cfg.blockDim = dim3(FA::NW * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return count_launch(), cudaLaunchKernelEx(&cfg, kern, a, b, a_ctas);
}

// ------------------------------------------------------------------ launch shapes
// One 16-warp CTA per SM owning up to the whole shared memory (deep rings).  Measured alternative (round 2, B200):
// 8-warp CTAs limited to half an SM (GemvCfg<8, 2..3, 2>), so that under programmatic dependent launch the next
// kernel's CTAs sit resident and prefetched beside the running ones — slower for every matrix class (bf16 3.15 vs
// 2.82 ms/token, Q4_K_M 2.13 vs 1.51): half the warps per SM hide less latency than the overlap wins back.
using S0r4 = GemvCfg<16, 2, 1>;  // 4-row units
using S0r2 = GemvCfg<16, 3, 1>;  // 2-row units
using S0q6 = GemvCfg<16, 2, 1>;

}  // namespace

// 4 rows per unit amortise the activation loads best and keep the most bytes in flight per warp; 2-row units give a
// matrix with few rows a unit for every warp of every SM.  Which wins was measured per format (B200_GEMV_WIDE_PCT =
// 4-row units when they cover at least that percentage of the 16 x SMs warps): 16-bit weights 60 % (8B qkv, 6144 rows,
// 2.685 -> 2.672 ms per token; o / down, 4096 rows = 43 %: 2-row units, 4-row ones cost 2 %), k-quants 40 % (4-row
// units everywhere: 1.438 -> 1.415 ms).  Where it matters most is the 70B model at TP-8, whose o and down keep all 8192
// rows with K cut to 1024 / 3584: 2-row units meant two rounds of short units per warp (o 9.0 us, down 10.9 us per
// layer), 4-row units one round (2.7 / 5.9 us): 189 -> 205 tok/s.
bool wide_units(int wtype, int64_t n_rows, int sm_count) {
  static const int env_pct = getenv("B200_GEMV_WIDE_PCT") ? atoi(getenv("B200_GEMV_WIDE_PCT")) : 0;
  const int pct = env_pct > 0 ? env_pct : ((wtype == T_Q4_K || wtype == T_Q6_K) ? 40 : 60);
  return (n_rows / 4) * 100 >= (int64_t)sm_count * 16 * pct;
}

// q,k rows in Q4_K and v rows in Q6_K (the Q4_K_M recipe's "more bits" layers), RMS-norm prologue, RoPE / KV-append
// epilogue, one launch.  Other combinations: cudaErrorNotSupported (the caller launches them one after the other).
cudaError_t launch_gemv_dual_qkv(int type_a, const GemvArgs& a, int type_b, const GemvArgs& b, cudaStream_t s, bool pdl, int sm_count) {
  static const bool off = getenv("B200_NO_DUAL_QKV") != nullptr;
  if (type_a != T_Q4_K || type_b != T_Q6_K || off) return cudaErrorNotSupported;
  if (!gemv_shape_ok(type_a, a.N, a.K, EPI_QKV) || !gemv_shape_ok(type_b, b.N, b.K, EPI_QKV)) return cudaErrorInvalidValue;
  const bool wide = wide_units(type_a, a.N, sm_count);
  if (wide) return launch_dual_t<FmtQ4K<4, S0r4>, FmtQ6K<S0q6>, PRO_RMSNORM, EPI_QKV>(a, b, s, pdl, sm_count);
  return launch_dual_t<FmtQ4K<2, S0r2>, FmtQ6K<S0q6>, PRO_RMSNORM, EPI_QKV>(a, b, s, pdl, sm_count);
}

size_t gemv_row_bytes(int wtype, int64_t K) {
  switch (wtype) {
    case T_BF16: return (size_t)K * 2;
    case T_F16: return (size_t)K * 2;
    case T_Q4_K: return (size_t)(K / QK_K) * Q4K_BYTES;
    case T_Q6_K: return (size_t)(K / QK_K) * Q6K_BYTES;
  }
  return 0;
}

bool gemv_shape_ok(int wtype, int64_t N, int64_t K, int epi) {
  if (N <= 0 || K <= 0 || K > 32768) return false;
  switch (wtype) {
    case T_BF16: return K % 256 == 0 && N % 4 == 0;
    case T_F16: return K % 256 == 0 && N % 4 == 0;
    case T_Q4_K: return K % (8 * QK_K) == 0 && N % 4 == 0;
    case T_Q6_K: return K % (8 * QK_K) == 0 && N % 2 == 0;
  }
  (void)epi;
  return false;
}

namespace {
template <class W, class Nr>
cudaError_t launch_by_width(bool wide, int pro, int epi, const GemvArgs& a, cudaStream_t s, bool pdl, int sm) {
  return wide ? launch_f<W>(pro, epi, a, s, pdl, sm) : launch_f<Nr>(pro, epi, a, s, pdl, sm);
}
}  // namespace

cudaError_t launch_gemv(int wtype, int pro, int epi, const GemvArgs& a, cudaStream_t s, bool pdl, int sm_count) {
  if (!gemv_shape_ok(wtype, a.N, a.K, epi)) return cudaErrorInvalidValue;
  const bool wide = wide_units(wtype, a.N, sm_count);
  switch (wtype) {
    case T_BF16:
      if (a.K % 512 == 0) return launch_by_width<FmtBF16<512, 4, S0r4>, FmtBF16<512, 2, S0r2>>(wide, pro, epi, a, s, pdl, sm_count);
      return launch_by_width<FmtBF16<256, 4, S0r4>, FmtBF16<256, 2, S0r2>>(wide, pro, epi, a, s, pdl, sm_count);
    case T_F16:  // F16 GGUF files: same data path, the widening is a convert instead of a shift
      return launch_by_width<FmtBF16<256, 4, S0r4, true>, FmtBF16<256, 2, S0r2, true>>(wide, pro, epi, a, s, pdl, sm_count);
    case T_Q4_K: return launch_by_width<FmtQ4K<4, S0r4>, FmtQ4K<2, S0r2>>(wide, pro, epi, a, s, pdl, sm_count);
    case T_Q6_K: return launch_f<FmtQ6K<S0q6>>(pro, epi, a, s, pdl, sm_count);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b200
