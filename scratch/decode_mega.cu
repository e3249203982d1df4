// decode_mega.cu — one token of batch-1 decode in ONE persistent kernel, sm_100a.  EXPERIMENT, opt-in
// (B200_MEGA=1): correct and bit-identical to the default path, but measured slower — see "Result" below.
//
// Idea.  The per-projection kernels of gemv.cu stream their weights at ~96 % of the HBM peak while they
// stream, but a decode step is 160+ launches and each one pays ~4-6 us of fill/drain around that (launch,
// RMS-norm / quantisation prologue, first TMA round trip, the tail of the slowest warp).  Here the whole step
// is one cooperative launch of one CTA per SM; the projections and the attention are *phases* separated by
// grid barriers, and every warp's TMA ring keeps running across the barriers: when a warp has consumed its
// last stage of phase p it immediately starts fetching its first stages of the next mat-vec phase.
//
// The arithmetic is the very same code as gemv.cu / attn.cu (gemv_formats.cuh, attn_decode.cuh): same
// per-row accumulation order, same epilogues, so a step through this kernel is bit-identical to a step
// through the per-projection kernels (tests/test_gpu_mega.py asserts that).
//
// Ring geometry: 16 warps x 2 stages x 4608 B (the largest stage of any format: Q4_K, 4 rows x 8
// super-blocks); formats with smaller stages simply use part of a slot.  Work units are dealt to warps as in
// gemv.cu (unit u -> global warp u % total_warps, CTA-first).
//
// Every wait in this kernel is bounded (~2 s): on a time-out it raises MegaArgs::error and carries on, so a
// bug or a lost CTA produces a wrong token and an error code, never a hung GPU.
//
// Result (B200, Llama-3.1-8B, ctx 576; per-phase %globaltimer trace, B200_MEGA_TRACE=<file>):
//   bf16 3.79 ms/token vs 3.01 with the PDL-chained kernels; Q4_K_M 2.71 vs 1.73.  Per layer the phases cost
//   qkv 13.7 + attention 14.0 + o 14.6 + gate/up 43.9 + down 27.4 = 113.6 us against 93 us for the chained
//   kernels.  A grid barrier under load costs 2.2 us after the last arrival (1.3 us in isolation), every phase
//   still pays its 1.5-2.4 us prologue, and a phase ends with its slowest warp (entry times into the barrier
//   spread over 2-5 us), so the cross-phase prefetch (19 MB of ring) buys less than the barriers cost; the
//   programmatic-dependent-launch chain already overlaps the next kernel's weight prefetch with the previous
//   kernel's tail at a lower fixed cost.  Kept for the next round's experiments (fewer, fatter phases).
#include "attn_decode.cuh"
#include "gemv_formats.cuh"

namespace b200 {
namespace {

constexpr int kSlot = 4608;
constexpr int kStages = 2;
constexpr long long kSpinLimit = 4000000000ll;  // clock64 ticks

using F0 = FmtBF16<512, 4>;
using F1 = FmtBF16<1024, 2>;
using F2 = FmtQ4K<4>;
using F3 = FmtQ4K<2>;
using F4 = FmtQ6K;
static_assert(F0::STAGE_BYTES <= kSlot && F1::STAGE_BYTES <= kSlot && F2::STAGE_BYTES <= kSlot && F3::STAGE_BYTES <= kSlot &&
                  F4::STAGE_BYTES <= kSlot, "ring slot too small");

DEVINL unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Self-resetting grid barrier (arrival counter + generation).  Cooperative launch guarantees co-residency.
DEVINL void grid_barrier(const MegaArgs& m, int ph = -1) {
  __syncthreads();
  if (m.trace && ph >= 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    m.trace[(size_t)m.n_phases * 4 + (size_t)ph * 160 + blockIdx.x] = t;
  }
  if (threadIdx.x == 0) {
    const unsigned int gen = ld_acquire_gpu(m.bar_gen);
    __threadfence();
    if (atomicAdd(m.bar_count, 1u) == gridDim.x - 1) {
      *m.bar_count = 0u;
      __threadfence();
      atomicAdd(m.bar_gen, 1u);
    } else {
      const long long t0 = clock64();
      while (ld_acquire_gpu(m.bar_gen) == gen) {
        // once any wait of the step has timed out nobody waits again: the step drains in bounded time
        if (clock64() - t0 > kSpinLimit || ld_acquire_gpu(m.error) != 0u) {
          *m.error = 1u;
          break;
        }
      }
    }
  }
  __syncthreads();
}

DEVINL void mbar_wait_bounded(uint32_t bar, uint32_t parity, unsigned int* error, bool& dead) {
  if (dead || mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > kSpinLimit) {
      *error = 1u;
      dead = true;
      return;
    }
  }
}

DEVINL void stamp(const MegaArgs& m, int ph, int k) {
  if (m.trace && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    m.trace[ph * 4 + k] = t;
  }
}

template <class F>
DEVINL void issue_t(uint32_t dst, uint32_t bar, const GemvArgs& a, int unit, int c, uint64_t pol) {
  mbar_expect_tx(bar, F::STAGE_BYTES);
  F::issue(dst, bar, a, unit, c, pol);
}

// the producer's position in the step's item sequence (phase, unit, chunk); warp-uniform
struct Cursor {
  int ph, unit, chunk, n_units, chunks, fmt;
};

// moves c.ph forward to the next mat-vec phase in which this warp owns a unit (c.ph == n_ph: none left)
DEVINL void cursor_enter(Cursor& c, const MegaArgs& m, int n_ph, int gwarp) {
  while (c.ph < n_ph) {
    const MegaPhase* P = m.phases + c.ph;
    if (P->kind == 0 && gwarp < P->n_units) {
      c.unit = gwarp; c.chunk = 0; c.n_units = P->n_units; c.chunks = P->chunks; c.fmt = P->fmt;
      return;
    }
    ++c.ph;
  }
}

template <int G>
__global__ void __launch_bounds__(kWarps * 32, 1) decode_mega_kernel(const MegaArgs m) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  uint8_t* xs = smem;
  uint8_t* ring = smem + m.x_bytes + (size_t)warp * kStages * kSlot;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + m.x_bytes + (size_t)kWarps * kStages * kSlot) + warp * kStages;
  float* red = reinterpret_cast<float*>(smem + m.x_bytes + (size_t)kWarps * kStages * kSlot + kWarps * kStages * 8);
  const int gwarp = warp * (int)gridDim.x + (int)blockIdx.x, twarps = (int)gridDim.x * kWarps;
  const uint32_t ring_s = smem_u32(ring), bars_s = smem_u32(bars);
  const uint64_t pol = l2_policy_evict_first();
  // while the prompt is still being fed token by token the logits are not needed: drop the lm_head phase
  const int n_ph = (m.lm_head_phase >= 0 && m.st->pos + 1 < m.st->n_prompt) ? m.lm_head_phase : m.n_phases;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) mbar_init(bars_s + 8 * s, 1);
    mbar_fence_init();
  }
  __syncwarp();

  Cursor pc;
  pc.ph = 0; pc.unit = 0; pc.chunk = 0; pc.n_units = 0; pc.chunks = 1; pc.fmt = 0;
  cursor_enter(pc, m, n_ph, gwarp);
  int p_stage = 0;
  auto produce = [&]() {
    if (pc.ph >= n_ph) return;
    if (lane == 0) {
      const uint32_t bar = bars_s + 8 * p_stage, dst = ring_s + p_stage * kSlot;
      const GemvArgs& a = m.phases[pc.ph].g;
      switch (pc.fmt) {
        case MF_BF16_R4: issue_t<F0>(dst, bar, a, pc.unit, pc.chunk, pol); break;
        case MF_BF16_R2: issue_t<F1>(dst, bar, a, pc.unit, pc.chunk, pol); break;
        case MF_Q4K_R4: issue_t<F2>(dst, bar, a, pc.unit, pc.chunk, pol); break;
        case MF_Q4K_R2: issue_t<F3>(dst, bar, a, pc.unit, pc.chunk, pol); break;
        default: issue_t<F4>(dst, bar, a, pc.unit, pc.chunk, pol); break;
      }
    }
    if (++pc.chunk == pc.chunks) {
      pc.chunk = 0;
      pc.unit += twarps;
      if (pc.unit >= pc.n_units) {
        ++pc.ph;
        cursor_enter(pc, m, n_ph, gwarp);
      }
    }
    p_stage ^= 1;
  };
  produce();
  produce();

  int c_stage = 0;
  uint32_t c_parity = 0;
  bool dead = false;  // a TMA wait of this warp timed out: stop waiting
  for (int ph = 0; ph < n_ph; ++ph) {
    const MegaPhase* P = m.phases + ph;
    stamp(m, ph, 0);
    if (P->kind == 1) {  // ---- attention: (KV head, split) items on warps 0..3 of the first n_kv * S CTAs
      const int n_kv = P->t.n_kv_heads, S = m.attn_splits;
      if ((int)blockIdx.x < n_kv * S && warp < 4) {
        const AttnArgs t = P->t;
        attn_decode_item<G>(t, *reinterpret_cast<AttnSmem<G>*>(xs), (int)blockIdx.x % n_kv, (int)blockIdx.x / n_kv, S, 0, n_kv,
                            (int)threadIdx.x, 1);
      }
      stamp(m, ph, 2);
      grid_barrier(m, ph);
      stamp(m, ph, 3);
      continue;
    }
    // ---- mat-vec phase
    const GemvArgs a = P->g;
    const int fmt = P->fmt, pro = P->pro, epi = P->epi, n_units = P->n_units, chunks = P->chunks;
    if (fmt <= MF_BF16_R2) {
      if (pro == PRO_RMSNORM) F0::template prologue<PRO_RMSNORM>(xs, a, red, warp, lane);
      else F0::template prologue<PRO_COPY>(xs, a, red, warp, lane);
    } else if (fmt <= MF_Q4K_R2) {
      if (pro == PRO_RMSNORM) F2::template prologue<PRO_RMSNORM>(xs, a, red, warp, lane);
      else F2::template prologue<PRO_COPY>(xs, a, red, warp, lane);
    } else {
      if (pro == PRO_RMSNORM) F4::template prologue<PRO_RMSNORM>(xs, a, red, warp, lane);
      else F4::template prologue<PRO_COPY>(xs, a, red, warp, lane);
    }
    stamp(m, ph, 1);
    const bool r4 = (fmt == MF_BF16_R4 || fmt == MF_Q4K_R4);
    float bestv = -INFINITY;
    uint32_t besti = 0;
    bool any = false;
    for (int unit = gwarp; unit < n_units; unit += twarps) {
      any = true;
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = 0; c < chunks; ++c) {
        mbar_wait_bounded(bars_s + 8 * c_stage, c_parity, m.error, dead);
        const uint8_t* st = ring + (size_t)c_stage * kSlot;
        switch (fmt) {
          case MF_BF16_R4: F0::consume(st, xs, c, a, lane, acc); break;
          case MF_BF16_R2: F1::consume(st, xs, c, a, lane, acc); break;
          case MF_Q4K_R4: F2::consume(st, xs, c, a, lane, acc); break;
          case MF_Q4K_R2: F3::consume(st, xs, c, a, lane, acc); break;
          default: F4::consume(st, xs, c, a, lane, acc); break;
        }
        __syncwarp();
        if (lane == 0) fence_proxy_async();
        produce();  // refills the stage just consumed (with this phase's next item, or a later phase's first)
        if (++c_stage == kStages) {
          c_stage = 0;
          c_parity ^= 1;
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = warp_sum(acc[r]);
      if (r4) {
        switch (epi) {
          case EPI_QKV: epilogue<EPI_QKV, 4>(a, unit, lane, acc, bestv, besti); break;
          case EPI_RESID: epilogue<EPI_RESID, 4>(a, unit, lane, acc, bestv, besti); break;
          case EPI_SWIGLU: epilogue<EPI_SWIGLU, 4>(a, unit, lane, acc, bestv, besti); break;
          default: epilogue<EPI_LOGITS, 4>(a, unit, lane, acc, bestv, besti); break;
        }
      } else {
        switch (epi) {
          case EPI_QKV: epilogue<EPI_QKV, 2>(a, unit, lane, acc, bestv, besti); break;
          case EPI_RESID: epilogue<EPI_RESID, 2>(a, unit, lane, acc, bestv, besti); break;
          case EPI_SWIGLU: epilogue<EPI_SWIGLU, 2>(a, unit, lane, acc, bestv, besti); break;
          default: epilogue<EPI_LOGITS, 2>(a, unit, lane, acc, bestv, besti); break;
        }
      }
    }
    if (epi == EPI_LOGITS && lane == 0 && any) atomicMax(a.argmax_key, argmax_key(bestv, besti));
    stamp(m, ph, 2);
    if (m.trace && blockIdx.x < 2 && lane == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      m.trace[(size_t)m.n_phases * 4 + (size_t)ph * 160 + 148 + warp % 12] = t;  // (debug) a few warps of CTA 0/1
    }
    if (ph + 1 < n_ph) grid_barrier(m, ph);
    stamp(m, ph, 3);
  }
}

template <class F>
bool fits(int64_t N, int64_t K) {
  return K % F::CHUNK == 0 && N % F::R == 0;
}

}  // namespace

bool mega_plan_gemv(int wtype, int64_t N, int64_t K, int sm_count, MegaPhase* p) {
  if (N <= 0 || K <= 0 || K > 32768) return false;
  const bool wide = (N / 4) >= (int64_t)sm_count * kWarps && N % 4 == 0;
  int fmt, R, chunk;
  switch (wtype) {
    case T_BF16:
      if (wide && fits<F0>(N, K)) { fmt = MF_BF16_R4; R = F0::R; chunk = F0::CHUNK; }
      else if (fits<F1>(N, K)) { fmt = MF_BF16_R2; R = F1::R; chunk = F1::CHUNK; }
      else return false;
      break;
    case T_Q4_K:
      if (wide && fits<F2>(N, K)) { fmt = MF_Q4K_R4; R = F2::R; chunk = F2::CHUNK; }
      else if (fits<F3>(N, K)) { fmt = MF_Q4K_R2; R = F3::R; chunk = F3::CHUNK; }
      else return false;
      break;
    case T_Q6_K:
      if (!fits<F4>(N, K)) return false;
      fmt = MF_Q6K; R = F4::R; chunk = F4::CHUNK;
      break;
    default: return false;
  }
  p->kind = 0; p->fmt = fmt; p->n_units = (int)(N / R); p->chunks = (int)(K / chunk);
  return true;
}

size_t mega_x_bytes(const MegaPhase* phases, int n, int group) {
  size_t b = 0;
  switch (group) {
    case 1: b = sizeof(AttnSmem<1>); break;
    case 2: b = sizeof(AttnSmem<2>); break;
    case 4: b = sizeof(AttnSmem<4>); break;
    case 8: b = sizeof(AttnSmem<8>); break;
    default: return 0;
  }
  for (int i = 0; i < n; ++i) {
    if (phases[i].kind != 0) continue;
    const size_t x = phases[i].fmt <= MF_BF16_R2 ? F0::x_bytes(phases[i].g.K) : QActLayout::bytes(phases[i].g.K);
    if (x > b) b = x;
  }
  return (b + 127) & ~(size_t)127;
}

cudaError_t launch_decode_mega(const MegaArgs& m, int group, int sm_count, cudaStream_t s) {
  const size_t smem = (size_t)m.x_bytes + (size_t)kWarps * kStages * kSlot + kWarps * kStages * 8 + kWarps * 4 + 64;
  constexpr int kMaxOptinSmem = 227 * 1024;
  if (smem > (size_t)kMaxOptinSmem || m.x_bytes % 128) return cudaErrorInvalidValue;
  void (*kern)(MegaArgs) = nullptr;
  switch (group) {
    case 1: kern = decode_mega_kernel<1>; break;
    case 2: kern = decode_mega_kernel<2>; break;
    case 4: kern = decode_mega_kernel<4>; break;
    case 8: kern = decode_mega_kernel<8>; break;
    default: return cudaErrorInvalidValue;
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxOptinSmem);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sm_count);
  cfg.blockDim = dim3(kWarps * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the grid barriers cannot dead-lock
  at[0].val.cooperative = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, m);
}

}  // namespace b200
