// gemv_formats.cuh — the weight-format policies (bf16, Q4_K, Q6_K), fused prologues and epilogues of the
// batch-1 decode mat-vec.  Used by gemv.cu (one kernel per projection;
// scratch/decode_mega.cu, a measured-slower persistent-kernel experiment, used them too); see gemv.cu for the data-movement and numerics description.
//
// Activation vectors (x, the residual stream) are read with ld.global.cg: in the persistent kernel they are
// produced by other SMs earlier in the same launch, and an L1 line from a previous phase would be stale.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace b200 {

namespace {

// Launch shape of a GEMV kernel: NW consumer warps per CTA (each is its own TMA producer), ST ring stages per warp,
// OCC CTAs that must fit on an SM.  OCC = 1: one fat CTA per SM owns all of its shared memory.  OCC = 2: a CTA uses
// at most half of the SM, so under programmatic dependent launch the NEXT kernel's CTA becomes resident beside the
// running one and has its weight ring full (and its launch latency behind it) before the running kernel ends.
template <int NW_, int ST_, int OCC_>
struct GemvCfg {
  static constexpr int NW = NW_, ST = ST_, OCC = OCC_;
};

// ------------------------------------------------------------------ block helpers
template <int NW>
DEVINL float block_sum(float v, float* red, int warp, int lane) {
  v = warp_sum(v);
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < NW; ++i) t += red[i];
  __syncthreads();
  return t;
}

// float4 per thread kept in registers by the single-fetch prologues: covers K <= 8192 whatever the CTA width
template <int NW>
struct ProU {
  static constexpr int value = 8192 / (NW * 32 * 4) < 1 ? 1 : 8192 / (NW * 32 * 4);
};

template <int PRO, int NW>
DEVINL float rms_inv(const GemvArgs& a, float* red, int warp, int lane) {
  if (PRO != PRO_RMSNORM) return 1.f;  // (PRO_TP_RMSNORM is bf16-only and has its own prologue)
  float ss = 0.f;
  for (int k = threadIdx.x * 4; k < a.K; k += NW * 32 * 4) {
    const float4 v = __ldcg(reinterpret_cast<const float4*>(a.x + k));
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = block_sum<NW>(ss, red, warp, lane);
  return 1.0f / sqrtf(ss / (float)a.K + a.eps);
}

// ------------------------------------------------------------------ format policies
// Each policy states: R rows per unit, CHUNK columns per stage, the stage layout, how lane 0 issues the
// bulk copies of a stage, how the block prepares x in shared memory, and how a warp consumes a stage.

DEVINL uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
DEVINL void st_relaxed_sys_v2(void* p, float v, unsigned int tag) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(tag) : "memory");
}

DEVINL uint4 ld_relaxed_gpu_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
DEVINL void st_relaxed_gpu_v4(void* p, uint4 v) {
  asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Consumer half of the fused all-reduce (see TpLink), in two phases so that no CTA reads more than it must:
//  1. the residual stream is cut into slices of 64 elements, slice i belongs to CTA i (mod grid): 16 of its threads
//     poll the `size` partials of their four elements in this GPU's receive buffer (all loads in flight together)
//     until every pair carries this exchange's number, add them to x in rank order (fixed order: every rank computes
//     the same bits), store x_new and publish it as {value, seq} pairs in a local buffer;
//  2. every CTA polls that local buffer for all d elements (8 bytes each instead of 8 x size), and the RMS-normalised
//     x_new goes to shared memory in the GEMV's permuted layout.
// (One phase — every CTA polling all partials — read 512 KB per CTA at TP-8, d = 8192: 13.4 us of prologue.)
// A wait that exceeds ~2 s sets the error flag and carries on (no hang).
template <int NW>
DEVINL void tp_prologue(float* xs, const GemvArgs& a, float* red, int warp, int lane) {
  const TpLink* t = a.tp;
  const int buf = a.tp_index & 1, size = t->size, d = a.K;
  const unsigned int seq = *t->epoch * (unsigned int)t->per_step + (unsigned int)a.tp_index + 1u;
  const float* rv = t->recv[buf];
  float* xsum = t->xsum[buf];
  const long long t0 = clock64();
  if (threadIdx.x < 16) {
    for (int sl = blockIdx.x; sl < d / 64; sl += gridDim.x) {
      const int k = sl * 64 + threadIdx.x * 4;
      float4 v = __ldcg(reinterpret_cast<const float4*>(a.x + k));
      uint4 p0[kMaxTp], p1[kMaxTp];
      unsigned int pending = (1u << size) - 1u;
      while (pending) {
#pragma unroll
        for (int s = 0; s < kMaxTp; ++s)
          if ((pending >> s) & 1u) {
            const float* e = rv + ((size_t)s * d + k) * 2;  // four {value, seq} pairs
            p0[s] = ld_relaxed_sys_v4(e);
            p1[s] = ld_relaxed_sys_v4(e + 4);
          }
#pragma unroll
        for (int s = 0; s < kMaxTp; ++s)
          if (((pending >> s) & 1u) && p0[s].y == seq && p0[s].w == seq && p1[s].y == seq && p1[s].w == seq) pending &= ~(1u << s);
        if (pending && clock64() - t0 > 4000000000ll) {
          *t->error = 1;
          break;
        }
      }
#pragma unroll
      for (int s = 0; s < kMaxTp; ++s)
        if (s < size) {
          v.x += __uint_as_float(p0[s].x); v.y += __uint_as_float(p0[s].z); v.z += __uint_as_float(p1[s].x); v.w += __uint_as_float(p1[s].z);
        }
      *reinterpret_cast<float4*>(a.x_out + k) = v;
      st_relaxed_gpu_v4(xsum + 2 * k, make_uint4(__float_as_uint(v.x), seq, __float_as_uint(v.y), seq));
      st_relaxed_gpu_v4(xsum + 2 * k + 4, make_uint4(__float_as_uint(v.z), seq, __float_as_uint(v.w), seq));
    }
  }
  float ss = 0.f;
  for (int k = threadIdx.x * 4; k < d; k += NW * 32 * 4) {
    uint4 p0, p1;
    for (;;) {
      p0 = ld_relaxed_gpu_v4(xsum + 2 * k);
      p1 = ld_relaxed_gpu_v4(xsum + 2 * k + 4);
      if (p0.y == seq && p0.w == seq && p1.y == seq && p1.w == seq) break;
      if (clock64() - t0 > 4000000000ll) {
        *t->error = 1;
        break;
      }
    }
    const float4 v = make_float4(__uint_as_float(p0.x), __uint_as_float(p0.z), __uint_as_float(p1.x), __uint_as_float(p1.z));
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    const int idx = (k & ~255) + ((k >> 2) & 1) * 128 + ((k & 255) >> 3) * 4;
    *reinterpret_cast<float4*>(xs + idx) = v;
  }
  ss = block_sum<NW>(ss, red, warp, lane);
  const float inv = 1.0f / sqrtf(ss / (float)d + a.eps);
  for (int k = threadIdx.x * 4; k < d; k += NW * 32 * 4) {
    const int idx = (k & ~255) + ((k >> 2) & 1) * 128 + ((k & 255) >> 3) * 4;
    float4 v = *reinterpret_cast<float4*>(xs + idx);
    const float4 w = *reinterpret_cast<const float4*>(a.norm_w + k);
    v.x = __fmul_rn(__fmul_rn(v.x, inv), w.x);
    v.y = __fmul_rn(__fmul_rn(v.y, inv), w.y);
    v.z = __fmul_rn(__fmul_rn(v.z, inv), w.z);
    v.w = __fmul_rn(__fmul_rn(v.w, inv), w.w);
    *reinterpret_cast<float4*>(xs + idx) = v;
  }
  __syncthreads();
}

// 16-bit float weights: bf16 (widening is a shift) or, F16_ = true, IEEE half as stored by F16 GGUF files (a convert)
template <int KC, int R_, class C, bool F16_ = false>  // KC in {256, 512}; R_ rows per work unit (4, or 2 for matrices with few rows)
struct FmtBF16 {
  static DEVINL float lo(uint32_t v) { return F16_ ? __half2float(__ushort_as_half((unsigned short)(v & 0xffffu))) : bf16lo(v); }
  static DEVINL float hi(uint32_t v) { return F16_ ? __half2float(__ushort_as_half((unsigned short)(v >> 16))) : bf16hi(v); }
  static constexpr int R = R_, CHUNK = KC, STAGES = C::ST, NW = C::NW, OCC = C::OCC;
  static constexpr int ROW_STAGE_BYTES = KC * 2;
  static constexpr int STAGE_BYTES = R * ROW_STAGE_BYTES;
  static __host__ __device__ size_t x_bytes(int K) { return (size_t)K * 4; }
  static __host__ __device__ size_t row_bytes(int K) { return (size_t)K * 2; }

  static DEVINL void issue(uint32_t dst, uint32_t bar, const GemvArgs& a, int unit, int c, uint64_t pol) {
    const uint8_t* src = a.W + ((size_t)unit * R * a.K + (size_t)c * KC) * 2;
#pragma unroll
    for (int r = 0; r < R; ++r) bulk_g2s(dst + r * ROW_STAGE_BYTES, src + (size_t)r * a.K * 2, ROW_STAGE_BYTES, bar, pol);
  }

  // x in shared memory, permuted per 256-group so that lane l's 8 values are two conflict-free float4:
  // element 8l+j  ->  (j<4 ? 4l+j : 128+4l+(j-4))
  template <int PRO, class AfterIssue>
  static DEVINL void prologue(uint8_t* xs_, const GemvArgs& a, float* red, int warp, int lane, const float4* nw, AfterIssue&& after_issue) {
    constexpr int kProU = ProU<NW>::value;
    float* xs = reinterpret_cast<float*>(xs_);
    if (PRO == PRO_TP_RMSNORM) {
      after_issue();
      tp_prologue<NW>(xs, a, red, warp, lane);
      return;
    }
    if (PRO == PRO_RMSNORM && a.K <= NW * 32 * 4 * kProU) {
      // x and the norm weights are fetched once and stay in registers between the sum of squares and the
      // scaling (same element -> thread mapping and summation order as rms_inv, so the same bits)
      float4 v[kProU], w[kProU];
#pragma unroll
      for (int i = 0; i < kProU; ++i) {
        const int k = threadIdx.x * 4 + i * NW * 32 * 4;
        const bool on = k < a.K;
        v[i] = on ? __ldcg(reinterpret_cast<const float4*>(a.x + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
        w[i] = nw[i];  // norm weights: fetched before the dependency wait (gemv_body)
      }
      after_issue();  // the activation loads are on their way: now the rest of the weight ring may queue up behind them
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < kProU; ++i)
        if (threadIdx.x * 4 + i * NW * 32 * 4 < a.K) ss += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
      ss = block_sum<NW>(ss, red, warp, lane);
      const float inv = 1.0f / sqrtf(ss / (float)a.K + a.eps);
#pragma unroll
      for (int i = 0; i < kProU; ++i) {
        const int k = threadIdx.x * 4 + i * NW * 32 * 4;
        if (k < a.K) {
          float4 o;
          o.x = __fmul_rn(__fmul_rn(v[i].x, inv), w[i].x);
          o.y = __fmul_rn(__fmul_rn(v[i].y, inv), w[i].y);
          o.z = __fmul_rn(__fmul_rn(v[i].z, inv), w[i].z);
          o.w = __fmul_rn(__fmul_rn(v[i].w, inv), w[i].w);
          const int idx = (k & ~255) + ((k >> 2) & 1) * 128 + ((k & 255) >> 3) * 4;
          *reinterpret_cast<float4*>(xs + idx) = o;
        }
      }
      __syncthreads();
      return;
    }
    if (PRO == PRO_RMSNORM) after_issue();
    const float inv = rms_inv<PRO, NW>(a, red, warp, lane);
    bool issued = PRO == PRO_RMSNORM;
    for (int k = threadIdx.x * 4; k < a.K; k += NW * 32 * 4) {
      float4 v = __ldcg(reinterpret_cast<const float4*>(a.x + k));
      if (!issued) {  // first activation load of this thread is on its way: now the rest of the weight ring
        after_issue();
        issued = true;
      }
      if (PRO == PRO_RMSNORM) {
        const float4 w = *reinterpret_cast<const float4*>(a.norm_w + k);
        v.x = __fmul_rn(__fmul_rn(v.x, inv), w.x);
        v.y = __fmul_rn(__fmul_rn(v.y, inv), w.y);
        v.z = __fmul_rn(__fmul_rn(v.z, inv), w.z);
        v.w = __fmul_rn(__fmul_rn(v.w, inv), w.w);
      }
      const int idx = (k & ~255) + ((k >> 2) & 1) * 128 + ((k & 255) >> 3) * 4;
      *reinterpret_cast<float4*>(xs + idx) = v;
    }
    if (!issued) after_issue();  // (a thread with no element of its own)
    __syncthreads();
  }

  static DEVINL void consume(const uint8_t* st, const uint8_t* xs_, int c, const GemvArgs& a, int lane, float* acc) {
    const float* xs = reinterpret_cast<const float*>(xs_) + (size_t)c * KC;
#pragma unroll
    for (int h = 0; h < KC / 256; ++h) {
      const float4 xa = *reinterpret_cast<const float4*>(xs + h * 256 + lane * 4);
      const float4 xb = *reinterpret_cast<const float4*>(xs + h * 256 + 128 + lane * 4);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint4 w = *reinterpret_cast<const uint4*>(st + r * ROW_STAGE_BYTES + h * 512 + lane * 16);
        float s = acc[r];
        s = fmaf(lo(w.x), xa.x, s);
        s = fmaf(hi(w.x), xa.y, s);
        s = fmaf(lo(w.y), xa.z, s);
        s = fmaf(hi(w.y), xa.w, s);
        s = fmaf(lo(w.z), xb.x, s);
        s = fmaf(hi(w.z), xb.y, s);
        s = fmaf(lo(w.w), xb.z, s);
        s = fmaf(hi(w.w), xb.w, s);
        acc[r] = s;
      }
    }
  }
};

// Shared by Q4_K / Q6_K: x -> int16 per 32-block ("Q16_32": d = amax/32767, X = rint(x/d)), kept in shared
// memory as int16 in a *lane-major* order chosen per weight format, so that every lane fetches the 64
// activations its 64 weights need with eight conflict-free 16-byte loads:
//     element k of a 2048-wide chunk  ->  ((i*32 + lane)*8 + r)   where (lane, i, r) = Fmt::xpos(k)
// The integer dot products use dp2a (two int16 x two int8 per instruction) and are exact.
// smem: X[K] int16 | dx[K/32] | dxs[K/32]  (dxs = dx * sum(X), used by Q4_K's "min" term)
struct QActLayout {
  static __host__ __device__ size_t bytes(int K) { return (size_t)K * 2 + (size_t)(K / 32) * 8; }
  static DEVINL const uint4* x(const uint8_t* xs) { return reinterpret_cast<const uint4*>(xs); }
  static DEVINL const float* dx(const uint8_t* xs, int K) { return reinterpret_cast<const float*>(xs + 2 * (size_t)K); }
  static DEVINL const float* dxs(const uint8_t* xs, int K) { return dx(xs, K) + K / 32; }
};

DEVINL int dp2a_lo_su(uint32_t x2, uint32_t q4, int c) {  // x2 = 2 x s16, q4 bytes 0,1 unsigned
  int d;
  asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(x2), "r"(q4), "r"(c));
  return d;
}
DEVINL int dp2a_hi_su(uint32_t x2, uint32_t q4, int c) {  // q4 bytes 2,3 unsigned
  int d;
  asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(x2), "r"(q4), "r"(c));
  return d;
}
DEVINL int dp2a_lo_ss(uint32_t x2, uint32_t q4, int c) {  // q4 bytes 0,1 signed
  int d;
  asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(x2), "r"(q4), "r"(c));
  return d;
}
DEVINL int dp2a_hi_ss(uint32_t x2, uint32_t q4, int c) {
  int d;
  asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(x2), "r"(q4), "r"(c));
  return d;
}
// small non-negative integer (< 2^23) -> float without the conversion pipe
DEVINL float small_u2f(uint32_t v) { return __uint_as_float(0x4B000000u | v) - 8388608.0f; }

template <class F, int PRO, class AfterIssue>
DEVINL void q16_prologue(uint8_t* xs, const GemvArgs& a, float* red, int warp, int lane, const float4* nw, AfterIssue&& after_issue) {
  int16_t* xq = reinterpret_cast<int16_t*>(xs);
  float* dx = reinterpret_cast<float*>(xs + 2 * (size_t)a.K);
  float* dxs = dx + a.K / 32;
  // lane = 4 consecutive elements (one float4), 8 lanes = one 32-element block, a warp = 4 blocks per round;
  // U rounds of loads are issued before the first dependent shuffle.  ~15 instructions per block instead of
  // ~120 with one element per lane, which matters because every CTA repeats this for the whole vector.
  constexpr int NW = F::NW;
  // float4 per thread and round.  Without a norm there are no weights to hold, so twice as many activations fit in the
  // same registers: K = 14336 (ffn_down's input) is then ONE round of loads, all in flight before the first shuffle
  constexpr int U = (PRO == PRO_RMSNORM) ? ProU<NW>::value : 2 * ProU<NW>::value;
  const int n4 = a.K / 4;
  // K <= 8192 with an RMS norm: x and the norm weights are fetched once, before the block-wide sum of
  // squares, and reused from registers below (thread -> element mapping and summation order as in rms_inv)
  const bool single = (PRO == PRO_RMSNORM) && n4 <= NW * 32 * U;
  float4 v[U], g[U];
  float inv = 1.f;
  if (single) {
    float ss = 0.f;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int gi = u * NW * 32 + (int)threadIdx.x;
      v[u] = gi < n4 ? __ldcg(reinterpret_cast<const float4*>(a.x + 4 * gi)) : make_float4(0.f, 0.f, 0.f, 0.f);
      g[u] = nw[u];  // norm weights: fetched before the dependency wait (gemv_body)
    }
    after_issue();  // activation loads issued: the rest of the weight ring may now queue up behind them
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (u * NW * 32 + (int)threadIdx.x < n4) ss += v[u].x * v[u].x + v[u].y * v[u].y + v[u].z * v[u].z + v[u].w * v[u].w;
    ss = block_sum<NW>(ss, red, warp, lane);
    inv = 1.0f / sqrtf(ss / (float)a.K + a.eps);
  } else {
    if (PRO == PRO_RMSNORM) after_issue();
    inv = rms_inv<PRO, NW>(a, red, warp, lane);
  }
  bool issued = single || PRO == PRO_RMSNORM;
  for (int g0 = warp * 32; g0 < n4; g0 += NW * 32 * U) {
    if (!single) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int gi = g0 + u * NW * 32 + lane;
        v[u] = gi < n4 ? __ldcg(reinterpret_cast<const float4*>(a.x + 4 * gi)) : make_float4(0.f, 0.f, 0.f, 0.f);
        g[u] = (PRO == PRO_RMSNORM && gi < n4) ? *reinterpret_cast<const float4*>(a.norm_w + 4 * gi) : make_float4(1.f, 1.f, 1.f, 1.f);
      }
      if (!issued) {
        after_issue();
        issued = true;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int gi = g0 + u * NW * 32 + lane;
      if (g0 + u * NW * 32 >= n4) break;  // warp-uniform
      float4 val = v[u];
      if (PRO == PRO_RMSNORM) {
        val.x = __fmul_rn(__fmul_rn(val.x, inv), g[u].x);
        val.y = __fmul_rn(__fmul_rn(val.y, inv), g[u].y);
        val.z = __fmul_rn(__fmul_rn(val.z, inv), g[u].z);
        val.w = __fmul_rn(__fmul_rn(val.w, inv), g[u].w);
      }
      float m = fmaxf(fmaxf(fabsf(val.x), fabsf(val.y)), fmaxf(fabsf(val.z), fabsf(val.w)));
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
      const float d = __fmul_rn(m, 1.0f / 32767.0f);
      const float id = d > 0.f ? __frcp_rn(d) : 0.f;
      const int X0 = (int)fminf(32767.f, fmaxf(-32767.f, rintf(__fmul_rn(val.x, id))));
      const int X1 = (int)fminf(32767.f, fmaxf(-32767.f, rintf(__fmul_rn(val.y, id))));
      const int X2 = (int)fminf(32767.f, fmaxf(-32767.f, rintf(__fmul_rn(val.z, id))));
      const int X3 = (int)fminf(32767.f, fmaxf(-32767.f, rintf(__fmul_rn(val.w, id))));
      int xsum = (X0 + X1) + (X2 + X3);
      xsum += __shfl_xor_sync(0xffffffffu, xsum, 1);
      xsum += __shfl_xor_sync(0xffffffffu, xsum, 2);
      xsum += __shfl_xor_sync(0xffffffffu, xsum, 4);
      if (gi < n4) {
        const int k = 4 * gi;
        uint2 pk;
        pk.x = (uint32_t)(X0 & 0xffff) | ((uint32_t)X1 << 16);
        pk.y = (uint32_t)(X2 & 0xffff) | ((uint32_t)X3 << 16);
        *reinterpret_cast<uint2*>(xq + (k & ~2047) + F::xpos(k & 2047)) = pk;
        if ((lane & 7) == 0) {
          dx[k >> 5] = d;
          dxs[k >> 5] = __fmul_rn(d, (float)xsum);
        }
      }
    }
  }
  __syncthreads();
}

// Q4_K: a stage is R rows x 8 super-blocks (GGUF byte order, 144 B each).  lane = (block, j): the 32 qs bytes
// of sub-block pair (2j, 2j+1) = 64 weights [64j, 64j+64) of that block; both sub-block sums are complete in
// one lane, so (d*sc)*(dx*isum) is formed exactly as in the oracle.
template <int R_, class C>
struct FmtQ4K {
  static constexpr int R = R_, BLOCKS = 8, CHUNK = BLOCKS * QK_K, STAGES = C::ST, NW = C::NW, OCC = C::OCC;
  static constexpr int ROW_STAGE_BYTES = BLOCKS * Q4K_BYTES;  // 1152
  static constexpr int STAGE_BYTES = R * ROW_STAGE_BYTES;
  static __host__ __device__ size_t x_bytes(int K) { return QActLayout::bytes(K); }
  static __host__ __device__ size_t row_bytes(int K) { return (size_t)(K / QK_K) * Q4K_BYTES; }
  // k in [0,2048): block = k>>8, j = (k>>6)&3, i = (k>>3)&7, r = k&7 ; lane = block*4 + j
  static DEVINL int xpos(int k) { return ((((k >> 3) & 7) * 32 + ((k >> 8) * 4 + ((k >> 6) & 3))) << 3) + (k & 7); }

  static DEVINL void issue(uint32_t dst, uint32_t bar, const GemvArgs& a, int unit, int c, uint64_t pol) {
    const size_t rb = row_bytes(a.K);
    const uint8_t* src = a.W + (size_t)unit * R * rb + (size_t)c * ROW_STAGE_BYTES;
#pragma unroll
    for (int r = 0; r < R; ++r) bulk_g2s(dst + r * ROW_STAGE_BYTES, src + (size_t)r * rb, ROW_STAGE_BYTES, bar, pol);
  }
  template <int PRO, class AfterIssue>
  static DEVINL void prologue(uint8_t* xs, const GemvArgs& a, float* red, int warp, int lane, const float4* nw, AfterIssue&& after_issue) {
    q16_prologue<FmtQ4K<R_, C>, PRO>(xs, a, red, warp, lane, nw, after_issue);
  }

  static DEVINL void consume(const uint8_t* st, const uint8_t* xs, int c, const GemvArgs& a, int lane, float* acc) {
    const int blk = lane >> 2, j = lane & 3;
    uint32_t xr[32];  // 64 int16: [0,16) low-nibble sub-block 2j, [16,32) high-nibble sub-block 2j+1
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint4 v = QActLayout::x(xs)[(size_t)c * (CHUNK / 8) + i * 32 + lane];
      xr[4 * i] = v.x; xr[4 * i + 1] = v.y; xr[4 * i + 2] = v.z; xr[4 * i + 3] = v.w;
    }
    const int xb = (c * CHUNK + blk * QK_K) / 32 + 2 * j;
    const float dxA = QActLayout::dx(xs, a.K)[xb], dxB = QActLayout::dx(xs, a.K)[xb + 1];
    const float dxsA = QActLayout::dxs(xs, a.K)[xb], dxsB = QActLayout::dxs(xs, a.K)[xb + 1];
    const int sh = (j & 1) * 16;
    const bool upper = j >= 2;
    int iA[R], iB[R];
    uint32_t scp[R], mnp[R], dd[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {  // phase 1: integer work of every row (independent chains)
      const uint8_t* b = st + r * ROW_STAGE_BYTES + blk * Q4K_BYTES;
      const uint4 hdr = *reinterpret_cast<const uint4*>(b);
      const uint4 q0 = *reinterpret_cast<const uint4*>(b + 16 + 32 * j);
      const uint4 q1 = *reinterpret_cast<const uint4*>(b + 32 + 32 * j);
      dd[r] = hdr.x;
      // six-bit (scale, min) of sub-blocks 2j, 2j+1 (ggml get_scale_min_k4), both variants then select
      const uint32_t u = hdr.y >> sh, v = hdr.z >> sh, t = hdr.w >> sh;
      const uint32_t sc_lo = u & 0x3F3Fu, mn_lo = v & 0x3F3Fu;
      const uint32_t sc_hi = (t & 0x0F0Fu) | ((u >> 2) & 0x3030u), mn_hi = ((t >> 4) & 0x0F0Fu) | ((v >> 2) & 0x3030u);
      scp[r] = upper ? sc_hi : sc_lo;
      mnp[r] = upper ? mn_hi : mn_lo;
      const uint32_t qw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      int sa = 0, sb = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        const uint32_t lo = qw[w] & 0x0F0F0F0Fu, hi = (qw[w] >> 4) & 0x0F0F0F0Fu;
        sa = dp2a_lo_su(xr[2 * w], lo, sa);
        sa = dp2a_hi_su(xr[2 * w + 1], lo, sa);
        sb = dp2a_lo_su(xr[16 + 2 * w], hi, sb);
        sb = dp2a_hi_su(xr[17 + 2 * w], hi, sb);
      }
      iA[r] = sa;
      iB[r] = sb;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {  // phase 2: scales in fp32
      const float d = __half2float(__ushort_as_half((unsigned short)(dd[r] & 0xffffu)));
      const float dmin = __half2float(__ushort_as_half((unsigned short)(dd[r] >> 16)));
      const float scA = small_u2f(scp[r] & 0xFFu), scB = small_u2f(scp[r] >> 8);
      const float mA = small_u2f(mnp[r] & 0xFFu), mB = small_u2f(mnp[r] >> 8);
      float t = acc[r];
      t = fmaf(__fmul_rn(d, scA), __fmul_rn(dxA, (float)iA[r]), t);
      t = fmaf(__fmul_rn(d, scB), __fmul_rn(dxB, (float)iB[r]), t);
      t = fmaf(-__fmul_rn(dmin, mA), dxsA, t);
      t = fmaf(-__fmul_rn(dmin, mB), dxsB, t);
      acc[r] = t;
    }
  }
};

// Q6_K rows are stored on the device chunk-wise (8 super-blocks = 1680 B, the same bytes as GGUF's eight
// 210-byte blocks, re-ordered lane-major): [qlA 32x16][qlB 32x16][qh 32x16][scales 32x4][d 8x2].
// lane = (block, half, lh): l in [16lh,16lh+16) of the four 32-wide runs t=0..3 of that half; every run is
// exactly one 16-weight scale group.  qlA = ql[half*64 + 16lh..], qlB = ql[half*64 + 32 + 16lh..].
template <class C>
struct FmtQ6K {
  static constexpr int R = 2, BLOCKS = 8, CHUNK = BLOCKS * QK_K, STAGES = C::ST, NW = C::NW, OCC = C::OCC;
  static constexpr int ROW_STAGE_BYTES = BLOCKS * Q6K_BYTES;  // 1680
  static constexpr int STAGE_BYTES = R * ROW_STAGE_BYTES;
  static __host__ __device__ size_t x_bytes(int K) { return QActLayout::bytes(K); }
  static __host__ __device__ size_t row_bytes(int K) { return (size_t)(K / QK_K) * Q6K_BYTES; }
  // k in [0,2048): block = k>>8, half = (k>>7)&1, t = (k>>5)&3, lh = (k>>4)&1, u = (k>>3)&1, r = k&7
  static DEVINL int xpos(int k) {
    const int lane = (k >> 8) * 4 + ((k >> 7) & 1) * 2 + ((k >> 4) & 1), i = ((k >> 5) & 3) * 2 + ((k >> 3) & 1);
    return ((i * 32 + lane) << 3) + (k & 7);
  }

  static DEVINL void issue(uint32_t dst, uint32_t bar, const GemvArgs& a, int unit, int c, uint64_t pol) {
    const size_t rb = row_bytes(a.K);
    const uint8_t* src = a.W + (size_t)unit * R * rb + (size_t)c * ROW_STAGE_BYTES;
#pragma unroll
    for (int r = 0; r < R; ++r) bulk_g2s(dst + r * ROW_STAGE_BYTES, src + (size_t)r * rb, ROW_STAGE_BYTES, bar, pol);
  }
  template <int PRO, class AfterIssue>
  static DEVINL void prologue(uint8_t* xs, const GemvArgs& a, float* red, int warp, int lane, const float4* nw, AfterIssue&& after_issue) {
    q16_prologue<FmtQ6K<C>, PRO>(xs, a, red, warp, lane, nw, after_issue);
  }

  static DEVINL void consume(const uint8_t* st, const uint8_t* xs, int c, const GemvArgs& a, int lane, float* acc) {
    const int blk = lane >> 2, half = (lane >> 1) & 1, lh = lane & 1;
    uint32_t xr[32];  // run t: xr[8t .. 8t+8) = 16 int16
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint4 v = QActLayout::x(xs)[(size_t)c * (CHUNK / 8) + i * 32 + lane];
      xr[4 * i] = v.x; xr[4 * i + 1] = v.y; xr[4 * i + 2] = v.z; xr[4 * i + 3] = v.w;
    }
    float dxt[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) dxt[t] = QActLayout::dx(xs, a.K)[(c * CHUNK + blk * QK_K + half * 128) / 32 + t];
    (void)lh;
    int is[R][4];
    uint32_t scw[R];
    float dv[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint8_t* b = st + r * ROW_STAGE_BYTES;
      const uint4 q0 = *reinterpret_cast<const uint4*>(b + lane * 16);
      const uint4 q1 = *reinterpret_cast<const uint4*>(b + 512 + lane * 16);
      const uint4 qh = *reinterpret_cast<const uint4*>(b + 1024 + lane * 16);
      scw[r] = *reinterpret_cast<const uint32_t*>(b + 1536 + lane * 4);
      dv[r] = __half2float(__ushort_as_half(*reinterpret_cast<const unsigned short*>(b + 1664 + blk * 2)));
      const uint32_t a0[4] = {q0.x, q0.y, q0.z, q0.w}, a1[4] = {q1.x, q1.y, q1.z, q1.w}, ah[4] = {qh.x, qh.y, qh.z, qh.w};
      int s[4] = {0, 0, 0, 0};
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        uint32_t qt[4];
        qt[0] = (a0[w] & 0x0F0F0F0Fu) | ((ah[w] << 4) & 0x30303030u);
        qt[1] = (a1[w] & 0x0F0F0F0Fu) | ((ah[w] << 2) & 0x30303030u);
        qt[2] = ((a0[w] >> 4) & 0x0F0F0F0Fu) | (ah[w] & 0x30303030u);
        qt[3] = ((a1[w] >> 4) & 0x0F0F0F0Fu) | ((ah[w] >> 2) & 0x30303030u);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const uint32_t qs = __vsub4(qt[t], 0x20202020u);  // bytes q-32 in [-32,31]
          s[t] = dp2a_lo_ss(xr[8 * t + 2 * w], qs, s[t]);
          s[t] = dp2a_hi_ss(xr[8 * t + 2 * w + 1], qs, s[t]);
        }
      }
#pragma unroll
      for (int t = 0; t < 4; ++t) is[r][t] = s[t];
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float tacc = acc[r];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float sc = (float)(int)(signed char)((scw[r] >> (8 * t)) & 0xFFu);
        tacc = fmaf(__fmul_rn(dv[r], sc), __fmul_rn(dxt[t], (float)is[r][t]), tacc);
      }
      acc[r] = tacc;
    }
  }
};

// ------------------------------------------------------------------ epilogues
template <int EPI, int R>
DEVINL void epilogue(const GemvArgs& a, int unit, int lane, const float* acc, float& bestv, uint32_t& besti) {
  const int row0 = unit * R;
  if (EPI == EPI_STORE || EPI == EPI_LOGITS) {
    if (lane == 0) {
      if (EPI == EPI_STORE || a.store_logits) {
#pragma unroll
        for (int r = 0; r < R; ++r) a.y[row0 + r] = acc[r];
      }
      if (EPI == EPI_LOGITS) {
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (acc[r] > bestv) {
            bestv = acc[r];
            besti = (uint32_t)(a.row_base + row0 + r);  // row_base: global id of row 0 (vocab-parallel lm_head)
          }
      }
    }
  } else if (EPI == EPI_TP_PUSH) {  // lane p sends this unit's partial sums to peer p over NVLink
    const TpLink* t = a.tp;
    if (lane < t->size) {
      const unsigned int seq = *t->epoch * (unsigned int)t->per_step + (unsigned int)a.tp_index + 1u;
      float* dst = t->peer_recv[a.tp_index & 1][lane] + ((size_t)t->rank * t->d + row0) * 2;
#pragma unroll
      for (int r = 0; r < R; ++r) st_relaxed_sys_v2(dst + 2 * r, acc[r], seq);
    }
  } else if (EPI == EPI_RESID) {  // acc already holds residual + W.x (the residual was fetched at the start of the unit)
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) a.y[row0 + r] = acc[r];
    }
  } else if (EPI == EPI_SWIGLU) {  // rows interleaved gate,up,gate,up
#pragma unroll
    for (int p = 0; p < R / 2; ++p)
      if (lane == p) {
        const float g = acc[2 * p], u = acc[2 * p + 1];
        a.y[row0 / 2 + p] = (g / (1.0f + expf(-g))) * u;
      }
  } else if (EPI == EPI_QKV) {
    const int pos = a.st->pos;
#pragma unroll
    for (int p = 0; p < R / 2; ++p)
      if (lane == p) {
        const int row = a.row_base + row0 + 2 * p;
        float x0 = acc[2 * p], x1 = acc[2 * p + 1];
        if (row < a.q_rows + a.kv_rows) {  // q or k: adjacent-pair rotation
          const int i = (row & (kHeadDim - 1)) >> 1;
          const float2 cs = *reinterpret_cast<const float2*>(a.rope + ((size_t)pos * (kHeadDim / 2) + i) * 2);
          const float r0 = __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y));
          const float r1 = __fadd_rn(__fmul_rn(x0, cs.y), __fmul_rn(x1, cs.x));
          x0 = r0;
          x1 = r1;
        }
        if (row < a.q_rows) {
          *reinterpret_cast<float2*>(a.y + row) = make_float2(x0, x1);
        } else {
          const bool isk = row < a.q_rows + a.kv_rows;
          const int kr = row - a.q_rows - (isk ? 0 : a.kv_rows);
          const int kvh = kr >> 7, dim = kr & (kHeadDim - 1);
          const int page = a.block_table[pos / kPageTokens];
          const size_t idx = (((size_t)page * a.n_kv_heads + kvh) * kPageTokens + (pos % kPageTokens)) * kHeadDim + dim;
          const uint32_t packed = (uint32_t)f32_to_bf16_bits(x0) | ((uint32_t)f32_to_bf16_bits(x1) << 16);
          *reinterpret_cast<uint32_t*>((isk ? a.k_cache : a.v_cache) + idx) = packed;
        }
      }
  }
}

}  // namespace
}  // namespace b200
