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
// 32-element block (hi/lo byte planes, dp4a) with exact integer sub-block sums.
#include "common.cuh"
#include "kernels.h"

namespace b200 {

namespace {

constexpr int kWarps = 16;  // consumer warps per CTA (each is its own TMA producer); one CTA per SM

// ------------------------------------------------------------------ block helpers
DEVINL float block_sum(float v, float* red, int warp, int lane) {
  v = warp_sum(v);
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < kWarps; ++i) t += red[i];
  __syncthreads();
  return t;
}

template <int PRO>
DEVINL float rms_inv(const GemvArgs& a, float* red, int warp, int lane) {
  if (PRO != PRO_RMSNORM) return 1.f;
  float ss = 0.f;
  for (int k = threadIdx.x * 4; k < a.K; k += kWarps * 32 * 4) {
    const float4 v = *reinterpret_cast<const float4*>(a.x + k);
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = block_sum(ss, red, warp, lane);
  return 1.0f / sqrtf(ss / (float)a.K + a.eps);
}

// ------------------------------------------------------------------ format policies
// Each policy states: R rows per unit, CHUNK columns per stage, the stage layout, how lane 0 issues the
// bulk copies of a stage, how the block prepares x in shared memory, and how a warp consumes a stage.

template <int KC>  // KC in {256, 512}
struct FmtBF16 {
  static constexpr int R = 4, CHUNK = KC, STAGES = 2;
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
  template <int PRO>
  static DEVINL void prologue(uint8_t* xs_, const GemvArgs& a, float* red, int warp, int lane) {
    float* xs = reinterpret_cast<float*>(xs_);
    const float inv = rms_inv<PRO>(a, red, warp, lane);
    for (int k = threadIdx.x * 4; k < a.K; k += kWarps * 32 * 4) {
      float4 v = *reinterpret_cast<const float4*>(a.x + k);
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
        s = fmaf(bf16lo(w.x), xa.x, s);
        s = fmaf(bf16hi(w.x), xa.y, s);
        s = fmaf(bf16lo(w.y), xa.z, s);
        s = fmaf(bf16hi(w.y), xa.w, s);
        s = fmaf(bf16lo(w.z), xb.x, s);
        s = fmaf(bf16hi(w.z), xb.y, s);
        s = fmaf(bf16lo(w.w), xb.z, s);
        s = fmaf(bf16hi(w.w), xb.w, s);
        acc[r] = s;
      }
    }
  }
};

// Shared by Q4_K / Q6_K: x -> int16 per 32-block (Q16_32), stored as a signed hi-byte plane and an
// unsigned lo-byte plane (X = 256*hi + lo) so that dp4a does the integer dot products exactly.
// smem: Xhi[K] | Xlo[K] | dx[K/32] | dxs[K/32]  (dxs = dx * sum(X), used by Q4_K's "min" term)
struct QActLayout {
  static __host__ __device__ size_t bytes(int K) { return (size_t)K * 2 + (size_t)(K / 32) * 8; }
  static DEVINL const int8_t* hi(const uint8_t* xs) { return reinterpret_cast<const int8_t*>(xs); }
  static DEVINL const uint8_t* lo(const uint8_t* xs, int K) { return xs + K; }
  static DEVINL const float* dx(const uint8_t* xs, int K) { return reinterpret_cast<const float*>(xs + 2 * (size_t)K); }
  static DEVINL const float* dxs(const uint8_t* xs, int K) { return dx(xs, K) + K / 32; }
};

template <int PRO>
DEVINL void q16_prologue(uint8_t* xs, const GemvArgs& a, float* red, int warp, int lane) {
  const float inv = rms_inv<PRO>(a, red, warp, lane);
  int8_t* xh = reinterpret_cast<int8_t*>(xs);
  uint8_t* xl = xs + a.K;
  float* dx = reinterpret_cast<float*>(xs + 2 * (size_t)a.K);
  float* dxs = dx + a.K / 32;
  for (int b = warp; b < a.K / 32; b += kWarps) {
    const int k = b * 32 + lane;
    float v = a.x[k];
    if (PRO == PRO_RMSNORM) v = __fmul_rn(__fmul_rn(v, inv), a.norm_w[k]);
    const float amax = warp_max(fabsf(v));
    const float d = __fdiv_rn(amax, 32767.0f);
    const float id = d > 0.f ? __fdiv_rn(1.0f, d) : 0.f;
    float r = rintf(__fmul_rn(v, id));
    r = fminf(32767.f, fmaxf(-32767.f, r));
    const int X = (int)r;
    xh[k] = (int8_t)(X >> 8);
    xl[k] = (uint8_t)(X & 255);
    const int xsum = __reduce_add_sync(0xffffffffu, X);
    if (lane == 0) {
      dx[b] = d;
      dxs[b] = __fmul_rn(d, (float)xsum);
    }
  }
  __syncthreads();
}

struct FmtQ4K {
  static constexpr int R = 4, BLOCKS = 4, CHUNK = BLOCKS * QK_K, STAGES = 3;
  static constexpr int ROW_STAGE_BYTES = BLOCKS * Q4K_BYTES;  // 576
  static constexpr int STAGE_BYTES = R * ROW_STAGE_BYTES;
  static __host__ __device__ size_t x_bytes(int K) { return QActLayout::bytes(K); }
  static __host__ __device__ size_t row_bytes(int K) { return (size_t)(K / QK_K) * Q4K_BYTES; }

  static DEVINL void issue(uint32_t dst, uint32_t bar, const GemvArgs& a, int unit, int c, uint64_t pol) {
    const size_t rb = row_bytes(a.K);
    const uint8_t* src = a.W + (size_t)unit * R * rb + (size_t)c * ROW_STAGE_BYTES;
#pragma unroll
    for (int r = 0; r < R; ++r) bulk_g2s(dst + r * ROW_STAGE_BYTES, src + (size_t)r * rb, ROW_STAGE_BYTES, bar, pol);
  }
  template <int PRO>
  static DEVINL void prologue(uint8_t* xs, const GemvArgs& a, float* red, int warp, int lane) {
    q16_prologue<PRO>(xs, a, red, warp, lane);
  }

  // lane -> (block = lane/8, s = lane%8): 16 bytes of qs = low nibbles of sub-block 2j (16 weights at
  // 64j+16h) and high nibbles of sub-block 2j+1 (16 weights at 64j+32+16h), j = s/2, h = s%2.
  static DEVINL void consume(const uint8_t* st, const uint8_t* xs, int c, const GemvArgs& a, int lane, float* acc) {
    const int blk = lane >> 3, s = lane & 7, j = s >> 1, h = s & 1;
    const int e0 = c * CHUNK + blk * QK_K + 64 * j + 16 * h;
    const uint4 xh0 = *reinterpret_cast<const uint4*>(QActLayout::hi(xs) + e0);
    const uint4 xl0 = *reinterpret_cast<const uint4*>(QActLayout::lo(xs, a.K) + e0);
    const uint4 xh1 = *reinterpret_cast<const uint4*>(QActLayout::hi(xs) + e0 + 32);
    const uint4 xl1 = *reinterpret_cast<const uint4*>(QActLayout::lo(xs, a.K) + e0 + 32);
    const int xb = (c * CHUNK + blk * QK_K) / 32;
    const float dxA = QActLayout::dx(xs, a.K)[xb + 2 * j];
    const float dxB = QActLayout::dx(xs, a.K)[xb + 2 * j + 1];
    const float dxsS = QActLayout::dxs(xs, a.K)[xb + s];
    const int sh = (j & 1) * 16;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint8_t* b = st + r * ROW_STAGE_BYTES + blk * Q4K_BYTES;
      const uint4 hdr = *reinterpret_cast<const uint4*>(b);
      const uint4 q = *reinterpret_cast<const uint4*>(b + 16 + 16 * s);
      const float d = __half2float(__ushort_as_half((unsigned short)(hdr.x & 0xffffu)));
      const float dmin = __half2float(__ushort_as_half((unsigned short)(hdr.x >> 16)));
      // six-bit (scale, min) pairs of sub-blocks 2j, 2j+1 (ggml get_scale_min_k4), two at a time
      uint32_t scp, mnp;
      if (j < 2) {
        scp = (hdr.y >> sh) & 0x3F3Fu;
        mnp = (hdr.z >> sh) & 0x3F3Fu;
      } else {
        const uint32_t t = (hdr.w >> sh) & 0xFFFFu, u = (hdr.y >> sh) & 0xFFFFu, v = (hdr.z >> sh) & 0xFFFFu;
        scp = (t & 0x0F0Fu) | ((u >> 2) & 0x3030u);
        mnp = ((t >> 4) & 0x0F0Fu) | ((v >> 2) & 0x3030u);
      }
      const float scA = (float)(scp & 0xFFu), scB = (float)(scp >> 8);
      const float mS = (float)(h ? (mnp >> 8) : (mnp & 0xFFu));
      int iAh = 0, iAl = 0, iBh = 0, iBl = 0;
      const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
      const uint32_t h0[4] = {xh0.x, xh0.y, xh0.z, xh0.w}, l0[4] = {xl0.x, xl0.y, xl0.z, xl0.w};
      const uint32_t h1[4] = {xh1.x, xh1.y, xh1.z, xh1.w}, l1[4] = {xl1.x, xl1.y, xl1.z, xl1.w};
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const int lo = (int)(qw[w] & 0x0F0F0F0Fu), hi = (int)((qw[w] >> 4) & 0x0F0F0F0Fu);
        iAh = dp4a_ss(lo, (int)h0[w], iAh);
        iAl = dp4a_su(lo, l0[w], iAl);
        iBh = dp4a_ss(hi, (int)h1[w], iBh);
        iBl = dp4a_su(hi, l1[w], iBl);
      }
      const float fa = (float)(iAh * 256 + iAl), fb = (float)(iBh * 256 + iBl);
      float t = acc[r];
      t = fmaf(__fmul_rn(d, scA), __fmul_rn(dxA, fa), t);
      t = fmaf(__fmul_rn(d, scB), __fmul_rn(dxB, fb), t);
      t = fmaf(-__fmul_rn(dmin, mS), dxsS, t);
      acc[r] = t;
    }
  }
};

// Q6_K rows are stored on the device section-wise: [ql: nb*128][qh: nb*64][scales: nb*16][d: nb*2]
// (same bytes as GGUF's 210-byte blocks, re-ordered so that every section is 16-byte aligned).
struct FmtQ6K {
  static constexpr int R = 2, BLOCKS = 8, CHUNK = BLOCKS * QK_K, STAGES = 2;
  static constexpr int QL = BLOCKS * 128, QH = BLOCKS * 64, SC = BLOCKS * 16, DD = BLOCKS * 2;
  static constexpr int ROW_STAGE_BYTES = QL + QH + SC + DD;  // 1680
  static constexpr int STAGE_BYTES = R * ROW_STAGE_BYTES;
  static __host__ __device__ size_t x_bytes(int K) { return QActLayout::bytes(K); }
  static __host__ __device__ size_t row_bytes(int K) { return (size_t)(K / QK_K) * Q6K_BYTES; }

  static DEVINL void issue(uint32_t dst, uint32_t bar, const GemvArgs& a, int unit, int c, uint64_t pol) {
    const int nb = a.K / QK_K;
    const size_t rb = row_bytes(a.K);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint8_t* row = a.W + ((size_t)unit * R + r) * rb;
      const uint32_t d = dst + r * ROW_STAGE_BYTES;
      bulk_g2s(d, row + (size_t)c * QL, QL, bar, pol);
      bulk_g2s(d + QL, row + (size_t)nb * 128 + (size_t)c * QH, QH, bar, pol);
      bulk_g2s(d + QL + QH, row + (size_t)nb * 192 + (size_t)c * SC, SC, bar, pol);
      bulk_g2s(d + QL + QH + SC, row + (size_t)nb * 208 + (size_t)c * DD, DD, bar, pol);
    }
  }
  template <int PRO>
  static DEVINL void prologue(uint8_t* xs, const GemvArgs& a, float* red, int warp, int lane) {
    q16_prologue<PRO>(xs, a, red, warp, lane);
  }

  // lane -> (block = lane/4, half = (lane%4)/2, lh = lane%2): 16 values of l in [16lh,16lh+16) for the four
  // 32-wide runs t=0..3 of that half; each run is exactly one 16-weight scale group.
  static DEVINL void consume(const uint8_t* st, const uint8_t* xs, int c, const GemvArgs& a, int lane, float* acc) {
    const int blk = lane >> 2, half = (lane >> 1) & 1, lh = lane & 1;
    const int e = c * CHUNK + blk * QK_K + half * 128 + 16 * lh;  // + 32 t
    uint32_t xh[4][4], xl[4][4];
    float dxt[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const uint4 vh = *reinterpret_cast<const uint4*>(QActLayout::hi(xs) + e + 32 * t);
      const uint4 vl = *reinterpret_cast<const uint4*>(QActLayout::lo(xs, a.K) + e + 32 * t);
      xh[t][0] = vh.x; xh[t][1] = vh.y; xh[t][2] = vh.z; xh[t][3] = vh.w;
      xl[t][0] = vl.x; xl[t][1] = vl.y; xl[t][2] = vl.z; xl[t][3] = vl.w;
      dxt[t] = QActLayout::dx(xs, a.K)[(e + 32 * t) / 32];
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint8_t* b = st + r * ROW_STAGE_BYTES;
      const uint4 q0 = *reinterpret_cast<const uint4*>(b + blk * 128 + half * 64 + 16 * lh);
      const uint4 q1 = *reinterpret_cast<const uint4*>(b + blk * 128 + half * 64 + 32 + 16 * lh);
      const uint4 qh = *reinterpret_cast<const uint4*>(b + QL + blk * 64 + half * 32 + 16 * lh);
      const int8_t* sc = reinterpret_cast<const int8_t*>(b + QL + QH + blk * 16 + half * 8 + lh);
      const float d = __half2float(__ushort_as_half(*reinterpret_cast<const unsigned short*>(b + QL + QH + SC + blk * 2)));
      const uint32_t a0[4] = {q0.x, q0.y, q0.z, q0.w}, a1[4] = {q1.x, q1.y, q1.z, q1.w}, ah[4] = {qh.x, qh.y, qh.z, qh.w};
      int ih[4] = {0, 0, 0, 0}, il[4] = {0, 0, 0, 0};
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        uint32_t qt[4];
        qt[0] = (a0[w] & 0x0F0F0F0Fu) | ((ah[w] << 4) & 0x30303030u);
        qt[1] = (a1[w] & 0x0F0F0F0Fu) | ((ah[w] << 2) & 0x30303030u);
        qt[2] = ((a0[w] >> 4) & 0x0F0F0F0Fu) | (ah[w] & 0x30303030u);
        qt[3] = ((a1[w] >> 4) & 0x0F0F0F0Fu) | ((ah[w] >> 2) & 0x30303030u);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int qs = (int)__vsub4(qt[t], 0x20202020u);  // bytes q-32 in [-32,31]
          ih[t] = dp4a_ss(qs, (int)xh[t][w], ih[t]);
          il[t] = dp4a_su(qs, xl[t][w], il[t]);
        }
      }
      float tacc = acc[r];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float f = (float)(ih[t] * 256 + il[t]);
        tacc = fmaf(__fmul_rn(d, (float)sc[2 * t]), __fmul_rn(dxt[t], f), tacc);
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
            besti = (uint32_t)(row0 + r);
          }
      }
    }
  } else if (EPI == EPI_RESID) {
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < R; ++r) a.y[row0 + r] += acc[r];
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

// ------------------------------------------------------------------ the kernel
template <class F, int PRO, int EPI>
__global__ void __launch_bounds__(kWarps * 32, 1) gemv_kernel(const GemvArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  // warp index through a shuffle so the compiler knows it is warp-uniform: the TMA issue below then takes
  // its addresses from uniform registers instead of a per-lane "waterfall" loop around every UBLKCP
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const size_t xb = (F::x_bytes(a.K) + 127) & ~(size_t)127;
  uint8_t* xs = smem;
  uint8_t* ring = smem + xb + (size_t)warp * F::STAGES * F::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + xb + (size_t)kWarps * F::STAGES * F::STAGE_BYTES) + warp * F::STAGES;
  float* red = reinterpret_cast<float*>(smem + xb + (size_t)kWarps * F::STAGES * F::STAGE_BYTES + kWarps * F::STAGES * 8);

  // unit u (R rows, full K) -> global warp u % twarps; consecutive units land on consecutive CTAs, so a
  // matrix with fewer units than warps still spreads over every SM
  const int n_units = a.N / F::R;
  const int chunks = a.K / F::CHUNK;
  const int gwarp = warp * (int)gridDim.x + (int)blockIdx.x, twarps = (int)gridDim.x * kWarps;
  const int my_units = gwarp < n_units ? (n_units - gwarp + twarps - 1) / twarps : 0;
  const int n_items = my_units * chunks;
  const uint32_t ring_s = smem_u32(ring), bars_s = smem_u32(bars);
  const uint64_t pol = l2_policy_evict_first();

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
#pragma unroll
  for (int s = 0; s < F::STAGES; ++s)
    if (p_item < n_items) produce();
  if (EPI != EPI_LOGITS) {
    pdl_launch_dependents();
    pdl_wait();
  }
  F::template prologue<PRO>(xs, a, red, warp, lane);
  if (EPI == EPI_LOGITS) pdl_launch_dependents();

  int c_stage = 0;
  uint32_t c_parity = 0;
  float bestv = -INFINITY;
  uint32_t besti = 0;
  for (int u = 0; u < my_units; ++u) {
    float acc[F::R];
#pragma unroll
    for (int r = 0; r < F::R; ++r) acc[r] = 0.f;
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
    epilogue<EPI, F::R>(a, gwarp + u * twarps, lane, acc, bestv, besti);
  }
  if (EPI == EPI_LOGITS && lane == 0 && my_units > 0) atomicMax(a.argmax_key, argmax_key(bestv, besti));
}

template <class F>
size_t smem_bytes(int K) {
  const size_t xb = (F::x_bytes(K) + 127) & ~(size_t)127;
  return xb + (size_t)kWarps * F::STAGES * F::STAGE_BYTES + kWarps * F::STAGES * 8 + kWarps * 4 + 32;
}

template <class F, int PRO, int EPI>
cudaError_t launch_t(const GemvArgs& a, cudaStream_t s, bool pdl, int sm_count) {
  auto kern = gemv_kernel<F, PRO, EPI>;
  const size_t smem = smem_bytes<F>(a.K);
  static thread_local size_t configured = 0;  // per instantiation
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  const int n_units = a.N / F::R;
  int grid = n_units < sm_count ? n_units : sm_count;  // units are dealt CTA-first (see gemv_kernel)
  if (grid < 1) grid = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kWarps * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, a);
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
#undef CASE
  return cudaErrorInvalidValue;
}

}  // namespace

size_t gemv_row_bytes(int wtype, int64_t K) {
  switch (wtype) {
    case T_BF16: return (size_t)K * 2;
    case T_Q4_K: return (size_t)(K / QK_K) * Q4K_BYTES;
    case T_Q6_K: return (size_t)(K / QK_K) * Q6K_BYTES;
  }
  return 0;
}

bool gemv_shape_ok(int wtype, int64_t N, int64_t K, int epi) {
  if (N <= 0 || K <= 0 || K > 32768) return false;
  switch (wtype) {
    case T_BF16: return K % 256 == 0 && N % 4 == 0;
    case T_Q4_K: return K % FmtQ4K::CHUNK == 0 && N % FmtQ4K::R == 0;
    case T_Q6_K: return K % FmtQ6K::CHUNK == 0 && N % FmtQ6K::R == 0;
  }
  (void)epi;
  return false;
}

cudaError_t launch_gemv(int wtype, int pro, int epi, const GemvArgs& a, cudaStream_t s, bool pdl, int sm_count) {
  if (!gemv_shape_ok(wtype, a.N, a.K, epi)) return cudaErrorInvalidValue;
  switch (wtype) {
    case T_BF16:
      return (a.K % 512 == 0) ? launch_f<FmtBF16<512>>(pro, epi, a, s, pdl, sm_count)
                              : launch_f<FmtBF16<256>>(pro, epi, a, s, pdl, sm_count);
    case T_Q4_K: return launch_f<FmtQ4K>(pro, epi, a, s, pdl, sm_count);
    case T_Q6_K: return launch_f<FmtQ6K>(pro, epi, a, s, pdl, sm_count);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b200
