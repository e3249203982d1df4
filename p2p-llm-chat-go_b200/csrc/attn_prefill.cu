// attn_prefill.cu — causal GQA attention over a whole prompt (one sequence, positions pos0..pos0+T-1), sm_100a.
//
// grid = (n_kv_heads, ceil(T/16)); one warp per query head of the KV head's group, 16 query rows per warp, so
// a CTA re-uses every K/V tile it stages for all G heads.  K/V tiles of 64 tokens are gathered through the
// block table into XOR-swizzled shared memory (bf16, exactly the cache contents); S = Q.K^T and O += P.V run on
// the tensor cores with mma.sync m16n8k16 (ldmatrix / ldmatrix.trans operands), online softmax in fp32
// registers.  Prefill attention is 1 % of the prefill FLOPs, so it does not need tcgen05; it needs to stop
// re-reading K/V per query, which the per-query decode kernel did (6.6 ms of a 16.8 ms TTFT).
//
// Numerics: Q (fp32) and P (fp32) are each fed as a bf16 "hi" plus a bf16 "lo" part (x = hi + lo to ~2^-17),
// two MMAs each, so the result matches fp32 attention over the bf16 cache to ~1e-5 and the CPU oracle needs
// no separate rounding rule for this kernel.  Output is bf16: it is the A operand of the o-projection GEMM.
#include "common.cuh"
#include "kernels.h"

namespace b200 {
namespace {

constexpr int kQRows = 16, kKvTile = 64;

DEVINL void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
DEVINL void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
DEVINL void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// (x0, x1) -> packed bf16 hi parts and packed bf16 lo parts (x = hi + lo up to 2^-17 relative)
DEVINL void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const uint16_t h0 = f32_to_bf16_bits(x0), h1 = f32_to_bf16_bits(x1);
  const float r0 = x0 - __uint_as_float((uint32_t)h0 << 16), r1 = x1 - __uint_as_float((uint32_t)h1 << 16);
  hi = (uint32_t)h0 | ((uint32_t)h1 << 16);
  lo = (uint32_t)f32_to_bf16_bits(r0) | ((uint32_t)f32_to_bf16_bits(r1) << 16);
}
// byte offset of 16-byte chunk c (0..15) of token row r in a [64][128] bf16 tile, XOR-swizzled per 8 rows
DEVINL uint32_t tile_off(int r, int c) { return (uint32_t)((r * 16 + (c ^ (r & 7))) * 16); }

template <int G>
__global__ void __launch_bounds__(G * 32) attn_prefill_kernel(const float* __restrict__ q, const uint16_t* __restrict__ k_cache,
                                                             const uint16_t* __restrict__ v_cache, const int32_t* __restrict__ block_table,
                                                             uint16_t* __restrict__ out, int T, int pos0, int n_heads, int n_kv_heads,
                                                             const AttnPrefillSegs segs) {
  extern __shared__ __align__(128) uint8_t tiles[];  // 2 stages x {K tile 16 KB, V tile 16 KB}
  if (segs.n > 0) {  // several prompts in one launch: blockIdx.z picks the prompt, its rows of q/out and its slot's block table
    const int4 sg = segs.seg[blockIdx.z];
    q += (size_t)sg.x * n_heads * kHeadDim;
    out += (size_t)sg.x * n_heads * kHeadDim;
    block_table += (size_t)sg.w * segs.pages_per_seq;
    T = sg.y; pos0 = sg.z;
    if ((int)(gridDim.y - 1 - blockIdx.y) * kQRows >= T) return;  // the grid is sized for the longest prompt
  }
  // heavy blocks first: under the causal mask the last query block does ~T/64 tiles, the first only one
  const int kvh = blockIdx.x, q0 = (gridDim.y - 1 - blockIdx.y) * kQRows;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int head = kvh * G + warp, qd = n_heads * kHeadDim;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  const int rowA = min(q0 + g, T - 1), rowB = min(q0 + g + 8, T - 1);  // clamp: rows past T compute, never store
  const int posA = pos0 + rowA, posB = pos0 + rowB;
  const int kv_len = pos0 + min(q0 + kQRows, T);  // causal: the last row of this block sees this many tokens
  const uint32_t tiles_s = smem_u32(tiles);

  // stage `st` <- K/V rows [t0, t0+64) gathered through the block table with 16-byte cp.async (zero-fill past kv_len)
  auto load_tile = [&](int st, int t0) {
    const uint32_t kb = tiles_s + st * (2 * kKvTile * 256), vb = kb + kKvTile * 256;
    for (int item = threadIdx.x; item < kKvTile * 16; item += G * 32) {
      const int r = item >> 4, c = item & 15, tok = t0 + r;
      const bool ok = tok < kv_len;
      const int page = ok ? block_table[tok / kPageTokens] : 0;
      const size_t row = (((size_t)page * n_kv_heads + kvh) * kPageTokens + (ok ? tok % kPageTokens : 0)) * kHeadDim;
      const uint32_t n = ok ? 16u : 0u;  // src-size 0: the 16 destination bytes are zero-filled
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(kb + tile_off(r, c)), "l"(k_cache + row + c * 8), "r"(n) : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(vb + tile_off(r, c)), "l"(v_cache + row + c * 8), "r"(n) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  load_tile(0, 0);

  // Q fragments (A operand), hi and lo parts: 8 k-steps of 16 dims
  uint32_t qh[8][4], ql[8][4];
  {
    const float* qa = q + (size_t)rowA * qd + (size_t)head * kHeadDim;
    const float* qb = q + (size_t)rowB * qd + (size_t)head * kHeadDim;
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
      const float2 a0 = *reinterpret_cast<const float2*>(qa + 16 * kk + 2 * t), a1 = *reinterpret_cast<const float2*>(qb + 16 * kk + 2 * t);
      const float2 a2 = *reinterpret_cast<const float2*>(qa + 16 * kk + 8 + 2 * t), a3 = *reinterpret_cast<const float2*>(qb + 16 * kk + 8 + 2 * t);
      split2(a0.x * scale, a0.y * scale, qh[kk][0], ql[kk][0]);
      split2(a1.x * scale, a1.y * scale, qh[kk][1], ql[kk][1]);
      split2(a2.x * scale, a2.y * scale, qh[kk][2], ql[kk][2]);
      split2(a3.x * scale, a3.y * scale, qh[kk][3], ql[kk][3]);
    }
  }
  float o[16][4];
#pragma unroll
  for (int j = 0; j < 16; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float mA = -INFINITY, mB = -INFINITY, lA = 0.f, lB = 0.f;

  int stage = 0;
  for (int t0 = 0; t0 < kv_len; t0 += kKvTile, stage ^= 1) {
    if (t0 + kKvTile < kv_len) {  // prefetch the next tile into the other stage, then wait for this one
      load_tile(stage ^ 1, t0 + kKvTile);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const uint32_t ks = tiles_s + stage * (2 * kKvTile * 256), vs = ks + kKvTile * 256;

    // ---- S = (Q*scale) . K^T : 8 n-tiles of 8 tokens
    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
      for (int kk = 0; kk < 8; kk += 2) {
        // four 8x8 matrices: tokens 8j..8j+7 x dim chunks 2kk, 2kk+1, 2kk+2, 2kk+3
        uint32_t b[4];
        ldsm_x4(b, ks + tile_off(8 * j + (lane & 7), 2 * kk + (lane >> 3)));
        mma_bf16(s[j], qh[kk], b[0], b[1]);
        mma_bf16(s[j], ql[kk], b[0], b[1]);
        mma_bf16(s[j], qh[kk + 1], b[2], b[3]);
        mma_bf16(s[j], ql[kk + 1], b[2], b[3]);
      }
    }
    // ---- causal mask + online softmax (rows g and g+8 of this warp's 16)
    float mxA = -INFINITY, mxB = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int tk = t0 + 8 * j + 2 * t;
      if (tk > posA) s[j][0] = -INFINITY;
      if (tk + 1 > posA) s[j][1] = -INFINITY;
      if (tk > posB) s[j][2] = -INFINITY;
      if (tk + 1 > posB) s[j][3] = -INFINITY;
      mxA = fmaxf(mxA, fmaxf(s[j][0], s[j][1]));
      mxB = fmaxf(mxB, fmaxf(s[j][2], s[j][3]));
    }
    mxA = fmaxf(mxA, __shfl_xor_sync(0xffffffffu, mxA, 1)); mxA = fmaxf(mxA, __shfl_xor_sync(0xffffffffu, mxA, 2));
    mxB = fmaxf(mxB, __shfl_xor_sync(0xffffffffu, mxB, 1)); mxB = fmaxf(mxB, __shfl_xor_sync(0xffffffffu, mxB, 2));
    const float nA = fmaxf(mA, mxA), nB = fmaxf(mB, mxB);  // finite from the first tile on (token 0 is always visible)
    const float cA = expf(mA - nA), cB = expf(mB - nB);   // exp(-inf) = 0 on the first tile
    mA = nA; mB = nB;
    lA *= cA; lB *= cB;
#pragma unroll
    for (int j = 0; j < 16; ++j) { o[j][0] *= cA; o[j][1] *= cA; o[j][2] *= cB; o[j][3] *= cB; }
    uint32_t ph[4][4], pl[4][4];  // P as A fragments: k-step kk2 covers tokens 16kk2..16kk2+15 = n-tiles 2kk2, 2kk2+1
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float p0 = expf(s[j][0] - mA), p1 = expf(s[j][1] - mA), p2 = expf(s[j][2] - mB), p3 = expf(s[j][3] - mB);
      lA += p0 + p1;
      lB += p2 + p3;
      split2(p0, p1, ph[j >> 1][(j & 1) * 2], pl[j >> 1][(j & 1) * 2]);
      split2(p2, p3, ph[j >> 1][(j & 1) * 2 + 1], pl[j >> 1][(j & 1) * 2 + 1]);
    }
    // ---- O += P . V : 16 n-tiles of 8 dims, 4 k-steps of 16 tokens
#pragma unroll
    for (int kk2 = 0; kk2 < 4; ++kk2) {
#pragma unroll
      for (int jd = 0; jd < 16; jd += 2) {
        // transposed 8x8 loads: tokens 16kk2+0..7 / +8..15 at dim chunk jd, then the same at chunk jd+1
        uint32_t b[4];
        ldsm_x4_trans(b, vs + tile_off(16 * kk2 + (lane & 7) + ((lane >> 3) & 1) * 8, jd + (lane >> 4)));
        mma_bf16(o[jd], ph[kk2], b[0], b[1]);
        mma_bf16(o[jd], pl[kk2], b[0], b[1]);
        mma_bf16(o[jd + 1], ph[kk2], b[2], b[3]);
        mma_bf16(o[jd + 1], pl[kk2], b[2], b[3]);
      }
    }
    __syncthreads();  // every warp is done with this stage before the next iteration's prefetch overwrites it
  }
  lA += __shfl_xor_sync(0xffffffffu, lA, 1); lA += __shfl_xor_sync(0xffffffffu, lA, 2);
  lB += __shfl_xor_sync(0xffffffffu, lB, 1); lB += __shfl_xor_sync(0xffffffffu, lB, 2);
  const float iA = 1.0f / lA, iB = 1.0f / lB;
  uint16_t* oa = out + (size_t)(q0 + g) * qd + (size_t)head * kHeadDim;
  uint16_t* ob = out + (size_t)(q0 + g + 8) * qd + (size_t)head * kHeadDim;
#pragma unroll
  for (int jd = 0; jd < 16; ++jd) {
    if (q0 + g < T)
      *reinterpret_cast<uint32_t*>(oa + 8 * jd + 2 * t) = (uint32_t)f32_to_bf16_bits(o[jd][0] * iA) | ((uint32_t)f32_to_bf16_bits(o[jd][1] * iA) << 16);
    if (q0 + g + 8 < T)
      *reinterpret_cast<uint32_t*>(ob + 8 * jd + 2 * t) = (uint32_t)f32_to_bf16_bits(o[jd][2] * iB) | ((uint32_t)f32_to_bf16_bits(o[jd][3] * iB) << 16);
  }
}


// ---------------------------------------------------------------------------------------------------------
// Batched decode: one query position per sequence (continuous batching), many sequences per launch.
//
// grid = (n_kv_heads, n_queries), 128 threads.  The G query heads that share a KV head are rows 0..G-1 of
// the 16-row MMA tile (rows G..15 are zero), so every K/V row a CTA stages is used by all G heads and the
// dot products run on the tensor cores: the CUDA-core kernel in attn.cu spends ~3.6 instructions per
// multiply-add on this shape (46 % of a 256-sequence decode step).  Each 64-token tile is split four ways:
// warp w owns tokens [16w, 16w+16) of every tile (two 8-token S tiles, one 16-token P.V step) and keeps its
// own online-softmax state; the four partial results are merged through shared memory at the end.
// Same numerics as the prefill kernel above (Q and P as bf16 hi + lo, fp32 accumulation).
template <int G>
__global__ void __launch_bounds__(128, 3) attn_decode_mma_kernel(const float* __restrict__ q, const uint16_t* __restrict__ k_cache,
                                                              const uint16_t* __restrict__ v_cache, const int32_t* __restrict__ block_table,
                                                              const int32_t* __restrict__ q_ctx, const int32_t* __restrict__ q_slot,
                                                              int pages_per_seq, int q_stride, int n_kv_heads, float* __restrict__ out,
                                                              uint16_t* __restrict__ out_bf16) {
  extern __shared__ __align__(128) uint8_t tiles[];  // 2 stages x {K tile 16 KB, V tile 16 KB}
  const int kvh = blockIdx.x, z = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int ctx = q_ctx[z];
  const int32_t* bt = block_table + (size_t)q_slot[z] * pages_per_seq;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  const uint32_t tiles_s = smem_u32(tiles);

  // thread -> 16-byte chunk c = tid & 15 of rows r = (tid >> 4) + 8 i, i = 0..7: a thread's eight rows fall into
  // the four 16-token pages of the tile two by two, so one block-table read serves two rows
  const int lc = threadIdx.x & 15, lr = threadIdx.x >> 4;
  auto load_tile = [&](int st, int t0) {
    const uint32_t kb = tiles_s + st * (2 * kKvTile * 256), vb = kb + kKvTile * 256;
#pragma unroll
    for (int pg = 0; pg < kKvTile / kPageTokens; ++pg) {
      const int tok0 = t0 + pg * kPageTokens;
      const int page = tok0 < ctx ? bt[tok0 / kPageTokens] : 0;
      const size_t base = ((size_t)page * n_kv_heads + kvh) * kPageTokens * kHeadDim + lc * 8;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int rp = lr + 8 * h, r = pg * kPageTokens + rp;  // row within the page / within the tile
        const uint32_t n = (tok0 + rp) < ctx ? 16u : 0u;      // src-size 0: zero-fill
        const size_t row = base + (size_t)rp * kHeadDim;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(kb + tile_off(r, lc)), "l"(k_cache + row), "r"(n) : "memory");
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(vb + tile_off(r, lc)), "l"(v_cache + row), "r"(n) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  load_tile(0, 0);

  // Q fragments: row g = head kvh*G + g for g < G, every other row of the 16 is zero
  uint32_t qh[8][2], ql[8][2];
  {
    const float* qa = q + (size_t)z * q_stride + (size_t)(kvh * G + (g < G ? g : 0)) * kHeadDim;
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
      float2 a0 = *reinterpret_cast<const float2*>(qa + 16 * kk + 2 * t), a2 = *reinterpret_cast<const float2*>(qa + 16 * kk + 8 + 2 * t);
      if (g >= G) a0 = a2 = make_float2(0.f, 0.f);
      split2(a0.x * scale, a0.y * scale, qh[kk][0], ql[kk][0]);
      split2(a2.x * scale, a2.y * scale, qh[kk][1], ql[kk][1]);
    }
  }
  float o[16][4];
#pragma unroll
  for (int j = 0; j < 16; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m = -INFINITY, l = 0.f;

  int stage = 0;
  for (int t0 = 0; t0 < ctx; t0 += kKvTile, stage ^= 1) {
    if (t0 + kKvTile < ctx) {
      load_tile(stage ^ 1, t0 + kKvTile);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const int tw = t0 + 16 * warp;  // this warp's 16 tokens of the tile
    if (tw < ctx) {                 // warp-uniform
      const uint32_t ks = tiles_s + stage * (2 * kKvTile * 256), vs = ks + kKvTile * 256;
      float s[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
        for (int kk = 0; kk < 8; kk += 2) {
          uint32_t b[4];
          ldsm_x4(b, ks + tile_off(16 * warp + 8 * j + (lane & 7), 2 * kk + (lane >> 3)));
          const uint32_t ah0[4] = {qh[kk][0], 0u, qh[kk][1], 0u}, al0[4] = {ql[kk][0], 0u, ql[kk][1], 0u};
          const uint32_t ah1[4] = {qh[kk + 1][0], 0u, qh[kk + 1][1], 0u}, al1[4] = {ql[kk + 1][0], 0u, ql[kk + 1][1], 0u};
          mma_bf16(s[j], ah0, b[0], b[1]);
          mma_bf16(s[j], al0, b[0], b[1]);
          mma_bf16(s[j], ah1, b[2], b[3]);
          mma_bf16(s[j], al1, b[2], b[3]);
        }
      }
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int tk = tw + 8 * j + 2 * t;
        if (tk >= ctx) s[j][0] = -INFINITY;
        if (tk + 1 >= ctx) s[j][1] = -INFINITY;
        mx = fmaxf(mx, fmaxf(s[j][0], s[j][1]));
      }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float mn = fmaxf(m, mx);  // finite: token tw < ctx is visible to every row
      const float corr = expf(m - mn);  // exp(-inf) = 0 on this warp's first tile
      m = mn;
      l *= corr;
#pragma unroll
      for (int j = 0; j < 16; ++j) { o[j][0] *= corr; o[j][1] *= corr; }
      const float p0 = expf(s[0][0] - m), p1 = expf(s[0][1] - m), p2 = expf(s[1][0] - m), p3 = expf(s[1][1] - m);
      l += (p0 + p1) + (p2 + p3);
      uint32_t ph[4] = {0u, 0u, 0u, 0u}, pl[4] = {0u, 0u, 0u, 0u};
      split2(p0, p1, ph[0], pl[0]);
      split2(p2, p3, ph[2], pl[2]);
#pragma unroll
      for (int jd = 0; jd < 16; jd += 2) {
        uint32_t b[4];
        ldsm_x4_trans(b, vs + tile_off(16 * warp + (lane & 7) + ((lane >> 3) & 1) * 8, jd + (lane >> 4)));
        mma_bf16(o[jd], ph, b[0], b[1]);
        mma_bf16(o[jd], pl, b[0], b[1]);
        mma_bf16(o[jd + 1], ph, b[2], b[3]);
        mma_bf16(o[jd + 1], pl, b[2], b[3]);
      }
    }
    __syncthreads();
  }
  // merge the four warps' (m, l, o) through shared memory (the tiles are no longer needed)
  l += __shfl_xor_sync(0xffffffffu, l, 1);
  l += __shfl_xor_sync(0xffffffffu, l, 2);
  float* os = reinterpret_cast<float*>(tiles);            // [4 warps][G][128]
  float* ms = os + 4 * G * kHeadDim;                       // [4][G]
  float* ls = ms + 4 * G;                                  // [4][G]
  if (g < G) {
#pragma unroll
    for (int jd = 0; jd < 16; ++jd)
      *reinterpret_cast<float2*>(os + ((size_t)warp * G + g) * kHeadDim + 8 * jd + 2 * t) = make_float2(o[jd][0], o[jd][1]);
    if (t == 0) {
      ms[warp * G + g] = m;
      ls[warp * G + g] = l;
    }
  }
  __syncthreads();
  const int dim = threadIdx.x;
#pragma unroll
  for (int h = 0; h < G; ++h) {
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) M = fmaxf(M, ms[w * G + h]);
    float num = 0.f, den = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float mw = ms[w * G + h];
      const float wt = (mw == -INFINITY) ? 0.f : expf(mw - M);  // a warp that saw no token of this sequence
      num = fmaf(wt, os[((size_t)w * G + h) * kHeadDim + dim], num);
      den = fmaf(wt, ls[w * G + h], den);
    }
    const float r = num / den;
    const size_t idx = (size_t)z * q_stride + (size_t)(kvh * G + h) * kHeadDim + dim;
    if (out_bf16) out_bf16[idx] = f32_to_bf16_bits(r);
    else out[idx] = r;
  }
}

}  // namespace

namespace {
cudaError_t launch_attn_prefill_any(const float* q, const uint16_t* k_cache, const uint16_t* v_cache, const int32_t* block_table, uint16_t* out,
                                    int T, int pos0, int n_heads, int n_kv_heads, const AttnPrefillSegs& segs, cudaStream_t s) {
  if (T < 1 || n_kv_heads < 1 || n_heads % n_kv_heads) return cudaErrorInvalidValue;
  dim3 grid(n_kv_heads, (T + kQRows - 1) / kQRows, segs.n > 0 ? segs.n : 1);
  constexpr int kSmem = 2 * 2 * kKvTile * 256;  // 64 KB
#define LAUNCH(G_)                                                                                                   \
  {                                                                                                                  \
    cudaError_t e_ = cudaFuncSetAttribute(attn_prefill_kernel<G_>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem); \
    if (e_ != cudaSuccess) return e_;                                                                                \
    count_launch(); attn_prefill_kernel<G_><<<grid, G_ * 32, kSmem, s>>>
  switch (n_heads / n_kv_heads) {
    case 1: LAUNCH(1)(q, k_cache, v_cache, block_table, out, T, pos0, n_heads, n_kv_heads, segs); } break;
    case 2: LAUNCH(2)(q, k_cache, v_cache, block_table, out, T, pos0, n_heads, n_kv_heads, segs); } break;
    case 4: LAUNCH(4)(q, k_cache, v_cache, block_table, out, T, pos0, n_heads, n_kv_heads, segs); } break;
    case 8: LAUNCH(8)(q, k_cache, v_cache, block_table, out, T, pos0, n_heads, n_kv_heads, segs); } break;
    default: return cudaErrorInvalidValue;
  }
#undef LAUNCH
  return cudaGetLastError();
}
}  // namespace

// q: [T][n_heads*128] fp32 (rotated); out: [T][n_heads*128] bf16.  K/V of positions [0, pos0+T) must be in the cache.
cudaError_t launch_attn_prefill(const float* q, const uint16_t* k_cache, const uint16_t* v_cache, const int32_t* block_table, uint16_t* out,
                                int T, int pos0, int n_heads, int n_kv_heads, cudaStream_t s) {
  AttnPrefillSegs none = {};
  return launch_attn_prefill_any(q, k_cache, v_cache, block_table, out, T, pos0, n_heads, n_kv_heads, none, s);
}

// Several prompts packed row-wise into q/out (segment i: rows [seg.x, seg.x + seg.y), first position seg.z, cache slot
// seg.w): one launch, grid.z = prompt, so short and long prompts fill the GPU together instead of one after the other.
cudaError_t launch_attn_prefill_segs(const float* q, const uint16_t* k_cache, const uint16_t* v_cache, const int32_t* block_table,
                                     int pages_per_seq, uint16_t* out, const int4* segs, int n_segs, int n_heads, int n_kv_heads, cudaStream_t s) {
  for (int i0 = 0; i0 < n_segs; i0 += AttnPrefillSegs::kMax) {
    AttnPrefillSegs a = {};
    a.n = n_segs - i0 < AttnPrefillSegs::kMax ? n_segs - i0 : AttnPrefillSegs::kMax;
    a.pages_per_seq = pages_per_seq;
    int t_max = 1;
    for (int i = 0; i < a.n; ++i) {
      a.seg[i] = segs[i0 + i];
      if (segs[i0 + i].y < 1) return cudaErrorInvalidValue;
      if (segs[i0 + i].y > t_max) t_max = segs[i0 + i].y;
    }
    cudaError_t e = launch_attn_prefill_any(q, k_cache, v_cache, block_table, out, t_max, 0, n_heads, n_kv_heads, a, s);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// One query per sequence: query z attends to the first a.q_ctx[z] tokens of the sequence in slot a.q_slot[z].
// Writes bf16 to out_bf16 when given (the o-projection GEMM's operand), else fp32 to a.out.
cudaError_t launch_attn_decode_batched(const AttnArgs& a, uint16_t* out_bf16, cudaStream_t s) {
  if (a.n_kv_heads < 1 || a.n_heads % a.n_kv_heads || a.n_queries < 1 || !a.q_ctx || !a.q_slot) return cudaErrorInvalidValue;
  dim3 grid(a.n_kv_heads, a.n_queries);
  constexpr int kSmem = 2 * 2 * kKvTile * 256;  // 64 KB
#define LAUNCH(G_)                                                                                                         \
  {                                                                                                                        \
    cudaError_t e_ = cudaFuncSetAttribute(attn_decode_mma_kernel<G_>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem); \
    if (e_ != cudaSuccess) return e_;                                                                                      \
    e_ = cudaFuncSetAttribute(attn_decode_mma_kernel<G_>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);            \
    if (e_ != cudaSuccess) return e_;                                                                                      \
    count_launch(); attn_decode_mma_kernel<G_><<<grid, 128, kSmem, s>>>(a.q, a.k_cache, a.v_cache, a.block_table, a.q_ctx, a.q_slot,       \
                                                        a.pages_per_seq, a.q_stride, a.n_kv_heads, a.out, out_bf16);        \
  }
  switch (a.n_heads / a.n_kv_heads) {
    case 1: LAUNCH(1) break;
    case 2: LAUNCH(2) break;
    case 4: LAUNCH(4) break;
    case 8: LAUNCH(8) break;
    default: return cudaErrorInvalidValue;
  }
#undef LAUNCH
  return cudaGetLastError();
}

}  // namespace b200
