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

// ---------------------------------------------------------------------------------------------------------
// Batch-1 decode step variant: the same split-KV work division, with everything that does not depend on the
// kernel before it (the qkv projection of this layer) moved in front of griddepcontrol.wait.  Under
// programmatic dependent launch these CTAs (128 threads, ~21 KB of shared memory) become resident next to the
// still-running projection kernel, so by the time the query vector exists the K / V rows of the context are
// already in shared memory: block-table look-up and the HBM round trip of the gather (~2.5 us) leave the
// critical path.  Only the row of the current position — written by that projection — is fetched after the wait.
//
// K tile layout: [32 tokens][16 chunks of 16 B], chunk c of token t stored at chunk (c ^ (t & 7)): 16-byte
// cp.async fills it, and lane = token reads whole chunks (LDS.128) without bank conflicts.
template <int G>
struct AttnFastSmem {
  alignas(16) uint16_t Ks[kTile * kHeadDim];
  alignas(16) uint16_t Vs[kTile * kHeadDim];
  alignas(16) float qs[G * kHeadDim];
  float P[G][kTile];
  float m_run[G], l_run[G], corr_s[G];
  int s_last;
};

DEVINL void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
DEVINL void cp_async_commit_wait_all() {
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int G>
__global__ void __launch_bounds__(128) attn_decode_fast_kernel(const AttnArgs a) {
  __shared__ AttnFastSmem<G> sm;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kvh = blockIdx.x, split = blockIdx.y, S = gridDim.y;
  const uint32_t ks_s = smem_u32(sm.Ks), vs_s = smem_u32(sm.Vs);
  // rows [row_lo, row_hi) of tile `tile` -> shared memory (thread = 16-byte chunk; 4 chunks of K and of V each)
  auto gather = [&](int tile, int row_lo, int row_hi) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int item = tid + 128 * j, t = item >> 4, c = item & 15;
      const int tok = tile * kTile + t;
      if (tok >= row_lo && tok < row_hi) {
        const int page = a.block_table[tok / kPageTokens];
        const size_t row = (((size_t)page * a.n_kv_heads + kvh) * kPageTokens + (tok % kPageTokens)) * kHeadDim;
        cp_async16(ks_s + (uint32_t)(t * kHeadDim * 2 + ((c ^ (t & 7)) << 4)), a.k_cache + row + c * 8);
        cp_async16(vs_s + (uint32_t)(t * kHeadDim * 2 + (c << 4)), a.v_cache + row + c * 8);
      }
    }
  };
  // ---- before the dependency: the first tile of this split, every token the cache has room for (rows at or
  // beyond the current position hold older data; they are masked below, and the current row is re-fetched)
  const bool tracer = a.trace != nullptr && kvh == 0 && split == 0 && tid == 0;
  if (tracer) a.trace[0] = global_timer_ns();
  gather(split, 0, a.prefetch_tokens);
  asm volatile("cp.async.commit_group;" ::: "memory");
  if (tid < G) {
    sm.m_run[tid] = -INFINITY;
    sm.l_run[tid] = 0.f;
  }
  pdl_wait();
  pdl_launch_dependents();
  if (tracer) a.trace[1] = global_timer_ns();
  const int pos = a.fixed_ctx > 0 ? a.fixed_ctx - 1 : a.st->pos;
  const int ctx = pos + 1;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();  // the prefetched rows have landed before the current row is written over its stale copy
  {
    const int lo = pos > a.prefetch_tokens ? a.prefetch_tokens : pos;  // rows not covered by the prefetch, and `pos`
    gather(split, lo, ctx);  // issued first: its round trip overlaps the fetch of the query vector
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int i = tid; i < G * kHeadDim / 4; i += 128)
    reinterpret_cast<float4*>(sm.qs)[i] = __ldcg(reinterpret_cast<const float4*>(a.q + (size_t)kvh * G * kHeadDim) + i);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  float acc[G];
#pragma unroll
  for (int h = 0; h < G; ++h) acc[h] = 0.f;
  __syncthreads();

  for (int tile = split; tile * kTile < ctx; tile += S) {
    const int t0 = tile * kTile;
    if (tile != split) {  // contexts beyond 32 * n_splits tokens: further tiles, fetched in place
      gather(tile, 0, ctx);
      cp_async_commit_wait_all();
      __syncthreads();
    }
    const int n_valid = ctx - t0 < kTile ? ctx - t0 : kTile;
    for (int h = warp; h < G; h += 4) {  // scores + online softmax: one head per warp, lane = token
      const uint4* kr = reinterpret_cast<const uint4*>(sm.Ks + lane * kHeadDim);
      const float4* qh = reinterpret_cast<const float4*>(sm.qs + h * kHeadDim);
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const uint4 k = kr[c ^ (lane & 7)];
        const float4 qa = qh[2 * c], qb = qh[2 * c + 1];
        s0 = fmaf(qa.x, bf16lo(k.x), s0);
        s1 = fmaf(qa.y, bf16hi(k.x), s1);
        s2 = fmaf(qa.z, bf16lo(k.y), s2);
        s3 = fmaf(qa.w, bf16hi(k.y), s3);
        s0 = fmaf(qb.x, bf16lo(k.z), s0);
        s1 = fmaf(qb.y, bf16hi(k.z), s1);
        s2 = fmaf(qb.z, bf16lo(k.w), s2);
        s3 = fmaf(qb.w, bf16hi(k.w), s3);
      }
      float sc = (s0 + s1) + (s2 + s3);
      const bool valid = lane < n_valid;
      sc = valid ? sc * scale : -INFINITY;
      const float m_old = sm.m_run[h];
      const float m_new = fmaxf(m_old, warp_max(sc));
      const float p = valid ? expf(sc - m_new) : 0.f;
      const float l_tile = warp_sum(p);
      sm.P[h][lane] = p;
      if (lane == 0) {
        const float corr = (m_old == -INFINITY) ? 0.f : expf(m_old - m_new);
        sm.corr_s[h] = corr;
        sm.l_run[h] = sm.l_run[h] * corr + l_tile;
        sm.m_run[h] = m_new;
      }
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < G; ++h) acc[h] *= sm.corr_s[h];
#pragma unroll 8
    for (int t = 0; t < kTile; ++t) {  // P.V: thread = output dim.  Fixed trip count so the loads pipeline; rows past the
      float v = __uint_as_float((uint32_t)sm.Vs[t * kHeadDim + tid] << 16);  // context hold older data (P is 0 there):
      v = t < n_valid ? v : 0.f;                                             // forced to 0 so nothing odd can leak in
#pragma unroll
      for (int h = 0; h < G; ++h) acc[h] = fmaf(sm.P[h][t], v, acc[h]);
    }
    __syncthreads();
  }

  if (tracer) a.trace[2] = global_timer_ns();
  // ---- split merge: as in attn_decode_item (partials in scratch, the last CTA of a KV head combines them)
  float* part = a.partial + ((size_t)kvh * S + split) * G * kPart;
#pragma unroll
  for (int h = 0; h < G; ++h) part[h * kPart + tid] = acc[h];
  if (tid < G) {
    part[tid * kPart + kHeadDim] = sm.m_run[tid];
    part[tid * kPart + kHeadDim + 1] = sm.l_run[tid];
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) sm.s_last = (atomicAdd(a.counters + kvh, 1u) == (unsigned)(S - 1));
  __syncthreads();
  if (!sm.s_last) return;
  __threadfence();
  const float* base = a.partial + (size_t)kvh * S * G * kPart;
  float* w_s = reinterpret_cast<float*>(sm.Ks);  // [G][kMaxSplits]
  for (int h = warp; h < G; h += 4) {
    const bool on = lane < S;
    const float* ps = base + ((size_t)(on ? lane : 0) * G + h) * kPart;
    const float ms = on ? __ldcg(ps + kHeadDim) : -INFINITY;
    const float ls = on ? __ldcg(ps + kHeadDim + 1) : 0.f;
    const float M = warp_max(ms);
    const float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
    const float den = warp_sum(ls * w);
    w_s[h * kMaxSplits + lane] = w;
    if (lane == 0) sm.corr_s[h] = 1.0f / den;
  }
  __syncthreads();
  float num[G];
#pragma unroll
  for (int h = 0; h < G; ++h) num[h] = 0.f;
  for (int sp = 0; sp < S; sp += 4) {
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
  for (int h = 0; h < G; ++h) a.out[((size_t)kvh * G + h) * kHeadDim + tid] = num[h] * sm.corr_s[h];
  if (tid == 0) a.counters[kvh] = 0u;
  if (a.trace != nullptr && tid == 0) atomicMax(a.trace + 3, global_timer_ns());
}

// ---------------------------------------------------------------------------------------------------------
// Cluster variant of the batch-1 decode step: the kCl = 16 splits of one KV head form ONE thread-block cluster, so
// their partial results are combined over distributed shared memory instead of through global scratch, a fence,
// an atomic ticket and two rounds of L2 loads by the last CTA (measured with B200_TRACE: that merge was 5.4 us of
// the kernel's 8 us).  Every CTA keeps (acc[G][128], m[G], l[G]) of its tiles in its own shared memory; after one
// cluster barrier CTA r combines output dims [8r, 8r+8) of all G heads from the 16 CTAs and stores them; a second
// barrier keeps every CTA's shared memory alive until its peers have read it.  Needs no scratch and no counters.
constexpr int kCl = 16;  // cluster size (non-portable: opted in at launch; the launcher falls back if refused)
constexpr int kClPre = 2;  // tiles per CTA held in shared memory: 2 x 16 x 32 = contexts up to 1024 tokens need no fetch after the wait

template <int G>
struct AttnClSmem {
  alignas(16) uint16_t Ks[kClPre][kTile * kHeadDim];  // the first kClPre tiles of this split, both fetched ahead of the dependency
  alignas(16) uint16_t Vs[kClPre][kTile * kHeadDim];
  alignas(16) float qs[G * kHeadDim];
  float P[G][kTile], P2[G][kTile];  // softmax numerators of the pass's first / second tile
  float m_run[G], l_run[G], corr_s[G];
  float mAll[G][kCl], lAll[G][kCl], wAll[G][kCl];
  float part[G][kCl][kHeadDim / kCl];   // [head][source CTA][dim]: written by the peers (this CTA's 8 output dims of every partial)
};

DEVINL uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
DEVINL void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `local_smem_addr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
DEVINL uint32_t map_to_rank(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t a;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(local_smem_addr), "r"(rank));
  return a;
}
DEVINL void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
DEVINL void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
DEVINL void st_dsmem(uint32_t addr, float v) { asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
DEVINL float ld_dsmem(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

template <int G>
__global__ void __launch_bounds__(128) attn_decode_cluster_kernel(const AttnArgs a) {
  __shared__ AttnClSmem<G> sm;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kvh = blockIdx.x, split = (int)cluster_ctarank(), S = kCl;
  // thread = one 16-byte chunk of 4 rows per tile: item j -> row t = (tid + 128 j) / 16 of the tile, chunk c = tid % 16
  const int c = tid & 15;
  auto row_addr = [&](int page, int tok) {
    return (((size_t)page * a.n_kv_heads + kvh) * kPageTokens + (tok % kPageTokens)) * kHeadDim + c * 8;
  };
  auto gather = [&](int buf, int tile, int row_lo, int row_hi) {  // with block-table look-ups (tiles beyond the first kClPre)
    const uint32_t ks_s = smem_u32(sm.Ks[buf]), vs_s = smem_u32(sm.Vs[buf]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = (tid + 128 * j) >> 4, tok = tile * kTile + t;
      if (tok >= row_lo && tok < row_hi) {
        const size_t row = row_addr(a.block_table[tok / kPageTokens], tok);
        cp_async16(ks_s + (uint32_t)(t * kHeadDim * 2 + ((c ^ (t & 7)) << 4)), a.k_cache + row);
        cp_async16(vs_s + (uint32_t)(t * kHeadDim * 2 + (c << 4)), a.v_cache + row);
      }
    }
  };
  const bool tracer = a.trace != nullptr && kvh == 0 && split == 0 && tid == 0;
  if (tracer) a.trace[0] = global_timer_ns();
  cluster_arrive();  // "this CTA has started": waited for just before the merge writes into the peers' shared memory
  // ---- before the dependency: everything that does not depend on the projection kernel still running.  The page of
  // every row this thread will fetch (kept in registers), and an L2 prefetch of those rows, so that the real fetch —
  // issued after the wait, because one of the rows is being written by that kernel — is a single round trip to L2
  // with no block-table look-up in front of it.
  int page[kClPre][4];
#pragma unroll
  for (int i = 0; i < kClPre; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = (tid + 128 * j) >> 4, tok = (split + i * S) * kTile + t;
      page[i][j] = -1;
      if (tok < a.prefetch_tokens) {
        page[i][j] = a.block_table[tok / kPageTokens];
        if ((c & 7) == 0) {  // one prefetch per 128-byte line
          const size_t row = row_addr(page[i][j], tok);
          asm volatile("prefetch.global.L2 [%0];" ::"l"(a.k_cache + row));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(a.v_cache + row));
        }
      }
    }
  if (tid < G) {
    sm.m_run[tid] = -INFINITY;
    sm.l_run[tid] = 0.f;
  }
  pdl_wait();
  pdl_launch_dependents();
  if (tracer) a.trace[1] = global_timer_ns();
  // ---- after the dependency: K/V rows, the query vector and the position, all in flight together
#pragma unroll
  for (int i = 0; i < kClPre; ++i) {
    const uint32_t ks_s = smem_u32(sm.Ks[i]), vs_s = smem_u32(sm.Vs[i]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = (tid + 128 * j) >> 4, tok = (split + i * S) * kTile + t;
      if (page[i][j] >= 0) {  // rows at or beyond the context hold older data: fetched anyway (no need to know the position yet), masked below
        const size_t row = row_addr(page[i][j], tok);
        cp_async16(ks_s + (uint32_t)(t * kHeadDim * 2 + ((c ^ (t & 7)) << 4)), a.k_cache + row);
        cp_async16(vs_s + (uint32_t)(t * kHeadDim * 2 + (c << 4)), a.v_cache + row);
      }
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int i = tid; i < G * kHeadDim / 4; i += 128)
    reinterpret_cast<float4*>(sm.qs)[i] = __ldcg(reinterpret_cast<const float4*>(a.q + (size_t)kvh * G * kHeadDim) + i);
  const int pos = a.fixed_ctx > 0 ? a.fixed_ctx - 1 : __ldcg(&a.st->pos);
  const int ctx = pos + 1;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  float acc[G];
#pragma unroll
  for (int h = 0; h < G; ++h) acc[h] = 0.f;
  __syncthreads();

  // Two tiles per pass (this CTA's tiles split + 2pS and split + (2p+1)S, 64 tokens): their scores run as independent
  // chains in the same loop and share one softmax update and one set of barriers.  Pass 0 uses the rows fetched above.
  static_assert(kClPre == 2, "a pass covers the two prefetched tiles");
  for (int pass = 0; (split + 2 * pass * S) * kTile < ctx; ++pass) {
    const int tileA = split + 2 * pass * S, tileB = tileA + S;
    if (pass > 0) {  // contexts beyond 1024 tokens: fetched in place
      gather(0, tileA, 0, ctx);
      gather(1, tileB, 0, ctx);
      cp_async_commit_wait_all();
      __syncthreads();
    }
    const int nA = min(ctx - tileA * kTile, kTile), nB = max(0, min(ctx - tileB * kTile, kTile));
    for (int h = warp; h < G; h += 4) {
      const uint4* ka = reinterpret_cast<const uint4*>(sm.Ks[0] + lane * kHeadDim);
      const uint4* kb = reinterpret_cast<const uint4*>(sm.Ks[1] + lane * kHeadDim);
      const float4* qh = reinterpret_cast<const float4*>(sm.qs + h * kHeadDim);
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const uint4 k = ka[c ^ (lane & 7)], j = kb[c ^ (lane & 7)];
        const float4 qa = qh[2 * c], qb = qh[2 * c + 1];
        a0 = fmaf(qa.x, bf16lo(k.x), a0); b0 = fmaf(qa.x, bf16lo(j.x), b0);
        a1 = fmaf(qa.y, bf16hi(k.x), a1); b1 = fmaf(qa.y, bf16hi(j.x), b1);
        a2 = fmaf(qa.z, bf16lo(k.y), a2); b2 = fmaf(qa.z, bf16lo(j.y), b2);
        a3 = fmaf(qa.w, bf16hi(k.y), a3); b3 = fmaf(qa.w, bf16hi(j.y), b3);
        a0 = fmaf(qb.x, bf16lo(k.z), a0); b0 = fmaf(qb.x, bf16lo(j.z), b0);
        a1 = fmaf(qb.y, bf16hi(k.z), a1); b1 = fmaf(qb.y, bf16hi(j.z), b1);
        a2 = fmaf(qb.z, bf16lo(k.w), a2); b2 = fmaf(qb.z, bf16lo(j.w), b2);
        a3 = fmaf(qb.w, bf16hi(k.w), a3); b3 = fmaf(qb.w, bf16hi(j.w), b3);
      }
      const bool va = lane < nA, vb = lane < nB;
      const float sa = va ? ((a0 + a1) + (a2 + a3)) * scale : -INFINITY;
      const float sb = vb ? ((b0 + b1) + (b2 + b3)) * scale : -INFINITY;
      const float m_old = sm.m_run[h];
      const float m_new = fmaxf(m_old, warp_max(fmaxf(sa, sb)));
      const float pa = va ? expf(sa - m_new) : 0.f, pb = vb ? expf(sb - m_new) : 0.f;
      const float l_tile = warp_sum(pa + pb);
      sm.P[h][lane] = pa;
      sm.P2[h][lane] = pb;
      if (lane == 0) {
        const float corr = (m_old == -INFINITY) ? 0.f : expf(m_old - m_new);
        sm.corr_s[h] = corr;
        sm.l_run[h] = sm.l_run[h] * corr + l_tile;
        sm.m_run[h] = m_new;
      }
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < G; ++h) acc[h] *= sm.corr_s[h];
#pragma unroll 8
    for (int t = 0; t < kTile; ++t) {  // fixed trip count (loads pipeline); rows past the context are forced to 0
      float v = __uint_as_float((uint32_t)sm.Vs[0][t * kHeadDim + tid] << 16);
      v = t < nA ? v : 0.f;
#pragma unroll
      for (int h = 0; h < G; ++h) acc[h] = fmaf(sm.P[h][t], v, acc[h]);
    }
    if (nB > 0) {
#pragma unroll 8
      for (int t = 0; t < kTile; ++t) {
        float v = __uint_as_float((uint32_t)sm.Vs[1][t * kHeadDim + tid] << 16);
        v = t < nB ? v : 0.f;
#pragma unroll
        for (int h = 0; h < G; ++h) acc[h] = fmaf(sm.P2[h][t], v, acc[h]);
      }
    }
    __syncthreads();
  }
  if (tracer) a.trace[2] = global_timer_ns();

  // ---- merge over the cluster: every CTA PUSHES what its peers need into their shared memory — CTA r combines output
  // dims [8r, 8r+8) of all heads, so dim `tid` of this partial goes to CTA tid / 8, and every CTA gets this split's
  // running max and sum — then ONE cluster barrier, after which each CTA works on local memory only and may exit
  // without a second barrier (a pull needed one more: peers could still be reading this CTA's memory).
  constexpr int kSlice = kHeadDim / kCl;  // 8 output dims per CTA
  cluster_wait();  // every CTA of the cluster has started (arrived at entry): its shared memory exists
  {
    const uint32_t dst = (uint32_t)(tid / kSlice);
#pragma unroll
    for (int h = 0; h < G; ++h) st_dsmem(map_to_rank(smem_u32(&sm.part[h][split][tid % kSlice]), dst), acc[h]);
    if (tid < G * kCl) {  // (head, destination CTA)
      const int h = tid / kCl;
      const uint32_t to = (uint32_t)(tid % kCl);
      st_dsmem(map_to_rank(smem_u32(&sm.mAll[h][split]), to), sm.m_run[h]);
      st_dsmem(map_to_rank(smem_u32(&sm.lAll[h][split]), to), sm.l_run[h]);
    }
  }
  cluster_sync_all();  // all pushes have landed
  if (tid < G * kCl) {  // softmax weight of every source for every head
    const int h = tid / kCl, s_ = tid % kCl;
    float M = -INFINITY;
#pragma unroll
    for (int j = 0; j < kCl; ++j) M = fmaxf(M, sm.mAll[h][j]);
    const float ms = sm.mAll[h][s_];
    sm.wAll[h][s_] = (ms == -INFINITY) ? 0.f : expf(ms - M);
  }
  __syncthreads();
  if (tid < G * kSlice) {
    const int h = tid / kSlice, dd = tid % kSlice;
    float num = 0.f, den = 0.f;
#pragma unroll
    for (int j = 0; j < kCl; ++j) {
      num += __fmul_rn(sm.part[h][j][dd], sm.wAll[h][j]);  // product rounded, then added: as when the product was stored first
      den = fmaf(sm.lAll[h][j], sm.wAll[h][j], den);
    }
    a.out[((size_t)kvh * G + h) * kHeadDim + split * kSlice + dd] = num / den;
  }
  if (a.trace != nullptr && tid == 0) atomicMax(a.trace + 3, global_timer_ns());
}

template <int G>
cudaError_t launch_cluster_t(const AttnArgs& a, cudaStream_t s, bool pdl) {
  auto kern = attn_decode_cluster_kernel<G>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a.n_kv_heads, kCl, 1);
  cfg.blockDim = dim3(128);
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = kCl; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 2 : 1;
  return count_launch(), cudaLaunchKernelEx(&cfg, kern, a);
}

// Can a 16-CTA cluster of this kernel be scheduled on the current device at all?  (asked once per group size)
template <int G>
bool cluster_ok_t() {
  auto kern = attn_decode_cluster_kernel<G>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { (void)cudaGetLastError(); return false; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(8, kCl, 1);
  cfg.blockDim = dim3(128);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = kCl; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { (void)cudaGetLastError(); return false; }
  return n >= 1;
}

}  // namespace

bool attn_cluster_supported(int group) {
  switch (group) {
    case 1: return cluster_ok_t<1>();
    case 2: return cluster_ok_t<2>();
    case 4: return cluster_ok_t<4>();
    case 8: return cluster_ok_t<8>();
  }
  return false;
}

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
  // the batch-1 decode step (one query, the engine's own block table, cache rows readable ahead of the dependency):
  // the cluster kernel when the engine asked for it (a.use_cluster; it checked attn_cluster_supported first)
  if (a.prefetch_tokens > 0 && a.use_cluster && a.n_queries <= 1 && !a.q_ctx && !a.q_slot) {
    switch (a.n_heads / a.n_kv_heads) {
      case 1: return launch_cluster_t<1>(a, s, pdl);
      case 2: return launch_cluster_t<2>(a, s, pdl);
      case 4: return launch_cluster_t<4>(a, s, pdl);
      case 8: return launch_cluster_t<8>(a, s, pdl);
    }
    return cudaErrorInvalidValue;
  }
  if (a.prefetch_tokens > 0 && a.n_queries <= 1 && !a.q_ctx && !a.q_slot) {
    switch (a.n_heads / a.n_kv_heads) {
      case 1: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_fast_kernel<1>, a);
      case 2: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_fast_kernel<2>, a);
      case 4: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_fast_kernel<4>, a);
      case 8: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_fast_kernel<8>, a);
    }
    return cudaErrorInvalidValue;
  }
  switch (a.n_heads / a.n_kv_heads) {
    case 1: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_kernel<1>, a);
    case 2: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_kernel<2>, a);
    case 4: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_kernel<4>, a);
    case 8: return count_launch(), cudaLaunchKernelEx(&cfg, attn_decode_kernel<8>, a);
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
