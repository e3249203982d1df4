// attn_decode.cuh — the split-KV decode attention work item (see attn.cu), used by the stand-alone kernel
// (a persistent-kernel experiment that also used it now lives in scratch/decode_mega.cu).
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace b200 {
namespace {

DEVINL void bar_sync128(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

constexpr int kTile = 32;
constexpr int kKStride = 65;  // words per K row in smem (64 data + 1 pad): lane=token reads are conflict-free
constexpr int kPart = kHeadDim + 2;  // acc[128], m, l
constexpr int kMaxSplits = 32;

// shared-memory working set of one (KV head, split) work item: 128 threads
template <int G>
struct AttnSmem {
  uint32_t Ks[kTile * kKStride];
  alignas(16) uint16_t Vs[kTile * kHeadDim];
  alignas(16) float qs[G * kHeadDim];
  float P[G][kTile];
  float m_run[G], l_run[G], corr_s[G];
  int s_last;
};

// One work item: KV head `kvh`, tiles split, split+S, ... of query `qi`; executed by 128 threads (tid 0..127)
// that own the barrier `bar_id` (bar.sync bar_id, 128).  n_kvh_grid = KV heads per query in the scratch layout.
template <int G>
DEVINL void attn_decode_item(const AttnArgs& a, AttnSmem<G>& sm, int kvh, int split, int S, int qi, int n_kvh_grid, int tid, int bar_id) {
  auto& Ks = sm.Ks; auto& Vs = sm.Vs; auto& qs = sm.qs; auto& P = sm.P;
  auto& m_run = sm.m_run; auto& l_run = sm.l_run; auto& corr_s = sm.corr_s; auto& s_last = sm.s_last;
  const int warp = tid >> 5, lane = tid & 31;
  const int ctx = a.q_ctx ? a.q_ctx[qi] : (a.fixed_ctx > 0 ? a.fixed_ctx + qi : a.st->pos + 1);
  const int32_t* bt = a.q_slot ? a.block_table + (size_t)a.q_slot[qi] * a.pages_per_seq : a.block_table;
  const float scale = 1.0f / sqrtf((float)kHeadDim);
  const float* q_in = a.q + (size_t)qi * a.q_stride;
  float* out = a.out + (size_t)qi * a.q_stride;
  float* partial = a.partial + (size_t)qi * n_kvh_grid * S * G * kPart;
  unsigned int* counters = a.counters + (size_t)qi * n_kvh_grid;

  for (int i = tid; i < G * kHeadDim; i += 128) qs[i] = __ldcg(q_in + (size_t)kvh * G * kHeadDim + i);
  if (tid < G) {
    m_run[tid] = -INFINITY;
    l_run[tid] = 0.f;
  }
  float acc[G];
#pragma unroll
  for (int h = 0; h < G; ++h) acc[h] = 0.f;
  bar_sync128(bar_id);

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
    bar_sync128(bar_id);
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
    bar_sync128(bar_id);
    // P.V: thread = output dim
#pragma unroll
    for (int h = 0; h < G; ++h) acc[h] *= corr_s[h];
#pragma unroll 8
    for (int t = 0; t < kTile; ++t) {
      const float v = __uint_as_float((uint32_t)Vs[t * kHeadDim + tid] << 16);
#pragma unroll
      for (int h = 0; h < G; ++h) acc[h] = fmaf(P[h][t], v, acc[h]);
    }
    bar_sync128(bar_id);
  }

  float* part = partial + ((size_t)kvh * S + split) * G * kPart;
#pragma unroll
  for (int h = 0; h < G; ++h) part[h * kPart + tid] = acc[h];
  if (tid < G) {
    part[tid * kPart + kHeadDim] = m_run[tid];
    part[tid * kPart + kHeadDim + 1] = l_run[tid];
  }
  __threadfence();
  bar_sync128(bar_id);
  if (tid == 0) s_last = (atomicAdd(counters + kvh, 1u) == (unsigned)(S - 1));
  bar_sync128(bar_id);
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
  bar_sync128(bar_id);
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
}  // namespace b200
