// prefill.cu — the element-wise kernels between the tensor-core GEMMs of a prefill / batched pass
// (GEMM numerics, oracle mode 1): token embedding gather, (residual add +) RMSNorm -> bf16, RoPE + KV-cache
// append, SwiGLU -> bf16, and the hand-over of the last position's residual to the decode path.
// All are HBM/L2-bound streaming kernels over [tokens x width] fp32; the bf16 rounding of GEMM inputs
// happens exactly once, here, with the same RNE recipe as the oracle.
#include <algorithm>

#include "common.cuh"
#include "kernels.h"

namespace b200 {
namespace {

DEVINL float block_sum256(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) t += red[i];
  __syncthreads();
  return t;
}

DEVINL float block_sum_any(float v, float* red) {  // any block size that is a multiple of 32, red[32]
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.f;
  for (int i = 0; i < nw; ++i) t += red[i];
  __syncthreads();
  return t;
}

// x[t][:] (+= sum of the split-K partials add[p][t][:], p ascending) ; xb[t][:] = bf16( (x * inv_rms) * w )
// grid = tokens, block = 256 .. 1024 (a multiple of 32)
template <int THREADS, int kMaxI, int MINB>  // <1024, 2, 1>: few rows (batched decode: latency-bound)
__global__ void __launch_bounds__(THREADS, MINB) add_rmsnorm_bf16_kernel(float* x, const float* add, int n_parts, long long part_stride, const float* w, uint16_t* xb, int d,
                                        float eps, const unsigned int* wait_flags, int n_wait, unsigned int wait_seq, int* wait_error) {
  __shared__ float red[32];
  if (wait_flags) {  // partials pushed by the peer GPUs: wait for every rank's sequence flag (bounded)
    if ((int)threadIdx.x < n_wait) {
      const long long t0 = clock64();
      unsigned int v;
      do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(wait_flags + threadIdx.x) : "memory");
        if (clock64() - t0 > 4000000000ll) { *wait_error = 1; break; }
      } while ((int)(v - wait_seq) < 0);
    }
    __syncthreads();
  }
  float* xr = x + (size_t)blockIdx.x * d;
  const float* ar = (add && n_parts > 0) ? add + (size_t)blockIdx.x * d : nullptr;
  // Each thread owns up to kMaxI float4 of the row and keeps them (and the norm weights) in registers between the sum
  // of squares and the scaling: with 1024 threads a row of 4096 is ONE float4 per thread, so the residual, the partials
  // (summed in ascending order, fetched eight at a time) and the weights are all in flight together.  With few rows
  // (batched decode) this kernel is a chain of L2 round trips, not bandwidth.
  const int step = (int)blockDim.x * 4;
  float4 v[kMaxI], g[kMaxI];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxI; ++i) {
    const int k = (int)threadIdx.x * 4 + i * step;
    if (k >= d) break;
    v[i] = *reinterpret_cast<const float4*>(xr + k);
    g[i] = *reinterpret_cast<const float4*>(w + k);
    if (ar) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int p0 = 0; p0 < n_parts; p0 += 8) {
        float4 b[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          b[j] = p0 + j < n_parts ? __ldcg(reinterpret_cast<const float4*>(ar + (size_t)(p0 + j) * part_stride + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (p0 + j < n_parts) {
            if (p0 + j == 0) a = b[0];
            else { a.x += b[j].x; a.y += b[j].y; a.z += b[j].z; a.w += b[j].w; }
          }
      }
      v[i].x += a.x; v[i].y += a.y; v[i].z += a.z; v[i].w += a.w;
      *reinterpret_cast<float4*>(xr + k) = v[i];
    }
    ss += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
  }
  for (int k = (int)threadIdx.x * 4 + kMaxI * step; k < d; k += step) {  // rows wider than 4 x kMaxI x threads: not held in registers
    float4 t = *reinterpret_cast<const float4*>(xr + k);
    if (ar) {
      float4 a = __ldcg(reinterpret_cast<const float4*>(ar + k));
      for (int p = 1; p < n_parts; ++p) {
        const float4 b = __ldcg(reinterpret_cast<const float4*>(ar + (size_t)p * part_stride + k));
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      }
      t.x += a.x; t.y += a.y; t.z += a.z; t.w += a.w;
      *reinterpret_cast<float4*>(xr + k) = t;
    }
    ss += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
  }
  ss = block_sum_any(ss, red);
  const float inv = 1.0f / sqrtf(ss / (float)d + eps);
  uint16_t* o = xb + (size_t)blockIdx.x * d;
  auto put = [&](int k, const float4& t, const float4& gg) {
    uint2 p;
    p.x = (uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(t.x, inv), gg.x)) | ((uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(t.y, inv), gg.y)) << 16);
    p.y = (uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(t.z, inv), gg.z)) | ((uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(t.w, inv), gg.w)) << 16);
    *reinterpret_cast<uint2*>(o + k) = p;
  };
#pragma unroll
  for (int i = 0; i < kMaxI; ++i) {
    const int k = (int)threadIdx.x * 4 + i * step;
    if (k >= d) break;
    put(k, v[i], g[i]);
  }
  for (int k = (int)threadIdx.x * 4 + kMaxI * step; k < d; k += step)
    put(k, *reinterpret_cast<const float4*>(xr + k), *reinterpret_cast<const float4*>(w + k));
}

// qkv[t][q_rows + 2 kv_rows] fp32 -> q[t][q_rows] (rotated), K/V appended to the paged bf16 cache at pos0+t
// Row t is position pos0+t of one sequence (prefill), or — batched decode — position pos_arr[t] of the
// sequence in slot slot_arr[t], whose block table is row slot_arr[t] of `block_table` (pages_per_seq wide).
__global__ void rope_kv_kernel(const float* qkv, float* q, uint16_t* k_cache, uint16_t* v_cache, const int32_t* block_table,
                               const float* rope, int q_rows, int kv_rows, int n_kv_heads, int pos0, const int32_t* pos_arr,
                               const int32_t* slot_arr, int pages_per_seq, int n_parts, long long part_stride) {
  const int t = blockIdx.x, pos = pos_arr ? pos_arr[t] : pos0 + t;
  const int width = q_rows + 2 * kv_rows;
  const float* in = qkv + (size_t)t * width;
  const int32_t* bt = slot_arr ? block_table + (size_t)slot_arr[t] * pages_per_seq : block_table;
  const int page = bt[pos / kPageTokens];
  for (int p = blockIdx.y * blockDim.x + threadIdx.x; p < width / 2; p += gridDim.y * blockDim.x) {
    const int row = 2 * p;
    float x0 = 0.f, x1 = 0.f;
    for (int p0 = 0; p0 < n_parts; p0 += 8) {  // split-K partials of the qkv GEMM: summed in ascending order, fetched 8 at a time
      float2 b[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        b[j] = p0 + j < n_parts ? __ldcg(reinterpret_cast<const float2*>(in + (size_t)(p0 + j) * part_stride + row)) : make_float2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (p0 + j < n_parts) {
          if (p0 + j == 0) { x0 = b[0].x; x1 = b[0].y; }
          else { x0 += b[j].x; x1 += b[j].y; }
        }
    }
    if (row < q_rows + kv_rows) {
      const int i = (row & (kHeadDim - 1)) >> 1;
      const float2 cs = *reinterpret_cast<const float2*>(rope + ((size_t)pos * (kHeadDim / 2) + i) * 2);
      const float r0 = __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y));
      const float r1 = __fadd_rn(__fmul_rn(x0, cs.y), __fmul_rn(x1, cs.x));
      x0 = r0;
      x1 = r1;
    }
    if (row < q_rows) {
      *reinterpret_cast<float2*>(q + (size_t)t * q_rows + row) = make_float2(x0, x1);
    } else {
      const bool isk = row < q_rows + kv_rows;
      const int kr = row - q_rows - (isk ? 0 : kv_rows);
      const int kvh = kr >> 7, dim = kr & (kHeadDim - 1);
      const size_t idx = (((size_t)page * n_kv_heads + kvh) * kPageTokens + (pos % kPageTokens)) * kHeadDim + dim;
      *reinterpret_cast<uint32_t*>((isk ? k_cache : v_cache) + idx) = (uint32_t)f32_to_bf16_bits(x0) | ((uint32_t)f32_to_bf16_bits(x1) << 16);
    }
  }
}

// The same operation for MANY rows (prefill): plenty of CTAs in flight, so the plain streaming form — 256 threads, the
// row re-read from L1/L2 for the scaling pass — is the faster one (8.8 vs 11.6 us at 512 rows).  grid = tokens, block = 256
__global__ void add_rmsnorm_bf16_rows_kernel(float* x, const float* add, int n_parts, long long part_stride, const float* w, uint16_t* xb, int d,
                                        float eps, const unsigned int* wait_flags, int n_wait, unsigned int wait_seq, int* wait_error) {
  __shared__ float red[8];
  if (wait_flags) {  // partials pushed by the peer GPUs: wait for every rank's sequence flag (bounded)
    if ((int)threadIdx.x < n_wait) {
      const long long t0 = clock64();
      unsigned int v;
      do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(wait_flags + threadIdx.x) : "memory");
        if (clock64() - t0 > 4000000000ll) { *wait_error = 1; break; }
      } while ((int)(v - wait_seq) < 0);
    }
    __syncthreads();
  }
  float* xr = x + (size_t)blockIdx.x * d;
  const float* ar = (add && n_parts > 0) ? add + (size_t)blockIdx.x * d : nullptr;
  float ss = 0.f;
  for (int k = threadIdx.x * 4; k < d; k += 1024) {
    float4 v = *reinterpret_cast<const float4*>(xr + k);
    if (ar) {
      float4 a = __ldcg(reinterpret_cast<const float4*>(ar + k));
      for (int p = 1; p < n_parts; ++p) {
        const float4 b = __ldcg(reinterpret_cast<const float4*>(ar + (size_t)p * part_stride + k));
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      }
      v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
      *reinterpret_cast<float4*>(xr + k) = v;
    }
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = block_sum256(ss, red);
  const float inv = 1.0f / sqrtf(ss / (float)d + eps);
  uint16_t* o = xb + (size_t)blockIdx.x * d;
  for (int k = threadIdx.x * 4; k < d; k += 1024) {
    const float4 v = *reinterpret_cast<const float4*>(xr + k);
    const float4 g = *reinterpret_cast<const float4*>(w + k);
    uint2 p;
    p.x = (uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(v.x, inv), g.x)) | ((uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(v.y, inv), g.y)) << 16);
    p.y = (uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(v.z, inv), g.z)) | ((uint32_t)f32_to_bf16_bits(__fmul_rn(__fmul_rn(v.w, inv), g.w)) << 16);
    *reinterpret_cast<uint2*>(o + k) = p;
  }
}

// Many rows (prefill): the plain loop over the partials (10 vs 17 us at 512 rows with the batched fetch below).
// qkv[t][q_rows + 2 kv_rows] fp32 -> q[t][q_rows] (rotated), K/V appended to the paged bf16 cache at pos0+t
// Row t is position pos0+t of one sequence (prefill), or — batched decode — position pos_arr[t] of the
// sequence in slot slot_arr[t], whose block table is row slot_arr[t] of `block_table` (pages_per_seq wide).
__global__ void rope_kv_rows_kernel(const float* qkv, float* q, uint16_t* k_cache, uint16_t* v_cache, const int32_t* block_table,
                               const float* rope, int q_rows, int kv_rows, int n_kv_heads, int pos0, const int32_t* pos_arr,
                               const int32_t* slot_arr, int pages_per_seq, int n_parts, long long part_stride) {
  const int t = blockIdx.x, pos = pos_arr ? pos_arr[t] : pos0 + t;
  const int width = q_rows + 2 * kv_rows;
  const float* in = qkv + (size_t)t * width;
  const int32_t* bt = slot_arr ? block_table + (size_t)slot_arr[t] * pages_per_seq : block_table;
  const int page = bt[pos / kPageTokens];
  for (int p = blockIdx.y * blockDim.x + threadIdx.x; p < width / 2; p += gridDim.y * blockDim.x) {
    const int row = 2 * p;
    float x0 = in[row], x1 = in[row + 1];
    for (int sp = 1; sp < n_parts; ++sp) {  // split-K partials of the qkv GEMM, ascending order
      x0 += in[(size_t)sp * part_stride + row];
      x1 += in[(size_t)sp * part_stride + row + 1];
    }
    if (row < q_rows + kv_rows) {
      const int i = (row & (kHeadDim - 1)) >> 1;
      const float2 cs = *reinterpret_cast<const float2*>(rope + ((size_t)pos * (kHeadDim / 2) + i) * 2);
      const float r0 = __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y));
      const float r1 = __fadd_rn(__fmul_rn(x0, cs.y), __fmul_rn(x1, cs.x));
      x0 = r0;
      x1 = r1;
    }
    if (row < q_rows) {
      *reinterpret_cast<float2*>(q + (size_t)t * q_rows + row) = make_float2(x0, x1);
    } else {
      const bool isk = row < q_rows + kv_rows;
      const int kr = row - q_rows - (isk ? 0 : kv_rows);
      const int kvh = kr >> 7, dim = kr & (kHeadDim - 1);
      const size_t idx = (((size_t)page * n_kv_heads + kvh) * kPageTokens + (pos % kPageTokens)) * kHeadDim + dim;
      *reinterpret_cast<uint32_t*>((isk ? k_cache : v_cache) + idx) = (uint32_t)f32_to_bf16_bits(x0) | ((uint32_t)f32_to_bf16_bits(x1) << 16);
    }
  }
}

// gu[t][2 ff] (rows interleaved gate_i, up_i) -> h[t][ff] = bf16( silu(g) * u )
// (gu: n_parts split-K partials, part_stride floats apart, summed in ascending order)
__global__ void swiglu_bf16_kernel(const float* gu, uint16_t* h, int64_t n_pairs, int n_parts, long long part_stride) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += stride) {
    float2 v = *reinterpret_cast<const float2*>(gu + 2 * i);
    for (int sp = 1; sp < n_parts; ++sp) {
      const float2 w = *reinterpret_cast<const float2*>(gu + (size_t)sp * part_stride + 2 * i);
      v.x += w.x; v.y += w.y;
    }
    h[i] = f32_to_bf16_bits(silu_mul(v.x, v.y));  // as the GEMM epilogue (gemm.cu), so every schedule gives the same h
  }
}

// out[:] = x[row][:] + sum of the partials add[p][row][:]   (residual of the last prompt position -> the
// decode path's stream)
__global__ void take_row_kernel(const float* x, const float* add, int n_parts, long long part_stride, int d, int row, float* out) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < d; k += gridDim.x * blockDim.x) {
    float a = add[(size_t)row * d + k];
    for (int p = 1; p < n_parts; ++p) a += add[(size_t)p * part_stride + (size_t)row * d + k];
    out[k] = x[(size_t)row * d + k] + a;
  }
}

// next[row] = argmax of logits[row][:] (lowest id on ties); one CTA per row
// keys != null (vocab-parallel lm_head): instead leave the packed (logit, ~global id) key of the row's maximum,
// id = id_base + local index, for a max all-reduce over the ranks
__global__ void argmax_rows_kernel(const float* logits, int vocab, int32_t* next, int id_base, unsigned long long* keys) {
  __shared__ unsigned long long red[32];
  const float* lg = logits + (size_t)blockIdx.x * vocab;
  unsigned long long best = 0;
  for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
    const unsigned long long k = argmax_key(lg[i], (uint32_t)(id_base + i));
    best = k > best ? k : best;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
    best = other > best ? other : best;
  }
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) best = red[w] > best ? red[w] : best;
    if (keys) keys[blockIdx.x] = best;
    else next[blockIdx.x] = (int32_t)argmax_key_index(best);
  }
}

__global__ void keys_to_ids_kernel(const unsigned long long* keys, int32_t* next, int rows) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rows) next[r] = (int32_t)argmax_key_index(keys[r]);
}

// history[slot][pos+1] = next token of every row (the repeat penalty of sampled requests reads it)
__global__ void append_history_kernel(const int32_t* next, const int32_t* slot_arr, const int32_t* pos_arr, int32_t* history, int max_ctx, int rows) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rows && pos_arr[r] + 1 < max_ctx) history[(size_t)slot_arr[r] * max_ctx + pos_arr[r] + 1] = next[r];
}

// this rank's partial [count floats] -> slot `rank` of every peer's receive buffer `buf`; the last CTA to finish
// raises this rank's sequence flag on every peer
__global__ void tp_push_kernel(const TpBatch b, const float* src, long long count, int buf, unsigned int seq) {
  const long long n4 = count / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = __ldcg(reinterpret_cast<const float4*>(src) + i);
    for (int p = 0; p < b.size; ++p) reinterpret_cast<float4*>(b.peer_recv[buf][p] + (size_t)b.rank * b.slot_floats)[i] = v;
  }
  __shared__ int last_cta;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const bool last = atomicAdd(b.ticket, 1u) == gridDim.x - 1;
    if (last) {
      *b.ticket = 0u;
      __threadfence_system();
    }
    last_cta = last ? 1 : 0;
  }
  __syncthreads();
  if (last_cta && threadIdx.x < b.size)  // one lane per peer (see gemv.cu)
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(b.peer_flags[threadIdx.x] + buf * kMaxTp + b.rank), "r"(seq) : "memory");
}

}  // namespace

cudaError_t launch_tp_push(const TpBatch& b, const float* src, int64_t count, int buf, unsigned int seq, cudaStream_t s) {
  if (count < 4 || count % 4 || count > b.slot_floats || (buf & ~1)) return cudaErrorInvalidValue;
  const int grid = (int)std::min<int64_t>(296, (count / 4 + 255) / 256);  // one float4 x `size` remote stores per thread
  count_launch(); tp_push_kernel<<<grid, 256, 0, s>>>(b, src, (long long)count, buf, seq);
  return cudaGetLastError();
}

cudaError_t launch_argmax_rows(const float* logits, int rows, int vocab, int32_t* next, cudaStream_t s) {
  count_launch(); argmax_rows_kernel<<<rows, 1024, 0, s>>>(logits, vocab, next, 0, nullptr);
  return cudaGetLastError();
}
cudaError_t launch_argmax_rows_keys(const float* logits, int rows, int vocab, int id_base, unsigned long long* keys, cudaStream_t s) {
  count_launch(); argmax_rows_kernel<<<rows, 1024, 0, s>>>(logits, vocab, nullptr, id_base, keys);
  return cudaGetLastError();
}
cudaError_t launch_keys_to_ids(const unsigned long long* keys, int32_t* next, int rows, cudaStream_t s) {
  count_launch(); keys_to_ids_kernel<<<(rows + 255) / 256, 256, 0, s>>>(keys, next, rows);
  return cudaGetLastError();
}
cudaError_t launch_append_history(const int32_t* next, const int32_t* slot_arr, const int32_t* pos_arr, int32_t* history, int max_ctx, int rows,
                                  cudaStream_t s) {
  count_launch(); append_history_kernel<<<(rows + 255) / 256, 256, 0, s>>>(next, slot_arr, pos_arr, history, max_ctx, rows);
  return cudaGetLastError();
}

cudaError_t launch_add_rmsnorm_bf16(float* x, const float* add, int n_parts, int64_t part_stride, const float* w, uint16_t* xb, int tokens,
                                    int d, float eps, cudaStream_t s, const unsigned int* wait_flags, int n_wait, unsigned int wait_seq,
                                    int* wait_error) {
  if (d % 4 || part_stride % 4 || (wait_flags && (n_wait < 1 || n_wait > 256 || !wait_error))) return cudaErrorInvalidValue;
  count_launch();
  if (tokens <= 64 && d >= 2048)  // few rows: one CTA per row is all the parallelism there is, so make it wide
    add_rmsnorm_bf16_kernel<1024, 2, 1><<<tokens, d >= 4096 ? 1024 : 512, 0, s>>>(x, add, n_parts, (long long)part_stride, w, xb, d, eps, wait_flags, n_wait, wait_seq,
                                                 wait_error);
  else
    add_rmsnorm_bf16_rows_kernel<<<tokens, 256, 0, s>>>(x, add, n_parts, (long long)part_stride, w, xb, d, eps, wait_flags, n_wait, wait_seq,
                                                 wait_error);
  return cudaGetLastError();
}
cudaError_t launch_rope_kv(const float* qkv, float* q, uint16_t* k_cache, uint16_t* v_cache, const int32_t* block_table, const float* rope,
                           int tokens, int q_rows, int kv_rows, int n_kv_heads, int pos0, const int32_t* pos_arr, const int32_t* slot_arr,
                           int pages_per_seq, int n_parts, int64_t part_stride, cudaStream_t s) {
  // few tokens (batched decode): spread a token's row pairs over several CTAs, above all when partials are summed
  const int pairs = (q_rows + 2 * kv_rows) / 2;
  const int full = (pairs + 255) / 256;  // CTAs per token at one pair per thread
  const int ny = tokens >= 256 ? 1 : (tokens >= 64 ? (full < 4 ? full : 4) : full);
  count_launch();
  if (tokens < 64) rope_kv_kernel<<<dim3(tokens, ny), 256, 0, s>>>(qkv, q, k_cache, v_cache, block_table, rope, q_rows, kv_rows, n_kv_heads, pos0, pos_arr, slot_arr,
                                        pages_per_seq, n_parts, (long long)part_stride);
  else rope_kv_rows_kernel<<<dim3(tokens, ny), 256, 0, s>>>(qkv, q, k_cache, v_cache, block_table, rope, q_rows, kv_rows, n_kv_heads, pos0, pos_arr, slot_arr,
                                        pages_per_seq, n_parts, (long long)part_stride);
  return cudaGetLastError();
}
cudaError_t launch_swiglu_bf16(const float* gu, uint16_t* h, int64_t n_pairs, int n_parts, int64_t part_stride, cudaStream_t s) {
  if (part_stride % 2) return cudaErrorInvalidValue;
  const int grid = (int)std::min<int64_t>(1184, (n_pairs + 255) / 256);
  count_launch(); swiglu_bf16_kernel<<<grid, 256, 0, s>>>(gu, h, n_pairs, n_parts, (long long)part_stride);
  return cudaGetLastError();
}
cudaError_t launch_take_row(const float* x, const float* add, int n_parts, int64_t part_stride, int d, int row, float* out, cudaStream_t s) {
  count_launch(); take_row_kernel<<<8, 256, 0, s>>>(x, add, n_parts, (long long)part_stride, d, row, out);
  return cudaGetLastError();
}

}  // namespace b200
