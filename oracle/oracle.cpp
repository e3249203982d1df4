// oracle/oracle.cpp — CPU restatement of the "Suggest a Reply" inference path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under p2p-llm-chat-go_b200/ may include, link or call this file;
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, and
// only as the checker or the timed CPU baseline — never as the shipped path.
//
// What it restates.  The reference (NajyFannoun/P2P-LLM-Chat-Go) contains no arithmetic for this path:
// web/streamlit_app.py:89-101 POSTs a prompt to an external Ollama server (README.md:62-70, untagged docker
// image, floating model tag "llama3.1").  Ollama/llama.cpp is NOT a go.mod dependency, is not vendored, has
// no pinned version, and cannot be installed here (no network, no Go).  This file therefore restates the
// *published* algorithms behind that call:
//   * GGUF block_q4_K / block_q6_K dequantisation (ggml-quants.c: dequantize_row_q4_K / _q6_K)
//   * the Llama-3.1 decoder (RMSNorm, RoPE "llama3" scaling, GQA attention, SwiGLU, untied lm_head)
//   * llama.cpp's default sampler chain order (penalties -> top-k -> top-p -> temperature -> dist)
// PARITY vs Ollama: UNPINNED (the reference holds no test, fixture or golden vector for this path).
// What IS pinned (tests/test_oracle_*.py, fixtures under tests/golden/ made by tests/golden/make_golden.py):
//   * dequantisation: bit-exact against gguf-py 0.19.0 quants.dequantize (numpy reference by the GGUF authors)
//   * whole-model logits: against HuggingFace transformers 5.5.0 LlamaForCausalLM (fp32, llama3 rope scaling)
//
// Numerics contract shared with the CUDA engine (DESIGN.md §"Numerics"):
//   mode 0 ("gemv", batch-1 decode): activations fp32; bf16 weights used exactly; Q4_K/Q6_K weights are
//       multiplied against activations quantised to int16 per 32-element block (d = amax*(1/32767),
//       X = rint(x/d)) with exact integer sub-block dot products and fp32 scale arithmetic.
//   mode 1 ("gemm", prefill / batched decode on tensor cores): GEMM inputs rounded to bf16 (activations,
//       and dequantised Q4_K/Q6_K weights), fp32 accumulate.
//   K (post-RoPE) and V are rounded to bf16 when written to the KV cache (kv_bf16=1).
// The oracle accumulates dot products in double so that it is the more accurate side of every comparison.

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <map>
#include <string>
#include <vector>

#define ORC_API extern "C" __attribute__((visibility("default")))

// ggml tensor type ids (GGUF spec) — the same codes the C-ABI uses.
enum { ORC_F32 = 0, ORC_F16 = 1, ORC_Q4_K = 12, ORC_Q6_K = 14, ORC_BF16 = 30 };
enum { QK_K = 256, Q4K_BYTES = 144, Q6K_BYTES = 210 };

struct orc_config {
  int32_t n_layers, d_model, n_heads, n_kv_heads, head_dim, d_ffn, vocab, max_ctx;
  float rms_eps, rope_theta;
  float rope_factor, rope_low_freq_factor, rope_high_freq_factor;  // llama3 scaling; factor<=1 disables
  int32_t rope_orig_ctx;
  int32_t kv_bf16;  // 1: round K,V to bf16 on cache write (engine behaviour); 0: keep fp32 (HF cross-check)
};

struct orc_sampling {
  float temperature, top_p, repeat_penalty;
  int32_t top_k, repeat_last_n;
  uint64_t seed;
};

namespace {

inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f32_to_bf16(float f) {  // round-to-nearest-even
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
inline float bf16_round(float f) { return bf16_to_f32(f32_to_bf16(f)); }

inline float fp16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000) << 16, e = (h >> 10) & 0x1f, m = h & 0x3ff, u;
  if (e == 0) {
    if (m == 0) u = sign;
    else {  // subnormal: value = m * 2^-24
      float f = (float)m * 5.9604644775390625e-8f;
      memcpy(&u, &f, 4);
      u |= sign;
    }
  } else if (e == 31) u = sign | 0x7f800000u | (m << 13);
  else u = sign | ((e + 112) << 23) | (m << 13);
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// ggml-quants.c get_scale_min_k4: 8 six-bit (scale,min) pairs packed in 12 bytes.
inline void q4k_scale_min(int j, const uint8_t* q, uint8_t* sc, uint8_t* mn) {
  if (j < 4) {
    *sc = q[j] & 63;
    *mn = q[j + 4] & 63;
  } else {
    *sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
    *mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
  }
}

// ggml-quants.c dequantize_row_q4_K: w = (d*sc)*q - (dmin*m), each product rounded to fp32 separately
// (gguf-py quants.py Q4_K.dequantize_blocks does the same in numpy float32).
void dequant_q4k_block(const uint8_t* blk, float* y) {
  uint16_t dh, mh;
  memcpy(&dh, blk, 2);
  memcpy(&mh, blk + 2, 2);
  const float d = fp16_to_f32(dh), dmin = fp16_to_f32(mh);
  const uint8_t* scales = blk + 4;
  const uint8_t* q = blk + 16;
  for (int j = 0; j < 4; ++j) {
    uint8_t sc, m;
    q4k_scale_min(2 * j, scales, &sc, &m);
    const float d1 = d * (float)sc, m1 = dmin * (float)m;
    q4k_scale_min(2 * j + 1, scales, &sc, &m);
    const float d2 = d * (float)sc, m2 = dmin * (float)m;
    for (int l = 0; l < 32; ++l) {
      const float p = d1 * (float)(q[l] & 0xF);
      y[l] = p - m1;
    }
    for (int l = 0; l < 32; ++l) {
      const float p = d2 * (float)(q[l] >> 4);
      y[32 + l] = p - m2;
    }
    q += 32;
    y += 64;
  }
}

// ggml-quants.c dequantize_row_q6_K: w = (d*sc)*(q-32)
void dequant_q6k_block(const uint8_t* blk, float* y) {
  const uint8_t* ql = blk;
  const uint8_t* qh = blk + 128;
  const int8_t* sc = (const int8_t*)(blk + 192);
  uint16_t dh;
  memcpy(&dh, blk + 208, 2);
  const float d = fp16_to_f32(dh);
  for (int n = 0; n < QK_K; n += 128) {
    for (int l = 0; l < 32; ++l) {
      int is = l / 16;
      const int q1 = (int)((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
      const int q2 = (int)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
      const int q3 = (int)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
      const int q4 = (int)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
      const float s0 = d * (float)sc[is + 0], s2 = d * (float)sc[is + 2], s4 = d * (float)sc[is + 4],
                     s6 = d * (float)sc[is + 6];
      y[l + 0] = s0 * (float)q1;
      y[l + 32] = s2 * (float)q2;
      y[l + 64] = s4 * (float)q3;
      y[l + 96] = s6 * (float)q4;
    }
    y += 128;
    ql += 64;
    qh += 32;
    sc += 8;
  }
}

// Activation quantisation "Q16_32": per 32 elements, d = amax * (1/32767), X = rint(x * (1/d)).
void quant_q16(const float* x, int64_t n, int16_t* q, float* d) {
  for (int64_t b = 0; b < n / 32; ++b) {
    float amax = 0.f;
    for (int i = 0; i < 32; ++i) amax = std::max(amax, std::fabs(x[b * 32 + i]));
    const float dd = amax * (1.0f / 32767.0f);
    const float id = dd > 0.f ? 1.0f / dd : 0.f;
    d[b] = dd;
    for (int i = 0; i < 32; ++i) {
      const float s = x[b * 32 + i] * id;
      float r = nearbyintf(s);
      r = std::min(32767.f, std::max(-32767.f, r));
      q[b * 32 + i] = (int16_t)r;
    }
  }
}

size_t row_bytes(int wtype, int64_t cols) {
  switch (wtype) {
    case ORC_F32: return (size_t)cols * 4;
    case ORC_BF16: return (size_t)cols * 2;
    case ORC_F16: return (size_t)cols * 2;
    case ORC_Q4_K: return (size_t)(cols / QK_K) * Q4K_BYTES;
    case ORC_Q6_K: return (size_t)(cols / QK_K) * Q6K_BYTES;
  }
  return 0;
}

void dequant_row(int wtype, const uint8_t* row, int64_t cols, float* out) {
  switch (wtype) {
    case ORC_F32: memcpy(out, row, (size_t)cols * 4); break;
    case ORC_F16:  // IEEE half, as F16 GGUF files store it
      for (int64_t k = 0; k < cols; ++k) {
        uint16_t h;
        memcpy(&h, row + 2 * k, 2);
        out[k] = fp16_to_f32(h);
      }
      break;
    case ORC_BF16:
      for (int64_t k = 0; k < cols; ++k) {
        uint16_t h;
        memcpy(&h, row + 2 * k, 2);
        out[k] = bf16_to_f32(h);
      }
      break;
    case ORC_Q4_K:
      for (int64_t b = 0; b < cols / QK_K; ++b) dequant_q4k_block(row + b * Q4K_BYTES, out + b * QK_K);
      break;
    case ORC_Q6_K:
      for (int64_t b = 0; b < cols / QK_K; ++b) dequant_q6k_block(row + b * Q6K_BYTES, out + b * QK_K);
      break;
  }
}

// One Q4_K row against one Q16-quantised activation vector (mode 0).
double dot_q4k_q16(const uint8_t* row, int64_t cols, const int16_t* X, const float* dx) {
  double acc = 0.0;
  for (int64_t b = 0; b < cols / QK_K; ++b) {
    const uint8_t* blk = row + b * Q4K_BYTES;
    uint16_t dh, mh;
    memcpy(&dh, blk, 2);
    memcpy(&mh, blk + 2, 2);
    const float d = fp16_to_f32(dh), dmin = fp16_to_f32(mh);
    const uint8_t* q = blk + 16;
    for (int j = 0; j < 8; ++j) {
      uint8_t sc, m;
      q4k_scale_min(j, blk + 4, &sc, &m);
      const int16_t* xs = X + b * QK_K + j * 32;
      const uint8_t* qq = q + (j / 2) * 32;
      int32_t isum = 0, xsum = 0;
      if (j & 1) for (int l = 0; l < 32; ++l) { isum += (int32_t)(qq[l] >> 4) * xs[l]; xsum += xs[l]; }
      else       for (int l = 0; l < 32; ++l) { isum += (int32_t)(qq[l] & 0xF) * xs[l]; xsum += xs[l]; }
      const float dxj = dx[b * 8 + j];
      const float a = (d * (float)sc) * (dxj * (float)isum);
      const float c = (dmin * (float)m) * (dxj * (float)xsum);
      acc += (double)a - (double)c;
    }
  }
  return acc;
}

double dot_q6k_q16(const uint8_t* row, int64_t cols, const int16_t* X, const float* dx) {
  double acc = 0.0;
  for (int64_t b = 0; b < cols / QK_K; ++b) {
    const uint8_t* blk = row + b * Q6K_BYTES;
    const uint8_t* ql = blk;
    const uint8_t* qh = blk + 128;
    const int8_t* sc = (const int8_t*)(blk + 192);
    uint16_t dh;
    memcpy(&dh, blk + 208, 2);
    const float d = fp16_to_f32(dh);
    // gather the 256 six-bit values in element order, then 16 sub-blocks of 16
    int8_t qv[QK_K];
    for (int n = 0; n < 2; ++n)
      for (int l = 0; l < 32; ++l) {
        const uint8_t* L = ql + n * 64;
        const uint8_t* H = qh + n * 32;
        qv[n * 128 + l + 0] = (int8_t)(((L[l] & 0xF) | (((H[l] >> 0) & 3) << 4)) - 32);
        qv[n * 128 + l + 32] = (int8_t)(((L[l + 32] & 0xF) | (((H[l] >> 2) & 3) << 4)) - 32);
        qv[n * 128 + l + 64] = (int8_t)(((L[l] >> 4) | (((H[l] >> 4) & 3) << 4)) - 32);
        qv[n * 128 + l + 96] = (int8_t)(((L[l + 32] >> 4) | (((H[l] >> 6) & 3) << 4)) - 32);
      }
    for (int s = 0; s < 16; ++s) {
      int32_t isum = 0;
      const int16_t* xs = X + b * QK_K + s * 16;
      for (int l = 0; l < 16; ++l) isum += (int32_t)qv[s * 16 + l] * xs[l];
      const float a = (d * (float)sc[s]) * (dx[b * 8 + s / 2] * (float)isum);
      acc += (double)a;
    }
  }
  return acc;
}

// Y[t][r] = sum_k W[r][k] X[t][k]  under the numerics contract of `mode`.
void matmul(int wtype, const void* W, int64_t rows, int64_t cols, const float* X, int n, int mode, float* Y,
            int64_t ldy) {
  const size_t rb = row_bytes(wtype, cols);
  const bool quant = (wtype == ORC_Q4_K || wtype == ORC_Q6_K);
  std::vector<int16_t> Xq;
  std::vector<float> dx, Xb;
  if (mode == 0 && quant) {
    Xq.resize((size_t)n * cols);
    dx.resize((size_t)n * cols / 32);
    for (int t = 0; t < n; ++t) quant_q16(X + (size_t)t * cols, cols, Xq.data() + (size_t)t * cols, dx.data() + (size_t)t * cols / 32);
  }
  if (mode == 1) {
    Xb.resize((size_t)n * cols);
    for (size_t i = 0; i < (size_t)n * cols; ++i) Xb[i] = bf16_round(X[i]);
  }
  const float* Xin = (mode == 1) ? Xb.data() : X;
#pragma omp parallel
  {
    std::vector<float> wrow;
    if (quant || wtype == ORC_F32 || wtype == ORC_F16 || n > 1) wrow.resize(cols);
#pragma omp for schedule(static)
    for (int64_t r = 0; r < rows; ++r) {
      const uint8_t* row = (const uint8_t*)W + (size_t)r * rb;
      if (mode == 0 && quant) {
        for (int t = 0; t < n; ++t) {
          const int16_t* xq = Xq.data() + (size_t)t * cols;
          const float* dxt = dx.data() + (size_t)t * cols / 32;
          Y[(size_t)t * ldy + r] =
              (float)(wtype == ORC_Q4_K ? dot_q4k_q16(row, cols, xq, dxt) : dot_q6k_q16(row, cols, xq, dxt));
        }
        continue;
      }
      if (wtype == ORC_BF16 && n == 1) {  // hot loop of the timed CPU baseline: bf16 row . fp32 x
        const uint16_t* wr = (const uint16_t*)row;
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        for (int64_t k = 0; k < cols; k += 4) {
          a0 += (double)(bf16_to_f32(wr[k + 0]) * Xin[k + 0]);
          a1 += (double)(bf16_to_f32(wr[k + 1]) * Xin[k + 1]);
          a2 += (double)(bf16_to_f32(wr[k + 2]) * Xin[k + 2]);
          a3 += (double)(bf16_to_f32(wr[k + 3]) * Xin[k + 3]);
        }
        Y[r] = (float)((a0 + a1) + (a2 + a3));
        continue;
      }
      dequant_row(wtype, row, cols, wrow.data());
      if (mode == 1 && (quant || wtype == ORC_F16))  // GEMM operands are bf16: k-quant and F16 weights are rounded to it
        for (int64_t k = 0; k < cols; ++k) wrow[k] = bf16_round(wrow[k]);
      // fp32 products, double accumulation.  The reduction clause lets the compiler keep one partial sum per vector
      // lane (the order of a double sum of fp32 products moves the result by ~1e-16 relative: far below fp32);
      // four tokens per pass share the loads of the weight row.
      const float* wr = wrow.data();
      int t = 0;
      for (; t + 3 < n; t += 4) {
        const float* x0 = Xin + (size_t)t * cols;
        const float* x1 = x0 + cols;
        const float* x2 = x1 + cols;
        const float* x3 = x2 + cols;
        double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma omp simd reduction(+ : s0, s1, s2, s3)
        for (int64_t k = 0; k < cols; ++k) {
          s0 += (double)(wr[k] * x0[k]);
          s1 += (double)(wr[k] * x1[k]);
          s2 += (double)(wr[k] * x2[k]);
          s3 += (double)(wr[k] * x3[k]);
        }
        Y[(size_t)t * ldy + r] = (float)s0;
        Y[(size_t)(t + 1) * ldy + r] = (float)s1;
        Y[(size_t)(t + 2) * ldy + r] = (float)s2;
        Y[(size_t)(t + 3) * ldy + r] = (float)s3;
      }
      for (; t < n; ++t) {
        const float* x = Xin + (size_t)t * cols;
        double s0 = 0;
#pragma omp simd reduction(+ : s0)
        for (int64_t k = 0; k < cols; ++k) s0 += (double)(wr[k] * x[k]);
        Y[(size_t)t * ldy + r] = (float)s0;
      }
    }
  }
}

void rmsnorm(const float* x, const float* w, int n, float eps, float* y) {
  double ss = 0.0;
  for (int i = 0; i < n; ++i) ss += (double)x[i] * x[i];
  const float inv = (float)(1.0 / std::sqrt(ss / n + (double)eps));
  for (int i = 0; i < n; ++i) y[i] = (x[i] * inv) * w[i];
}

// cos/sin table [max_ctx][head_dim/2][2], computed in double, stored fp32.
// Llama-3.1 "llama3" rope scaling (transformers modeling_rope_utils._compute_llama3_parameters).
void rope_table(const orc_config& c, float* tab) {
  const int half = c.head_dim / 2;
  std::vector<double> inv(half);
  for (int i = 0; i < half; ++i) {
    double f = 1.0 / std::pow((double)c.rope_theta, (2.0 * i) / c.head_dim);
    if (c.rope_factor > 1.0f) {
      const double lo_wl = (double)c.rope_orig_ctx / c.rope_low_freq_factor;
      const double hi_wl = (double)c.rope_orig_ctx / c.rope_high_freq_factor;
      const double wl = 2.0 * M_PI / f;
      if (wl > lo_wl) f = f / c.rope_factor;
      else if (wl >= hi_wl) {
        const double smooth = ((double)c.rope_orig_ctx / wl - c.rope_low_freq_factor) /
                              ((double)c.rope_high_freq_factor - c.rope_low_freq_factor);
        f = (1.0 - smooth) * f / c.rope_factor + smooth * f;
      }
    }
    inv[i] = f;
  }
  for (int p = 0; p < c.max_ctx; ++p)
    for (int i = 0; i < half; ++i) {
      const double a = (double)p * inv[i];
      tab[((size_t)p * half + i) * 2 + 0] = (float)std::cos(a);
      tab[((size_t)p * half + i) * 2 + 1] = (float)std::sin(a);
    }
}

// adjacent-pair rotation (GGUF "llama" arch convention): (x[2i], x[2i+1]) rotated by angle(pos, i)
void rope_apply(float* v, int n_heads, int head_dim, const float* tab_pos) {
  for (int h = 0; h < n_heads; ++h)
    for (int i = 0; i < head_dim / 2; ++i) {
      float* p = v + (size_t)h * head_dim + 2 * i;
      const float c = tab_pos[2 * i], s = tab_pos[2 * i + 1];
      const float a = p[0], b = p[1];
      p[0] = a * c - b * s;
      p[1] = a * s + b * c;
    }
}

uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

struct Tensor {
  int type = -1;
  const void* data = nullptr;
  int64_t rows = 0, cols = 0;
};

}  // namespace

struct orc_model {
  orc_config c;
  std::map<std::string, Tensor> t;
  std::vector<float> rope;        // [max_ctx][hd/2][2]
  std::vector<float> kc, vc;      // [layer][pos][n_kv*hd]
  int n_past = 0;
  const Tensor& get(const std::string& name) const {
    auto it = t.find(name);
    if (it == t.end() || !it->second.data) {
      fprintf(stderr, "oracle: tensor '%s' not set\n", name.c_str());
      abort();
    }
    return it->second;
  }
};

// ------------------------------------------------------------------ C entry points (ctypes) ------

ORC_API void orc_dequant_q4k(const void* blocks, int64_t nb, float* out) {
  for (int64_t b = 0; b < nb; ++b) dequant_q4k_block((const uint8_t*)blocks + b * Q4K_BYTES, out + b * QK_K);
}
ORC_API void orc_dequant_q6k(const void* blocks, int64_t nb, float* out) {
  for (int64_t b = 0; b < nb; ++b) dequant_q6k_block((const uint8_t*)blocks + b * Q6K_BYTES, out + b * QK_K);
}
ORC_API void orc_quant_q16(const float* x, int64_t n, int16_t* q, float* d) { quant_q16(x, n, q, d); }
ORC_API void orc_f32_to_bf16(const float* x, int64_t n, uint16_t* y) {
  for (int64_t i = 0; i < n; ++i) y[i] = f32_to_bf16(x[i]);
}
ORC_API void orc_matmul(int wtype, const void* W, int64_t rows, int64_t cols, const float* X, int n, int mode,
                        float* Y) {
  matmul(wtype, W, rows, cols, X, n, mode, Y, rows);
}
ORC_API void orc_rmsnorm(const float* x, const float* w, int n, float eps, float* y) { rmsnorm(x, w, n, eps, y); }
ORC_API void orc_rope_table(const orc_config* c, float* tab) { rope_table(*c, tab); }
ORC_API void orc_rope_apply(float* v, int n_heads, int head_dim, const float* tab_pos) {
  rope_apply(v, n_heads, head_dim, tab_pos);
}
ORC_API void orc_swiglu(const float* g, const float* u, int n, float* h) {
  for (int i = 0; i < n; ++i) h[i] = (g[i] / (1.0f + expf(-g[i]))) * u[i];
}
// One query position of GQA attention against a contiguous K/V history [n_ctx][n_kv*hd] (fp32 values).
ORC_API void orc_attention(const float* q, const float* K, const float* V, int n_ctx, int n_heads,
                           int n_kv_heads, int head_dim, float* out) {
  const int grp = n_heads / n_kv_heads, kvd = n_kv_heads * head_dim;
  const float scale = 1.0f / sqrtf((float)head_dim);
  std::vector<float> s(n_ctx);
  for (int h = 0; h < n_heads; ++h) {
    const int kh = h / grp;
    const float* qh = q + (size_t)h * head_dim;
    float mx = -INFINITY;
    for (int t = 0; t < n_ctx; ++t) {
      const float* k = K + (size_t)t * kvd + (size_t)kh * head_dim;
      double a = 0.0;
      for (int i = 0; i < head_dim; ++i) a += (double)(qh[i] * k[i]);
      s[t] = (float)a * scale;
      mx = std::max(mx, s[t]);
    }
    double den = 0.0;
    for (int t = 0; t < n_ctx; ++t) {
      s[t] = expf(s[t] - mx);
      den += s[t];
    }
    for (int i = 0; i < head_dim; ++i) {
      double a = 0.0;
      for (int t = 0; t < n_ctx; ++t) a += (double)(s[t] * V[(size_t)t * kvd + (size_t)kh * head_dim + i]);
      out[(size_t)h * head_dim + i] = (float)(a / den);
    }
  }
}

ORC_API orc_model* orc_model_new(const orc_config* c) {
  orc_model* m = new orc_model();
  m->c = *c;
  m->rope.resize((size_t)c->max_ctx * (c->head_dim / 2) * 2);
  rope_table(*c, m->rope.data());
  const size_t kv = (size_t)c->n_layers * c->max_ctx * c->n_kv_heads * c->head_dim;
  m->kc.assign(kv, 0.f);
  m->vc.assign(kv, 0.f);
  return m;
}
ORC_API void orc_model_free(orc_model* m) { delete m; }
ORC_API void orc_model_reset(orc_model* m) { m->n_past = 0; }
ORC_API int orc_model_n_past(orc_model* m) { return m->n_past; }
// Rewind to the first n cached positions (n <= n_past): lets a timed sample decode a few tokens again and again from
// the same long context without paying its prefill each time.
ORC_API int orc_model_truncate(orc_model* m, int n) {
  if (n < 0 || n > m->n_past) return -1;
  m->n_past = n;
  return 0;
}

// Borrow (no copy) a tensor in GGUF naming: token_embd.weight, blk.N.{attn_norm,attn_q,attn_k,attn_v,
// attn_output,ffn_norm,ffn_gate,ffn_up,ffn_down}.weight, output_norm.weight, output.weight.
ORC_API int orc_model_set_tensor(orc_model* m, const char* name, int type, const void* data, int64_t rows,
                                 int64_t cols) {
  if (type != ORC_F32 && type != ORC_F16 && type != ORC_BF16 && type != ORC_Q4_K && type != ORC_Q6_K) return -1;
  if ((type == ORC_Q4_K || type == ORC_Q6_K) && cols % QK_K) return -2;
  Tensor t;
  t.type = type;
  t.data = data;
  t.rows = rows;
  t.cols = cols;
  m->t[name] = t;
  return 0;
}

// Run n tokens at positions [n_past, n_past+n).  logits: [n][vocab] if all_logits else [vocab] (last token).
ORC_API int orc_model_forward(orc_model* m, const int32_t* toks, int n, int mode, float* logits, int all_logits) {
  const orc_config& c = m->c;
  const int d = c.d_model, hd = c.head_dim, nh = c.n_heads, nkv = c.n_kv_heads, kvd = nkv * hd, qd = nh * hd,
            ff = c.d_ffn;
  if (m->n_past + n > c.max_ctx) return -1;
  const int pos0 = m->n_past;
  std::vector<float> x((size_t)n * d), xn((size_t)n * d), q((size_t)n * qd), k((size_t)n * kvd),
      v((size_t)n * kvd), att((size_t)n * qd), proj((size_t)n * d), g((size_t)n * ff), u((size_t)n * ff);
  const Tensor& emb = m->get("token_embd.weight");
  {
    const size_t rb = row_bytes(emb.type, emb.cols);
    for (int t = 0; t < n; ++t) {
      if (toks[t] < 0 || toks[t] >= c.vocab) return -2;
      dequant_row(emb.type, (const uint8_t*)emb.data + (size_t)toks[t] * rb, d, &x[(size_t)t * d]);
    }
  }
  char name[64];
  auto T = [&](int l, const char* s) -> const Tensor& {
    snprintf(name, sizeof name, "blk.%d.%s.weight", l, s);
    return m->get(name);
  };
  for (int l = 0; l < c.n_layers; ++l) {
    const float* an = (const float*)T(l, "attn_norm").data;
    for (int t = 0; t < n; ++t) rmsnorm(&x[(size_t)t * d], an, d, c.rms_eps, &xn[(size_t)t * d]);
    const Tensor &wq = T(l, "attn_q"), &wk = T(l, "attn_k"), &wv = T(l, "attn_v"), &wo = T(l, "attn_output");
    matmul(wq.type, wq.data, qd, d, xn.data(), n, mode, q.data(), qd);
    matmul(wk.type, wk.data, kvd, d, xn.data(), n, mode, k.data(), kvd);
    matmul(wv.type, wv.data, kvd, d, xn.data(), n, mode, v.data(), kvd);
    float* KC = &m->kc[(size_t)l * c.max_ctx * kvd];
    float* VC = &m->vc[(size_t)l * c.max_ctx * kvd];
    for (int t = 0; t < n; ++t) {
      const float* tab = &m->rope[(size_t)(pos0 + t) * (hd / 2) * 2];
      rope_apply(&q[(size_t)t * qd], nh, hd, tab);
      rope_apply(&k[(size_t)t * kvd], nkv, hd, tab);
      for (int i = 0; i < kvd; ++i) {
        const float kk = k[(size_t)t * kvd + i], vv = v[(size_t)t * kvd + i];
        KC[(size_t)(pos0 + t) * kvd + i] = c.kv_bf16 ? bf16_round(kk) : kk;
        VC[(size_t)(pos0 + t) * kvd + i] = c.kv_bf16 ? bf16_round(vv) : vv;
      }
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < n; ++t)
      orc_attention(&q[(size_t)t * qd], KC, VC, pos0 + t + 1, nh, nkv, hd, &att[(size_t)t * qd]);
    matmul(wo.type, wo.data, d, qd, att.data(), n, mode, proj.data(), d);
    for (size_t i = 0; i < (size_t)n * d; ++i) x[i] += proj[i];
    const float* fn = (const float*)T(l, "ffn_norm").data;
    for (int t = 0; t < n; ++t) rmsnorm(&x[(size_t)t * d], fn, d, c.rms_eps, &xn[(size_t)t * d]);
    const Tensor &wg = T(l, "ffn_gate"), &wu = T(l, "ffn_up"), &wd = T(l, "ffn_down");
    matmul(wg.type, wg.data, ff, d, xn.data(), n, mode, g.data(), ff);
    matmul(wu.type, wu.data, ff, d, xn.data(), n, mode, u.data(), ff);
    for (size_t i = 0; i < (size_t)n * ff; ++i) g[i] = (g[i] / (1.0f + expf(-g[i]))) * u[i];
    matmul(wd.type, wd.data, d, ff, g.data(), n, mode, proj.data(), d);
    for (size_t i = 0; i < (size_t)n * d; ++i) x[i] += proj[i];
  }
  m->n_past += n;
  const float* on = (const float*)m->get("output_norm.weight").data;
  const Tensor& wo = m->get("output.weight");
  if (all_logits) {
    for (int t = 0; t < n; ++t) rmsnorm(&x[(size_t)t * d], on, d, c.rms_eps, &xn[(size_t)t * d]);
    matmul(wo.type, wo.data, c.vocab, d, xn.data(), n, mode, logits, c.vocab);
  } else {
    // The engine always computes the last position's logits with the batch-1 (mode 0) lm_head.
    rmsnorm(&x[(size_t)(n - 1) * d], on, d, c.rms_eps, xn.data());
    matmul(wo.type, wo.data, c.vocab, d, xn.data(), 1, 0, logits, c.vocab);
  }
  return 0;
}

ORC_API int orc_argmax(const float* logits, int n, float* margin) {
  int best = 0;
  float b1 = -INFINITY, b2 = -INFINITY;
  for (int i = 0; i < n; ++i) {
    if (logits[i] > b1) { b2 = b1; b1 = logits[i]; best = i; }
    else if (logits[i] > b2) b2 = logits[i];
  }
  if (margin) *margin = b1 - b2;
  return best;
}

// llama.cpp default sampler chain order: repeat penalty (last_n) -> top-k -> top-p -> temperature -> dist.
// Deterministic RNG: u = top 24 bits of splitmix64(seed + step * golden) * 2^-24.  Ties: lower id first.
ORC_API int orc_sample(const float* logits_in, int vocab, const orc_sampling* sp, const int32_t* history,
                       int n_hist, uint64_t step) {
  std::vector<float> lg(logits_in, logits_in + vocab);
  if (sp->repeat_penalty != 1.0f && sp->repeat_last_n != 0) {
    const int from = (sp->repeat_last_n < 0) ? 0 : std::max(0, n_hist - sp->repeat_last_n);
    std::vector<char> seen(vocab, 0);
    for (int i = from; i < n_hist; ++i)
      if (history[i] >= 0 && history[i] < vocab) seen[history[i]] = 1;
    for (int i = 0; i < vocab; ++i)
      if (seen[i]) lg[i] = lg[i] > 0.f ? lg[i] / sp->repeat_penalty : lg[i] * sp->repeat_penalty;
  }
  if (sp->temperature <= 0.f) return orc_argmax(lg.data(), vocab, nullptr);
  int k = sp->top_k <= 0 ? vocab : std::min(sp->top_k, vocab);
  std::vector<int> idx(vocab);
  for (int i = 0; i < vocab; ++i) idx[i] = i;
  std::partial_sort(idx.begin(), idx.begin() + k, idx.end(),
                    [&](int a, int b) { return lg[a] > lg[b] || (lg[a] == lg[b] && a < b); });
  idx.resize(k);
  // top-p on softmax of the k kept logits (already sorted descending)
  std::vector<float> p(k);
  float mx = lg[idx[0]], sum = 0.f;
  for (int i = 0; i < k; ++i) { p[i] = expf(lg[idx[i]] - mx); sum += p[i]; }
  int keep = k;
  if (sp->top_p < 1.0f) {
    float cum = 0.f;
    for (int i = 0; i < k; ++i) {
      cum += p[i] / sum;
      if (cum >= sp->top_p) { keep = i + 1; break; }
    }
  }
  // temperature, softmax over kept, inverse-CDF draw
  float sum2 = 0.f;
  for (int i = 0; i < keep; ++i) { p[i] = expf((lg[idx[i]] - mx) / sp->temperature); sum2 += p[i]; }
  const uint64_t r = splitmix64(sp->seed + step * 0x9E3779B97F4A7C15ull);
  const float uu = (float)(r >> 40) * 5.9604644775390625e-8f * sum2;
  float cum = 0.f;
  for (int i = 0; i < keep; ++i) {
    cum += p[i];
    if (uu < cum) return idx[i];
  }
  return idx[keep - 1];
}

// Greedy generation: prefill under `prefill_mode` (1 = tensor-core numerics, 0 = token-by-token gemv
// numerics), then batch-1 decode (mode 0).  margins[i] = top1-top2 logit gap at step i (may be NULL).
ORC_API int orc_generate_greedy(orc_model* m, const int32_t* prompt, int n_prompt, int n_new, int prefill_mode,
                                int32_t* out, float* margins) {
  std::vector<float> lg(m->c.vocab);
  orc_model_reset(m);
  int rc;
  if (prefill_mode == 1) rc = orc_model_forward(m, prompt, n_prompt, 1, lg.data(), 0);
  else {
    rc = 0;
    for (int i = 0; i < n_prompt && !rc; ++i) rc = orc_model_forward(m, prompt + i, 1, 0, lg.data(), 0);
  }
  if (rc) return rc;
  for (int i = 0; i < n_new; ++i) {
    float mg;
    const int tok = orc_argmax(lg.data(), m->c.vocab, &mg);
    out[i] = tok;
    if (margins) margins[i] = mg;
    if (i + 1 < n_new) {
      rc = orc_model_forward(m, &tok, 1, 0, lg.data(), 0);
      if (rc) return rc;
    }
  }
  return 0;
}
