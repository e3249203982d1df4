// engine.cu — Engine: weights (fused, GPU-friendly layouts), paged KV pool, the CUDA-graph decode step,
// greedy generation and the request worker.  Host code only; every number is produced by the kernels in
// gemv.cu / attn.cu / misc.cu / gemm.cu / sampler.cu.
#include "engine.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

void tp_destroy(void* comm);

static std::atomic<unsigned long long> g_launches{0};
static thread_local bool g_launch_pause = false;
void count_launch(unsigned long long n) { if (!g_launch_pause) g_launches.fetch_add(n, std::memory_order_relaxed); }
unsigned long long launch_count() { return g_launches.load(std::memory_order_relaxed); }
void count_launches_pause(bool on) { g_launch_pause = on; }

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  (void)cudaGetLastError();
  return (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? B200_ENODEV
         : (e == cudaErrorMemoryAllocation)                           ? B200_ENOMEM
                                                                      : B200_ECUDA;
}

static size_t tensor_row_bytes(int type, int64_t cols) {
  return type == T_F32 ? (size_t)cols * 4 : gemv_row_bytes(type, cols);
}

// llama.cpp's Q4_K_M recipe (llama_tensor_get_type, use_more_bits): which layers keep attn_v / ffn_down in Q6_K
static bool more_bits(int i, int n) { return i < n / 8 || i >= 7 * n / 8 || (i - n / 8) % 3 == 2; }

Engine::~Engine() {
  {
    std::lock_guard<std::mutex> lk(qmu);
    stopping = true;
  }
  qcv.notify_all();
  if (worker.joinable()) worker.join();
  if (stream) cudaStreamSynchronize(stream);
  if (step_graph) cudaGraphExecDestroy(step_graph);
  for (auto& g : batch_graphs) cudaGraphExecDestroy(g.second.first);
  if (b_in_host) cudaFreeHost(b_in_host);
  for (void* p : tp_ipc_opened) cudaIpcCloseMemHandle(p);
  tp_destroy(nccl_comm);
  for (void* p : allocs) cudaFree(p);
  if (out_tokens) cudaFreeHost(out_tokens);
  if (host_n_out) cudaFreeHost((void*)host_n_out);
  if (stream) cudaStreamDestroy(stream);
}

int Engine::alloc(void** p, size_t bytes) {
  if (bytes == 0) bytes = 16;
  cudaError_t e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
  allocs.push_back(*p);
  return 0;
}

int Engine::init(const b200_config& c, const void* nccl_id) {
  cfg = c;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("no CUDA device available (%s); libb200llm has no CPU fallback", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return B200_ENODEV;
  }
  if (c.device < 0 || c.device >= ndev) {
    set_error("device %d out of range (%d devices)", c.device, ndev);
    return B200_EINVAL;
  }
  B200_CUDA(cudaSetDevice(c.device));
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, c.device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library only contains sm_100a code", c.device, prop.major, prop.minor);
    return B200_ENODEV;
  }
  sm_count = prop.multiProcessorCount;
  use_pdl = getenv("B200_NO_PDL") == nullptr;
  use_graph = getenv("B200_NO_GRAPH") == nullptr;
  use_batch_graph = use_graph && getenv("B200_NO_BATCH_GRAPH") == nullptr;
  use_fa_prefill = getenv("B200_NO_FA_PREFILL") == nullptr;
  use_mma_decode = getenv("B200_NO_MMA_DECODE") == nullptr;
  use_attn_prefetch = getenv("B200_NO_ATTN_PREFETCH") == nullptr;
  use_attn_cluster = use_attn_prefetch && getenv("B200_NO_ATTN_CLUSTER") == nullptr && attn_cluster_supported(c.n_heads / c.n_kv_heads);
  use_prefix_cache = getenv("B200_NO_PREFIX_CACHE") == nullptr && (c.reserved & 2) == 0 && tp_size == 1;
  // L2 prefetch chain between the decode GEMVs (see enqueue_step): B200_L2_CHAIN=<MB> overrides the amount, 0 = off
  l2_chain_mb = getenv("B200_L2_CHAIN") ? atoi(getenv("B200_L2_CHAIN")) : -1;

  const int tp = tp_size, R = tp_rank;
  if (tp < 1 || R < 0 || R >= tp || c.n_heads % tp || c.n_kv_heads % tp || c.d_ffn % tp || c.vocab % tp ||
      (tp > 1 && c.weight_format != B200_WEIGHTS_BF16)) {
    set_error("tensor parallel degree %d unsupported for this config (heads, kv heads, d_ffn, vocab must divide; bf16)", tp);
    return B200_EINVAL;
  }
  lc = c;
  lc.n_heads /= tp; lc.n_kv_heads /= tp; lc.d_ffn /= tp; lc.vocab /= tp;
  // local (per-rank) sizes; upper-case = the full model
  const int d = c.d_model, qd = lc.n_heads * c.head_dim, kvd = lc.n_kv_heads * c.head_dim, ff = lc.d_ffn, voc = lc.vocab;
  const int QD = c.n_heads * c.head_dim, KVD = c.n_kv_heads * c.head_dim, FF = c.d_ffn, VOC = c.vocab;
  if (c.head_dim != kHeadDim || c.n_layers <= 0 || c.n_kv_heads <= 0 || c.n_heads % c.n_kv_heads ||
      c.max_ctx <= 0 || c.vocab <= 0 || c.max_seqs < 1) {
    set_error("unsupported config (head_dim must be 128, n_heads %% n_kv_heads == 0, sizes > 0)");
    return B200_EINVAL;
  }
  const int G = c.n_heads / c.n_kv_heads;
  if (G != 1 && G != 2 && G != 4 && G != 8) {
    set_error("GQA group %d unsupported (1,2,4,8)", G);
    return B200_EINVAL;
  }
  const bool q4 = c.weight_format == B200_WEIGHTS_Q4_K_M;
  if (!q4 && c.weight_format != B200_WEIGHTS_BF16) {
    set_error("unknown weight_format %d", c.weight_format);
    return B200_EINVAL;
  }
  const int tmat = q4 ? T_Q4_K : T_BF16, tbig = q4 ? T_Q6_K : T_BF16;
  B200_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  if (tp_size > 1) {
    if (!nccl_id) { set_error("tensor parallelism needs the NCCL unique id of rank 0"); return B200_EINVAL; }
    int rc = tp_init(nccl_id);
    if (rc) return rc;
  }

  auto need = [&](int type, int64_t N, int64_t K, const char* what) -> int {
    if (!gemv_shape_ok(type, N, K, 0)) {
      set_error("shape %lldx%lld of %s not supported for tensor type %d (bf16: K%%256, Q4_K/Q6_K: K%%2048, N%%4)",
                (long long)N, (long long)K, what, type);
      return B200_EINVAL;
    }
    return 0;
  };
  auto add_slot = [&](const std::string& name, int type, int64_t rows, int64_t cols, uint8_t* base, int64_t row0,
                      int64_t step, int64_t full_rows = -1, int64_t full_cols = -1, int64_t src_row0 = 0, int64_t src_col0 = 0) {
    Slot s;
    s.type = type; s.rows = rows; s.cols = cols; s.base = base; s.row0 = row0; s.row_step = step;
    s.full_rows = full_rows < 0 ? rows : full_rows; s.full_cols = full_cols < 0 ? cols : full_cols;
    s.src_row0 = src_row0; s.src_col0 = src_col0;
    slots[name] = s;
  };

  // storage type of a tensor: what a GGUF file says (type_override, filled by the loader) or the recipe
  auto ty = [&](const std::string& name, int recipe) {
    auto it = type_override.find(name);
    return it == type_override.end() ? recipe : it->second;
  };
  layers.resize(c.n_layers);
  weight_bytes = 0;
  for (int l = 0; l < c.n_layers; ++l) {
    Layer& L = layers[l];
    const std::string p = "blk." + std::to_string(l) + ".";
    const int t_more = (q4 && more_bits(l, c.n_layers)) ? T_Q6_K : tmat;
    const int tq = ty(p + "attn_q.weight", tmat), tk = ty(p + "attn_k.weight", tmat), tv = ty(p + "attn_v.weight", t_more);
    const int to = ty(p + "attn_output.weight", tmat), tg = ty(p + "ffn_gate.weight", tmat), tu = ty(p + "ffn_up.weight", tmat);
    const int tdown = ty(p + "ffn_down.weight", t_more);
    if (tq != tk || tg != tu) { set_error("layer %d: attn_q/attn_k (and ffn_gate/ffn_up) must share a storage type", l); return B200_EINVAL; }
    int rc;
    if ((rc = need(tq, qd + kvd, d, "attn_q/k")) || (rc = need(tv, kvd, d, "attn_v")) || (rc = need(to, d, qd, "attn_output")) ||
        (rc = need(tg, 2 * ff, d, "ffn_gate/up")) || (rc = need(tdown, d, ff, "ffn_down")))
      return rc;
    L.qkv_type = tq;
    const size_t rb = gemv_row_bytes(tq, d);
    if (tv == tq) {
      if ((rc = alloc((void**)&L.qkv, (size_t)(qd + 2 * kvd) * rb))) return rc;
      add_slot(p + "attn_v.weight", tv, kvd, d, L.qkv, qd + kvd, 1, KVD, d, (int64_t)R * kvd, 0);
    } else {
      if ((rc = alloc((void**)&L.qkv, (size_t)(qd + kvd) * rb))) return rc;
      if ((rc = alloc((void**)&L.v, (size_t)kvd * gemv_row_bytes(tv, d)))) return rc;
      L.v_type = tv;
      add_slot(p + "attn_v.weight", tv, kvd, d, L.v, 0, 1, KVD, d, (int64_t)R * kvd, 0);
    }
    add_slot(p + "attn_q.weight", tq, qd, d, L.qkv, 0, 1, QD, d, (int64_t)R * qd, 0);
    add_slot(p + "attn_k.weight", tq, kvd, d, L.qkv, qd, 1, KVD, d, (int64_t)R * kvd, 0);
    L.o_type = to;
    if ((rc = alloc((void**)&L.o, (size_t)d * gemv_row_bytes(to, qd)))) return rc;
    add_slot(p + "attn_output.weight", to, d, qd, L.o, 0, 1, d, QD, 0, (int64_t)R * qd);
    L.gu_type = tg;
    const size_t rbg = gemv_row_bytes(tg, d);
    if ((rc = alloc((void**)&L.gate_up, (size_t)2 * ff * rbg))) return rc;
    add_slot(p + "ffn_gate.weight", tg, ff, d, L.gate_up, 0, 2, FF, d, (int64_t)R * ff, 0);
    add_slot(p + "ffn_up.weight", tg, ff, d, L.gate_up, 1, 2, FF, d, (int64_t)R * ff, 0);
    L.down_type = tdown;
    if ((rc = alloc((void**)&L.down, (size_t)d * gemv_row_bytes(tdown, ff)))) return rc;
    add_slot(p + "ffn_down.weight", tdown, d, ff, L.down, 0, 1, d, FF, 0, (int64_t)R * ff);
    if ((rc = alloc((void**)&L.attn_norm, (size_t)d * 4)) || (rc = alloc((void**)&L.ffn_norm, (size_t)d * 4))) return rc;
    add_slot(p + "attn_norm.weight", T_F32, 1, d, (uint8_t*)L.attn_norm, 0, 1);
    add_slot(p + "ffn_norm.weight", T_F32, 1, d, (uint8_t*)L.ffn_norm, 0, 1);
    weight_bytes += (int64_t)(qd + kvd) * rb + (int64_t)kvd * gemv_row_bytes(tv, d) + (int64_t)d * gemv_row_bytes(to, qd) +
                    (int64_t)2 * ff * rbg + (int64_t)d * gemv_row_bytes(tdown, ff) + 2 * (int64_t)d * 4;
  }
  {
    int rc;
    embd_type = ty("token_embd.weight", tmat);
    const int tout = ty("output.weight", tbig);
    if (embd_type != T_BF16 && embd_type != T_F16 && embd_type != T_F32 && d % QK_K) { set_error("d_model must be a multiple of 256 for k-quant embeddings"); return B200_EINVAL; }
    if ((rc = need(tout, voc, d, "output"))) return rc;
    if ((rc = alloc((void**)&token_embd, (size_t)c.vocab * tensor_row_bytes(embd_type, d)))) return rc;
    add_slot("token_embd.weight", embd_type, c.vocab, d, token_embd, 0, 1);
    output_type = tout;
    if ((rc = alloc((void**)&output, (size_t)voc * gemv_row_bytes(tout, d)))) return rc;
    add_slot("output.weight", tout, voc, d, output, 0, 1, VOC, d, (int64_t)R * voc, 0);
    if ((rc = alloc((void**)&output_norm, (size_t)d * 4))) return rc;
    add_slot("output_norm.weight", T_F32, 1, d, (uint8_t*)output_norm, 0, 1);
    weight_bytes += (int64_t)voc * gemv_row_bytes(tout, d) + (int64_t)d * 4;
  }

  // paged KV pool: page = 16 tokens of one kv head; static ownership (sequence slot s owns a page run)
  pages_per_seq = (c.max_ctx + kPageTokens - 1) / kPageTokens;
  const size_t n_pages = (size_t)pages_per_seq * c.max_seqs;
  const size_t layer_kv = n_pages * lc.n_kv_heads * kPageTokens * kHeadDim * 2;
  for (auto& L : layers) {
    int rc;
    if ((rc = alloc((void**)&L.k_cache, layer_kv)) || (rc = alloc((void**)&L.v_cache, layer_kv))) return rc;
    B200_CUDA(cudaMemsetAsync(L.k_cache, 0, layer_kv, stream));
    B200_CUDA(cudaMemsetAsync(L.v_cache, 0, layer_kv, stream));
  }
  {
    int rc;
    if ((rc = alloc((void**)&block_table, n_pages * 4))) return rc;
    std::vector<int32_t> bt(n_pages);
    for (size_t i = 0; i < n_pages; ++i) bt[i] = (int32_t)i;
    B200_CUDA(cudaMemcpy(block_table, bt.data(), n_pages * 4, cudaMemcpyHostToDevice));
    if ((rc = alloc((void**)&active_table, (size_t)pages_per_seq * 4))) return rc;
    B200_CUDA(cudaMemcpy(active_table, block_table, (size_t)pages_per_seq * 4, cudaMemcpyDeviceToDevice));
    seqs.resize(c.max_seqs);
    slot_kv.assign(c.max_seqs, {});
  }
  // activations + sequence state
  n_splits = 20;
  {
    int rc;
    if ((rc = alloc((void**)&x, (size_t)d * 4)) || (rc = alloc((void**)&q, (size_t)qd * 4)) || (rc = alloc((void**)&att, (size_t)qd * 4)) ||
        (rc = alloc((void**)&h, (size_t)ff * 4)) || (rc = alloc((void**)&logits, (size_t)c.vocab * 4)) ||
        (rc = alloc((void**)&rope, (size_t)c.max_ctx * (kHeadDim / 2) * 2 * 4)) ||
        (rc = alloc((void**)&attn_partial, attn_partial_floats(lc.n_heads, lc.n_kv_heads, n_splits) * 4)) ||
        (rc = alloc((void**)&tp_partial, (size_t)d * 4)) ||
        (rc = alloc((void**)&attn_counters, (size_t)lc.n_kv_heads * 4)) || (rc = alloc((void**)&st, sizeof(StepState))) ||
        (rc = alloc((void**)&sample_part, sample_part_bytes(cfg.vocab))) ||
        (rc = alloc((void**)&sp_dev, sizeof(SampleParams))) ||
        (rc = alloc((void**)&prompt_dev, (size_t)c.max_ctx * 4)))
      return rc;
    B200_CUDA(cudaMemset(attn_counters, 0, (size_t)lc.n_kv_heads * 4));
    B200_CUDA(cudaMemset(st, 0, sizeof(StepState)));
  }
  B200_CUDA(cudaHostAlloc((void**)&out_tokens, (size_t)c.max_ctx * 4, cudaHostAllocMapped));
  B200_CUDA(cudaHostGetDevicePointer((void**)&out_tokens_dev, out_tokens, 0));
  B200_CUDA(cudaHostAlloc((void**)&host_n_out, 64, cudaHostAllocMapped));
  B200_CUDA(cudaHostGetDevicePointer((void**)&host_n_out_dev, (void*)host_n_out, 0));
  *host_n_out = 0;

  {  // RoPE table in double on the host ("llama3" scaling), stored fp32: [max_ctx][64][2]
    const int half = kHeadDim / 2;
    std::vector<double> inv(half);
    for (int i = 0; i < half; ++i) {
      double f = 1.0 / std::pow((double)c.rope_theta, (2.0 * i) / kHeadDim);
      if (c.rope_factor > 1.0f) {
        const double lo_wl = (double)c.rope_orig_ctx / c.rope_low_freq_factor, hi_wl = (double)c.rope_orig_ctx / c.rope_high_freq_factor;
        const double wl = 2.0 * M_PI / f;
        if (wl > lo_wl) f = f / c.rope_factor;
        else if (wl >= hi_wl) {
          const double sm = ((double)c.rope_orig_ctx / wl - c.rope_low_freq_factor) / ((double)c.rope_high_freq_factor - c.rope_low_freq_factor);
          f = (1.0 - sm) * f / c.rope_factor + sm * f;
        }
      }
      inv[i] = f;
    }
    std::vector<float> tab((size_t)c.max_ctx * half * 2);
    for (int p = 0; p < c.max_ctx; ++p)
      for (int i = 0; i < half; ++i) {
        const double a = (double)p * inv[i];
        tab[((size_t)p * half + i) * 2] = (float)std::cos(a);
        tab[((size_t)p * half + i) * 2 + 1] = (float)std::sin(a);
      }
    B200_CUDA(cudaMemcpy(rope, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice));
  }
  if (tp_size > 1 && getenv("B200_TP_NCCL") == nullptr) {
    int rc = tp_setup_link();
    if (rc) return rc;
  }
  if (c.seed >= 0) {
    int rc = synthesize();
    if (rc) return rc;
  }
  B200_CUDA(cudaStreamSynchronize(stream));
  worker = std::thread([this] { worker_loop(); });
  return 0;
}

// `src` holds the FULL tensor in GGUF layout; the slot takes its window of it (the whole tensor unless TP).
int Engine::place_tensor(const Slot& s, const uint8_t* src) {
  const size_t rb = tensor_row_bytes(s.type, s.cols), full_rb = tensor_row_bytes(s.type, s.full_cols);
  uint8_t* dst = s.base + (size_t)s.row0 * rb;
  const bool window = s.full_rows != s.rows || s.full_cols != s.cols;
  if (window && s.type != T_BF16 && s.type != T_F16 && s.type != T_F32) { set_error("tensor-parallel windows need bf16 / f32 tensors"); return B200_EINVAL; }
  const uint8_t* from = src + (size_t)s.src_row0 * full_rb + tensor_row_bytes(s.type, s.src_col0);
  if (s.type == T_Q6_K && s.base != token_embd) {
    B200_CUDA(launch_q6k_repack(from, dst, s.rows, s.cols / QK_K, 1, stream));
  } else {
    B200_CUDA(cudaMemcpy2DAsync(dst, (size_t)s.row_step * rb, from, full_rb, rb, (size_t)s.rows, cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}

// The slot's (local) tensor back in GGUF layout, rows x cols.
int Engine::fetch_tensor(const Slot& s, uint8_t* dst) {
  const size_t rb = tensor_row_bytes(s.type, s.cols);
  const uint8_t* src = s.base + (size_t)s.row0 * rb;
  if (s.type == T_Q6_K && s.base != token_embd) {
    B200_CUDA(launch_q6k_repack(src, dst, s.rows, s.cols / QK_K, 0, stream));
  } else {
    B200_CUDA(cudaMemcpy2DAsync(dst, rb, src, (size_t)s.row_step * rb, rb, (size_t)s.rows, cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}

bool Engine::all_set() const {
  for (auto& kv : slots)
    if (!kv.second.set) return false;
  return true;
}

int Engine::synthesize() {
  size_t maxb = 0;
  for (auto& kv : slots) maxb = std::max(maxb, (size_t)kv.second.full_rows * tensor_row_bytes(kv.second.type, kv.second.full_cols));
  uint8_t* tmp = nullptr;
  B200_CUDA(cudaMalloc((void**)&tmp, maxb));
  uint32_t tid = 0;
  int rc = 0;
  for (auto& kv : slots) {  // std::map: deterministic (name-sorted) tensor ids
    Slot& s = kv.second;
    // std 1/sqrt(fan_in); the two projections that write into the residual stream are scaled by
    // 1/sqrt(2*n_layers) (GPT-2 / Megatron random-init convention), which keeps a deep random network from
    // amplifying rounding differences chaotically layer after layer
    float sigma = kv.first == "token_embd.weight" ? 1.0f : 1.0f / sqrtf((float)s.full_cols);
    if (kv.first.find("attn_output") != std::string::npos || kv.first.find("ffn_down") != std::string::npos)
      sigma /= sqrtf(2.0f * (float)cfg.n_layers);
    cudaError_t e = launch_synth(s.type, tmp, s.full_rows, s.full_cols, (uint64_t)cfg.seed, tid++, sigma, stream);
    if (e != cudaSuccess) { rc = cuda_fail(e, "synth"); break; }
    if ((rc = place_tensor(s, tmp))) break;
    s.set = true;
  }
  cudaError_t e = cudaStreamSynchronize(stream);
  cudaFree(tmp);
  if (!rc && e != cudaSuccess) rc = cuda_fail(e, "synthesize sync");
  return rc;
}

void Engine::lm_head_args(GemvArgs* out, int* pro_out, bool consume_tp) {
  const b200_config& c = lc;
  GemvArgs lm = {};
  lm.W = output; lm.N = c.vocab; lm.K = c.d_model; lm.x = x; lm.norm_w = output_norm; lm.eps = c.rms_eps;
  int pro = PRO_RMSNORM;
  if (consume_tp) {  // the last down projection of the step is still sitting in the receive buffers
    pro = PRO_TP_RMSNORM;
    lm.tp = tp_link; lm.tp_index = 2 * c.n_layers - 1; lm.x = x_alt; lm.x_out = x;  // index odd -> reads buffer 1
  }
  lm.y = logits + (size_t)tp_rank * c.vocab;  // vocab-parallel: this rank's slice of the logits
  lm.row_base = tp_rank * c.vocab;             // global token id of row 0 (argmax key)
  lm.st = st; lm.store_logits = 1; lm.argmax_key = &st->argmax_key;
  *out = lm;
  *pro_out = pro;
}

// lm_head GEMV (logits + argmax), sampler (no-op for greedy requests), finish_step (token hand-over)
int Engine::enqueue_tail(cudaStream_t s, bool pdl, int* n, bool consume_tp, bool with_lm_head) {
  const b200_config& c = lc;
  if (with_lm_head) {
    GemvArgs lm;
    int pro;
    lm_head_args(&lm, &pro, consume_tp);
    B200_CUDA(launch_gemv(output_type, pro, EPI_LOGITS, lm, s, pdl, sm_count)); ++*n;
  }
  if (tp_size > 1) {  // global argmax = max over ranks of the packed (logit, ~id) keys
    int rc = tp_allreduce_max_u64(&st->argmax_key, s);
    if (rc) return rc;
    ++*n;
  }
  SampleArgs sa = {};
  sa.logits = logits; sa.vocab = cfg.vocab; sa.params = sp_dev; sa.history = prompt_dev; sa.st = st; sa.out = &st->sampled; sa.n_hist = -1;
  sa.part = sample_part;
  B200_CUDA(launch_sample(sa, s, pdl && tp_size == 1 && with_lm_head));
  *n += (sample_part && cfg.vocab >= 4096) ? 2 : 1;
  FinishArgs f = {};
  f.st = st; f.prompt = prompt_dev; f.out_tokens = out_tokens_dev; f.host_n_out = host_n_out_dev; f.embd = token_embd;
  f.embd_type = embd_type; f.d_model = c.d_model; f.x = x; f.max_ctx = c.max_ctx;
  f.tp_epoch = tp_fused ? tp_link_host.epoch : nullptr;
  B200_CUDA(launch_finish_step(f, s, pdl)); ++*n;
  return 0;
}

int Engine::dump_trace() {
  const char* path = getenv("B200_TRACE");
  if (!trace_dev || !path || trace_names.empty()) return 0;
  std::vector<unsigned long long> t(trace_names.size() * 4);
  B200_CUDA(cudaMemcpy(t.data(), trace_dev, t.size() * 8, cudaMemcpyDeviceToHost));
  const std::string file = tp_size > 1 ? std::string(path) + ".r" + std::to_string(tp_rank) : std::string(path);  // one file per rank
  FILE* f = fopen(file.c_str(), "w");
  if (!f) return 0;
  for (size_t i = 0; i < trace_names.size(); ++i)
    fprintf(f, "%zu %s %llu %llu %llu %llu\n", i, trace_names[i].c_str(), t[4 * i], t[4 * i + 1], t[4 * i + 2], t[4 * i + 3]);
  fclose(f);
  return 0;
}

int Engine::enqueue_step(cudaStream_t s) {
  const b200_config& c = lc;  // per-rank shapes (== cfg without tensor parallelism)
  const int d = c.d_model, qd = c.n_heads * kHeadDim, kvd = c.n_kv_heads * kHeadDim, ff = c.d_ffn;
  const bool tp = tp_size > 1;
  int n = 0;
  trace_names.clear();
  auto tr = [&](const char* name) -> unsigned long long* {  // stamps of the launch about to be enqueued
    if (!trace_dev) return nullptr;
    trace_names.push_back(name);
    return trace_dev + (trace_names.size() - 1) * 4;
  };
  for (int l = 0; l < c.n_layers; ++l) {
    Layer& L = layers[l];
    // fused all-reduce: residual buffers alternate, all-reduce i is consumed from xbuf[i&1] into xbuf[(i+1)&1]
    float* xbuf[2] = {x, x_alt};
    const bool fused = tp && tp_fused;
    GemvArgs a = {};
    a.W = L.qkv; a.N = L.v ? qd + kvd : qd + 2 * kvd; a.K = d; a.x = x; a.norm_w = L.attn_norm; a.eps = c.rms_eps; a.y = q;
    int qkv_pro = PRO_RMSNORM;
    if (fused && l > 0) {  // consumes the previous layer's down projection (all-reduce 2l-1)
      qkv_pro = PRO_TP_RMSNORM;
      a.tp = tp_link; a.tp_index = 2 * l - 1; a.x = xbuf[(2 * l - 1) & 1]; a.x_out = xbuf[(2 * l) & 1];
    }
    a.row_base = 0; a.q_rows = qd; a.kv_rows = kvd; a.n_kv_heads = c.n_kv_heads; a.rope = rope;
    a.k_cache = L.k_cache; a.v_cache = L.v_cache; a.block_table = active_table; a.st = st;
    // L2 prefetch chain: qkv -> o (waits in L2 while attention runs), o -> gate/up, gate/up -> down, down -> next qkv
    // A warp that has streamed its last unit asks L2 for a slice of the first megabytes of the NEXT kernel's weights
    // (GemvArgs::next_W), so that kernel's first ring fills come from L2 while HBM would otherwise idle across the
    // kernel boundary.  The amount matters: 40 MB (round 1) made the step slower, 12 MB makes the bf16 step 3.5 %
    // faster (2.78 -> 2.68 ms), 6 MB the Q4_K_M step 0.5 % (1.461 -> 1.454); more displaces what is about to be used.
    // resident_early: the next kernel (o) is already on the SMs, filling its own ring, while attention runs between the
    // two — a third of the amount for 16-bit weights, none for k-quants (measured: 2.677 -> 2.667 ms, 1.411 -> 1.405 ms)
    auto chain = [&](GemvArgs& g, const uint8_t* w, int type, int64_t rows, int64_t k, bool resident_early = false) {
      const bool kq = type == T_Q4_K || type == T_Q6_K;
      if (resident_early && l2_chain_mb < 0) {
        g.next_W = w;
        g.next_bytes = kq ? 0ull : std::min<unsigned long long>(4ull << 20, (unsigned long long)rows * gemv_row_bytes(type, k)) & ~127ull;
        return;
      }
      const unsigned long long cap = l2_chain_mb >= 0 ? (unsigned long long)l2_chain_mb << 20
                                                      : ((type == T_Q4_K || type == T_Q6_K) ? (6ull << 20) : (12ull << 20));
      g.next_W = w;
      g.next_bytes = std::min<unsigned long long>(cap, (unsigned long long)rows * gemv_row_bytes(type, k)) & ~127ull;
    };
    chain(a, L.o, L.o_type, d, qd, true);
    bool dual = false;
    a.trace = tr("qkv");
    if (L.v && qkv_pro == PRO_RMSNORM) {  // q,k and a differently-typed v: one launch when the pair is supported
      GemvArgs bv = a;
      bv.next_bytes = 0;
      bv.W = L.v; bv.N = kvd; bv.row_base = qd + kvd;
      const cudaError_t de = launch_gemv_dual_qkv(L.qkv_type, a, L.v_type, bv, s, use_pdl, sm_count);
      if (de == cudaSuccess) { dual = true; ++n; }
      else if (de != cudaErrorNotSupported) return cuda_fail(de, "launch_gemv_dual_qkv");
      else (void)cudaGetLastError();
    }
    if (!dual) {
      B200_CUDA(launch_gemv(L.qkv_type, qkv_pro, EPI_QKV, a, s, use_pdl, sm_count)); ++n;
      if (L.v) {
        a.next_bytes = 0;
        a.W = L.v; a.N = kvd; a.row_base = qd + kvd;
        a.trace = tr("v");
        B200_CUDA(launch_gemv(L.v_type, PRO_RMSNORM, EPI_QKV, a, s, use_pdl, sm_count)); ++n;
      }
    }
    AttnArgs t = {};
    t.q = q; t.k_cache = L.k_cache; t.v_cache = L.v_cache; t.block_table = active_table; t.st = st;
    t.n_heads = c.n_heads; t.n_kv_heads = c.n_kv_heads; t.out = att; t.partial = attn_partial; t.counters = attn_counters;
    t.n_splits = n_splits; t.fixed_ctx = 0;
    t.prefetch_tokens = use_attn_prefetch ? pages_per_seq * kPageTokens : 0;
    t.use_cluster = use_attn_cluster ? 1 : 0;
    t.trace = tr("attn");
    B200_CUDA(launch_attn_decode(t, s, use_pdl)); ++n;
    // row-parallel projections: without TP the residual add is the GEMV epilogue; with TP every rank holds a
    // K-slice, so the partial products are summed over NVLink (NCCL all-reduce) before they join the stream
    GemvArgs o = {};
    o.W = L.o; o.N = d; o.K = qd; o.x = att; o.y = tp ? tp_partial : x;
    chain(o, L.gate_up, L.gu_type, 2 * ff, d);
    if (fused) { o.tp = tp_link; o.tp_index = 2 * l; o.y = nullptr; }
    o.trace = tr("o");
    B200_CUDA(launch_gemv(L.o_type, PRO_COPY, fused ? EPI_TP_PUSH : (tp ? EPI_STORE : EPI_RESID), o, s, use_pdl, sm_count)); ++n;
    if (tp && !fused) {
      int rc = tp_allreduce_sum(tp_partial, (size_t)d, s);
      if (rc) return rc;
      B200_CUDA(launch_add_inplace(x, tp_partial, d, s)); n += 2;
    }
    GemvArgs g = {};
    g.W = L.gate_up; g.N = 2 * ff; g.K = d; g.x = x; g.norm_w = L.ffn_norm; g.eps = c.rms_eps; g.y = h;
    chain(g, L.down, L.down_type, d, ff);
    g.trace = tr("gate_up");
    int gu_pro = PRO_RMSNORM;
    if (fused) {  // consumes this layer's o projection (all-reduce 2l): reads xbuf[0], publishes xbuf[1]
      gu_pro = PRO_TP_RMSNORM;
      g.tp = tp_link; g.tp_index = 2 * l; g.x = xbuf[(2 * l) & 1]; g.x_out = xbuf[(2 * l + 1) & 1];
    }
    B200_CUDA(launch_gemv(L.gu_type, gu_pro, EPI_SWIGLU, g, s, use_pdl && (!tp || fused), sm_count)); ++n;
    GemvArgs dn = {};
    dn.W = L.down; dn.N = d; dn.K = ff; dn.x = h; dn.y = tp ? tp_partial : x;
    if (l + 1 < c.n_layers) chain(dn, layers[l + 1].qkv, layers[l + 1].qkv_type, layers[l + 1].v ? qd + kvd : qd + 2 * kvd, d);
    else chain(dn, output, output_type, c.vocab, d);
    if (fused) { dn.tp = tp_link; dn.tp_index = 2 * l + 1; dn.y = nullptr; }
    dn.trace = tr("down");
    B200_CUDA(launch_gemv(L.down_type, PRO_COPY, fused ? EPI_TP_PUSH : (tp ? EPI_STORE : EPI_RESID), dn, s, use_pdl, sm_count)); ++n;
    if (tp && !fused) {
      int rc = tp_allreduce_sum(tp_partial, (size_t)d, s);
      if (rc) return rc;
      B200_CUDA(launch_add_inplace(x, tp_partial, d, s)); n += 2;
    }
  }
  {
    int rc = enqueue_tail(s, use_pdl, &n, tp && tp_fused);
    if (rc) return rc;
  }
  launches_per_step = n;
  return 0;
}

int Engine::ensure_graph() {
  if (!all_set()) { set_error("not all tensors are set"); return B200_ESTATE; }
  if (getenv("B200_TRACE") && !trace_dev) {  // before any capture: allocation is not allowed inside one
    const size_t bytes = (size_t)(6 * lc.n_layers + 8) * 4 * 8;
    int rc = alloc((void**)&trace_dev, bytes);
    if (rc) return rc;
    B200_CUDA(cudaMemset(trace_dev, 0, bytes));
  }
  if (!use_graph || step_graph) return 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamBeginCapture");
    count_launches_pause(true);
    int rc = enqueue_step(stream);
    count_launches_pause(false);
    e = cudaStreamEndCapture(stream, &g);
    if (rc == 0 && e == cudaSuccess) {
      size_t n_nodes = 0;
      if (cudaGraphGetNodes(g, nullptr, &n_nodes) == cudaSuccess) step_graph_nodes = (int)n_nodes;
      e = cudaGraphInstantiate(&step_graph, g, 0);
      cudaGraphDestroy(g);
      if (e == cudaSuccess) return 0;
    } else if (g) {
      cudaGraphDestroy(g);
    }
    (void)cudaGetLastError();
    step_graph = nullptr;
    if (use_pdl) {
      use_pdl = false;  // retry the capture with plain stream-order edges
      continue;
    }
    return rc ? rc : cuda_fail(e, "graph capture/instantiate");
  }
  return B200_ECUDA;
}

int Engine::launch_steps(int n) {
  if (!all_set()) { set_error("not all tensors are set"); return B200_ESTATE; }
  int rc = ensure_graph();
  if (rc) return rc;
  for (int i = 0; i < n; ++i) {
    if (step_graph) { B200_CUDA(cudaGraphLaunch(step_graph, stream)); count_launch((unsigned long long)step_graph_nodes); }
    else if ((rc = enqueue_step(stream))) return rc;
  }
  return 0;
}

int Engine::begin_sequence(const int32_t* prompt, int n_prompt, int max_out, int greedy, const b200_sampling* sp) {
  if (n_prompt < 1 || n_prompt > cfg.max_ctx || max_out < 1 || n_prompt + max_out - 1 > cfg.max_ctx) {
    set_error("prompt of %d tokens + %d new tokens does not fit max_ctx %d", n_prompt, max_out, cfg.max_ctx);
    return B200_EINVAL;
  }
  for (int i = 0; i < n_prompt; ++i)
    if (prompt[i] < 0 || prompt[i] >= cfg.vocab) { set_error("token id %d out of range", prompt[i]); return B200_EINVAL; }
  B200_CUDA(cudaMemcpyAsync(prompt_dev, prompt, (size_t)n_prompt * 4, cudaMemcpyHostToDevice, stream));
  StepState s0 = {};
  s0.pos = 0; s0.cur_token = prompt[0]; s0.n_prompt = n_prompt; s0.n_out = 0; s0.max_out = max_out; s0.eos = cfg.eos_token;
  s0.done = 0; s0.greedy = greedy; s0.argmax_key = 0ull;
  B200_CUDA(cudaMemcpyAsync(st, &s0, sizeof s0, cudaMemcpyHostToDevice, stream));
  SampleParams sp0 = {};
  if (sp) {
    sp0.temperature = sp->temperature; sp0.top_p = sp->top_p; sp0.repeat_penalty = sp->repeat_penalty;
    sp0.top_k = sp->top_k; sp0.repeat_last_n = sp->repeat_last_n; sp0.seed = sp->seed;
  }
  B200_CUDA(cudaMemcpyAsync(sp_dev, &sp0, sizeof sp0, cudaMemcpyHostToDevice, stream));
  B200_CUDA(cudaStreamSynchronize(stream));  // s0 / sp0 / prompt are stack / caller memory
  *host_n_out = 0;
  B200_CUDA(launch_embed(token_embd, embd_type, cfg.d_model, &st->cur_token, x, stream));
  return 0;
}

// The synchronous hooks borrow slot 0 and the single-sequence state; they refuse to run under live requests.
static int claim_slot0(Engine* e) {
  for (auto& q : e->seqs)
    if (q.r) { set_error("engine is serving requests; synchronous hooks need an idle engine"); return B200_ESTATE; }
  e->bs1_owner = -1;
  e->invalidate_prefix_cache(0);  // the hook is about to overwrite slot 0's K/V
  return e->set_active_slot(0);
}

int Engine::generate_greedy(const int32_t* prompt, int n_prompt, int32_t* ids, int n_new, int prefill_mode, float* ttft_ms,
                            float* decode_ms) {
  int rc = claim_slot0(this);
  if (rc) return rc;
  if ((rc = ensure_graph())) return rc;
  if (prefill_mode == B200_MODE_GEMM && n_prompt > 1 && (rc = ensure_prefill())) return rc;
  cudaEvent_t e0, e1, e2;
  B200_CUDA(cudaEventCreate(&e0));
  B200_CUDA(cudaEventCreate(&e1));
  B200_CUDA(cudaEventCreate(&e2));
  B200_CUDA(cudaEventRecord(e0, stream));
  rc = run_prompt(prompt, n_prompt, n_new, 1, prefill_mode);  // after this the first generated token exists
  if (!rc) {
    cudaEventRecord(e1, stream);
    rc = launch_steps(n_new - 1);
    cudaEventRecord(e2, stream);
  }
  cudaError_t e = cudaStreamSynchronize(stream);
  if (!rc && e != cudaSuccess) rc = cuda_fail(e, "generate sync");
  if (!rc) rc = tp_check_error();
  if (!rc) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, e0, e1);
    cudaEventElapsedTime(&b, e1, e2);
    if (ttft_ms) *ttft_ms = a;
    if (decode_ms) *decode_ms = b;
    const int got = *host_n_out;
    for (int i = 0; i < n_new; ++i) ids[i] = i < got ? out_tokens[i] : -1;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
  return rc;
}

int Engine::forward_logits(const int32_t* ids, int n, int mode, float* logits_host) {
  int rc = claim_slot0(this);
  if (rc) return rc;
  if ((rc = run_prompt(ids, n, 1, 1, mode))) return rc;
  if (tp_size > 1) {  // every rank holds its vocab slice in place; exchange the slices
    int rc2 = tp_allgather(logits + (size_t)tp_rank * lc.vocab, logits, (size_t)lc.vocab, stream);
    if (rc2) return rc2;
  }
  B200_CUDA(cudaMemcpyAsync(logits_host, logits, (size_t)cfg.vocab * 4, cudaMemcpyDeviceToHost, stream));
  B200_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

// Teacher-forced decode (parity hook): the prompt goes through `prefill_mode`, then forced[k] is fed at position
// n_prompt + k through the batch-1 decode step whatever the model itself would have picked.  logits_host receives
// n_forced + 1 rows of `vocab` floats: row 0 after the prompt, row k + 1 after forced[k].
int Engine::forward_logits_forced(const int32_t* prompt, int n_prompt, int prefill_mode, const int32_t* forced, int n_forced,
                                  float* logits_host) {
  int rc = claim_slot0(this);
  if (rc) return rc;
  if (tp_size > 1) { set_error("forward_logits_forced: single-GPU engines only"); return B200_ESTATE; }
  for (int k = 0; k < n_forced; ++k)
    if (forced[k] < 0 || forced[k] >= cfg.vocab) { set_error("token id %d out of range", forced[k]); return B200_EINVAL; }
  if ((rc = ensure_graph())) return rc;
  if ((rc = run_prompt(prompt, n_prompt, n_forced + 1, 1, prefill_mode))) return rc;
  const size_t row = (size_t)cfg.vocab;
  B200_CUDA(cudaMemcpyAsync(logits_host, logits, row * 4, cudaMemcpyDeviceToHost, stream));
  for (int k = 0; k < n_forced; ++k) {
    // the step about to run processes position n_prompt + k: replace the token the model picked for it
    B200_CUDA(cudaMemcpyAsync(&st->cur_token, forced + k, 4, cudaMemcpyHostToDevice, stream));
    B200_CUDA(cudaMemcpyAsync(prompt_dev + n_prompt + k, forced + k, 4, cudaMemcpyHostToDevice, stream));
    B200_CUDA(launch_embed(token_embd, embd_type, cfg.d_model, &st->cur_token, x, stream));
    if ((rc = launch_steps(1))) return rc;
    B200_CUDA(cudaMemcpyAsync(logits_host + (size_t)(k + 1) * row, logits, row * 4, cudaMemcpyDeviceToHost, stream));
  }
  B200_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

}  // namespace b200

// ------------------------------------------------------------------------------------------------------
// Prefill on the tensor cores (GEMM numerics, oracle mode 1).
namespace b200 {

// bf16 shadow (tensor-core operand) of one GGUF-named matrix, rebuilt from the tensor as it is stored now.
// Called for every matrix when the shadows are first built and again whenever b200_set_tensor replaces a tensor
// afterwards, so the GEMM path never runs on stale weights.  No-op for names without a shadow.
int Engine::refresh_shadow(const std::string& name) {
  if (!has_shadows) return 0;
  const b200_config& c = lc;
  const int qd = c.n_heads * kHeadDim, kvd = c.n_kv_heads * kHeadDim;
  auto it = slots.find(name);
  if (it == slots.end()) return 0;
  const Slot& s = it->second;
  uint16_t* base = nullptr;
  int64_t row0 = 0, step = 1;
  if (name == "output.weight") {
    if (!output16 || output_type == T_BF16) return 0;
    base = output16;
  } else if (name.compare(0, 4, "blk.") == 0) {
    const size_t dot = name.find('.', 4);
    if (dot == std::string::npos) return 0;
    const int l = atoi(name.c_str() + 4);
    if (l < 0 || l >= (int)layers.size()) return 0;
    Layer& L = layers[l];
    const std::string t = name.substr(dot + 1);
    if (t == "attn_q.weight") { base = L.qkv16; }
    else if (t == "attn_k.weight") { base = L.qkv16; row0 = qd; }
    else if (t == "attn_v.weight") { base = L.qkv16; row0 = qd + kvd; }
    else if (t == "attn_output.weight") { base = L.o16; }
    else if (t == "ffn_gate.weight") { base = L.gu16; step = 2; }
    else if (t == "ffn_up.weight") { base = L.gu16; row0 = 1; step = 2; }
    else if (t == "ffn_down.weight") { base = L.down16; }
    else return 0;
  } else {
    return 0;
  }
  if (!base) return 0;
  int rc;
  if (name == "output.weight") {  // larger than the per-layer staging buffers: its own temporary
    const size_t bytes = (size_t)s.rows * gemv_row_bytes(s.type, s.cols);
    uint8_t* tmp = nullptr;
    B200_CUDA(cudaMalloc((void**)&tmp, bytes));
    rc = fetch_tensor(s, tmp);
    cudaError_t e = rc ? cudaSuccess
                       : (s.type == T_F16 ? launch_f16_to_bf16((const uint16_t*)tmp, output16, s.rows * s.cols, stream)
                                          : launch_dequant(s.type, tmp, s.rows * s.cols / QK_K, nullptr, output16, stream));
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(tmp);
    if (rc) return rc;
    return e == cudaSuccess ? 0 : cuda_fail(e, "lm_head shadow");
  }
  if (!shadow_tmp) return 0;  // all-bf16 layers: the "shadows" are the GEMV buffers themselves
  if ((rc = fetch_tensor(s, shadow_tmp))) return rc;  // GGUF block order
  const uint16_t* src16 = (const uint16_t*)shadow_tmp;  // bf16 tensors are copied as they are
  if (s.type == T_F16) {
    B200_CUDA(launch_f16_to_bf16((const uint16_t*)shadow_tmp, shadow_tmp16, s.rows * s.cols, stream));
    src16 = shadow_tmp16;
  } else if (s.type != T_BF16) {
    B200_CUDA(launch_dequant(s.type, shadow_tmp, s.rows * s.cols / QK_K, nullptr, shadow_tmp16, stream));
    src16 = shadow_tmp16;
  }
  B200_CUDA(cudaMemcpy2DAsync(base + (size_t)row0 * s.cols, (size_t)step * s.cols * 2, src16, (size_t)s.cols * 2, (size_t)s.cols * 2,
                              (size_t)s.rows, cudaMemcpyDeviceToDevice, stream));
  return 0;
}

// Up to 8 prompts (4096 rows) share one pass: at 512 rows the 256 x 256 GEMM tiles of gate/up fill 1.5 waves of the
// 148 SMs (a quarter of the machine idles in the second), at 2048-4096 rows 6-12 waves.
int Engine::prefill_rows_cap() const {
  if (cfg.max_seqs <= 1 || tp_size > 1) return 0;
  const int64_t want = (int64_t)cfg.max_ctx * std::min(cfg.max_seqs, 8);
  return (int)std::min<int64_t>(4096, want);
}

int Engine::ensure_prefill() {
  if (pf_ready) return 0;
  const b200_config& c = lc;
  const int d = c.d_model, qd = c.n_heads * kHeadDim, kvd = c.n_kv_heads * kHeadDim, ff = c.d_ffn, T = std::max(std::max(c.max_ctx, c.max_seqs), prefill_rows_cap());
  if (!gemm_shape_ok(qd + 2 * kvd, d, 1) || !gemm_shape_ok(d, qd, 1) || !gemm_shape_ok(2 * ff, d, 1) || !gemm_shape_ok(d, ff, 1)) {
    set_error("model dims not supported by the tensor-core path (rows %%128, K %%64)");
    return B200_EINVAL;
  }
  int rc;
  bool q4 = false;  // any k-quant matrix -> bf16 shadow copies (same fused row order) for the tensor-core path
  for (auto& L : layers) q4 = q4 || L.qkv_type != T_BF16 || (L.v && L.v_type != T_BF16) || L.o_type != T_BF16 || L.gu_type != T_BF16 || L.down_type != T_BF16;
  if (q4) {
    size_t maxb = 0;
    for (auto& kv : slots) maxb = std::max(maxb, (size_t)kv.second.rows * tensor_row_bytes(kv.second.type, kv.second.cols));
    if ((rc = alloc((void**)&shadow_tmp, maxb)) || (rc = alloc((void**)&shadow_tmp16, (size_t)2 * ff * std::max(d, qd) * 2 + (size_t)d * ff * 2))) return rc;
    for (int l = 0; l < c.n_layers; ++l) {
      Layer& L = layers[l];
      if ((rc = alloc((void**)&L.qkv16, (size_t)(qd + 2 * kvd) * d * 2)) || (rc = alloc((void**)&L.o16, (size_t)d * qd * 2)) ||
          (rc = alloc((void**)&L.gu16, (size_t)2 * ff * d * 2)) || (rc = alloc((void**)&L.down16, (size_t)d * ff * 2)))
        return rc;
    }
    has_shadows = true;
    for (int l = 0; l < c.n_layers; ++l)
      for (const char* t : {"attn_q", "attn_k", "attn_v", "attn_output", "ffn_gate", "ffn_up", "ffn_down"})
        if ((rc = refresh_shadow("blk." + std::to_string(l) + "." + t + ".weight"))) return rc;
    B200_CUDA(cudaStreamSynchronize(stream));
  } else {
    for (auto& L : layers) {
      L.qkv16 = (uint16_t*)L.qkv; L.o16 = (uint16_t*)L.o; L.gu16 = (uint16_t*)L.gate_up; L.down16 = (uint16_t*)L.down;
    }
  }
  const size_t wmax = (size_t)std::max(std::max(d, qd), ff);
  if ((rc = alloc((void**)&pf_x, (size_t)T * d * 4)) || (rc = alloc((void**)&pf_xb, (size_t)T * wmax * 2)) ||
      (rc = alloc((void**)&pf_qkv, (size_t)T * (qd + 2 * kvd) * 4)) || (rc = alloc((void**)&pf_q, (size_t)T * qd * 4)) ||
      (rc = alloc((void**)&pf_att, (size_t)T * qd * 4)) ||
      (rc = alloc((void**)&pf_proj, (size_t)(pf_proj_floats = (int64_t)std::max(2 * T, 2560) * d) * 4)) ||
      (rc = alloc((void**)&pf_gu, (size_t)T * 2 * ff * 4)) || (rc = alloc((void**)&pf_hb, (size_t)T * ff * 2)) ||
      (rc = alloc((void**)&pf_partial, (size_t)T * attn_partial_floats(c.n_heads, c.n_kv_heads, pf_splits) * 4)) ||
      (rc = alloc((void**)&pf_counters, (size_t)T * c.n_kv_heads * 4)) || (rc = alloc((void**)&pfb_tok, (size_t)T * 4)) ||
      (rc = alloc((void**)&pfb_pos, (size_t)T * 4)) || (rc = alloc((void**)&pfb_slot, (size_t)T * 4)))
    return rc;
  B200_CUDA(cudaMemsetAsync(pf_counters, 0, (size_t)T * c.n_kv_heads * 4, stream));
  gemm_ws.sm_count = sm_count;
  gemm_ws.ws_bytes = (size_t)sm_count * 256 * 256 * 4;
  if ((rc = alloc((void**)&gemm_ws.ws, gemm_ws.ws_bytes)) || (rc = alloc((void**)&gemm_ws.flags, (size_t)(sm_count + 8) * 4))) return rc;
  B200_CUDA(cudaMemsetAsync(gemm_ws.flags, 0, (size_t)(sm_count + 8) * 4, stream));
  pf_rows = T;
  pf_ready = true;
  return 0;
}

// The layer stack on the tensor cores over n rows.  pos_arr == nullptr: rows are positions 0..n-1 of the
// sequence whose block table is `active_table` (prefill).  Otherwise row t is position pos_arr[t] of the
// sequence in slot slot_arr[t], which then sees ctx_arr[t] tokens (batched decode).
// Leaves the residual stream of every row in pf_x (+ pf_proj still to be added).
int Engine::forward_gemm(int n, const int32_t* tokens_dev, const int32_t* pos_arr, const int32_t* ctx_arr, const int32_t* slot_arr, int pos0,
                         const std::vector<PrefillSeg>* segs) {
  const b200_config& c = lc;
  const int d = c.d_model, qd = c.n_heads * kHeadDim, kvd = c.n_kv_heads * kHeadDim, ff = c.d_ffn;
  cudaStream_t s = stream;
  const int32_t* table = pos_arr ? block_table : active_table;
  B200_CUDA(launch_embed_tokens(token_embd, embd_type, d, tokens_dev, n, pf_x, s));
  for (int l = 0; l < c.n_layers; ++l) {
    Layer& L = layers[l];
    { int rc = add_rmsnorm_proj(L.attn_norm, n, l > 0); if (rc) return rc; }
    int qkv_splits = max_wide_splits(n);
    const int64_t qkv_stride = (int64_t)n * (qd + 2 * kvd);
    B200_CUDA(launch_gemm_bf16_splitk(L.qkv16, qd + 2 * kvd, d, pf_xb, n, pf_qkv, qd + 2 * kvd, qkv_stride, &qkv_splits, s, &gemm_ws));
    B200_CUDA(launch_rope_kv(pf_qkv, pf_q, L.k_cache, L.v_cache, table, rope, n, qd, kvd, c.n_kv_heads, pos0, pos_arr, slot_arr, pages_per_seq,
                             qkv_splits, qkv_stride, s));
    if (segs) {  // several prompts: causal attention per prompt over its own rows and its own slot's cache
      static_assert(sizeof(PrefillSeg) == sizeof(int4), "PrefillSeg is passed to the kernel as int4 (row0, n, pos0, slot)");
      B200_CUDA(launch_attn_prefill_segs(pf_q, L.k_cache, L.v_cache, block_table, pages_per_seq, pf_xb,
                                         reinterpret_cast<const int4*>(segs->data()), (int)segs->size(), c.n_heads, c.n_kv_heads, s));
    } else if (!pos_arr && (use_fa_prefill || pos0 > 0)) {  // one sequence, causal: tensor-core attention, bf16 straight into the GEMM operand
      B200_CUDA(launch_attn_prefill(pf_q, L.k_cache, L.v_cache, table, pf_xb, n, pos0, c.n_heads, c.n_kv_heads, s));
    } else {                            // one row per sequence (batched decode): the split-KV decode kernel, a query per z
      AttnArgs t = {};
      t.q = pf_q; t.k_cache = L.k_cache; t.v_cache = L.v_cache; t.block_table = table; t.st = nullptr;
      t.n_heads = c.n_heads; t.n_kv_heads = c.n_kv_heads; t.out = pf_att; t.partial = pf_partial; t.counters = pf_counters;
      t.n_splits = pf_splits; t.fixed_ctx = 1; t.n_queries = n; t.q_stride = qd;
      t.q_ctx = ctx_arr; t.q_slot = slot_arr; t.pages_per_seq = pages_per_seq;
      if (use_mma_decode) {  // tensor cores, bf16 straight into the GEMM operand
        B200_CUDA(launch_attn_decode_batched(t, pf_xb, s));
      } else {
        B200_CUDA(launch_attn_decode(t, s, false));
        B200_CUDA(launch_f32_to_bf16(pf_att, pf_xb, (int64_t)n * qd, s));
      }
    }
    // o and down have few output tiles (d_model/256 x tokens/256): K is split when that under-fills the GPU; the
    // partials sit n*d floats apart in pf_proj and the next consumer adds them in a fixed order
    pf_proj_splits = max_proj_splits(n);
    pf_proj_stride = (int64_t)n * d;
    B200_CUDA(launch_gemm_bf16_splitk(L.o16, d, qd, pf_xb, n, pf_proj, d, pf_proj_stride, &pf_proj_splits, s, &gemm_ws));
    { int rc = reduce_proj(n, pos_arr != nullptr); if (rc) return rc; }
    { int rc = add_rmsnorm_proj(L.ffn_norm, n, true); if (rc) return rc; }
    // gate/up with SwiGLU in the GEMM's epilogue: bf16 h leaves the kernel (pf_gu only when K had to be split)
    B200_CUDA(launch_gemm_bf16_swiglu(L.gu16, 2 * ff, d, pf_xb, n, pf_hb, ff, pf_gu, max_wide_splits(n), s, &gemm_ws));
    pf_proj_splits = max_proj_splits(n);
    B200_CUDA(launch_gemm_bf16_splitk(L.down16, d, ff, pf_hb, n, pf_proj, d, pf_proj_stride, &pf_proj_splits, s, &gemm_ws));
    { int rc = reduce_proj(n, pos_arr != nullptr); if (rc) return rc; }
  }
  return 0;
}

// A row-parallel projection (o, down) has just written its [n][d_model] output (or split-K partials) to pf_proj.
// One GPU: nothing to do, the consumer adds the partials.  Tensor parallel: batched decode of <= kTpBatchRows rows
// pushes the partial into every peer's receive buffer (the consumer waits for the flags and adds the tp slots in
// rank order); anything else is an NCCL all-reduce in place.
int Engine::reduce_proj(int n, bool decode) {
  const int d = lc.d_model;
  pf_proj_src = pf_proj; pf_wait_flags = nullptr;
  if (tp_size <= 1) return 0;
  if (decode && tp_fused && tp_batch_push && n <= kTpBatchRows) {
    const int buf = (int)(tp_bseq & 1u);
    ++tp_bseq;
    B200_CUDA(launch_tp_push(tp_batch, pf_proj, (int64_t)n * d, buf, tp_bseq, stream));
    pf_proj_src = tp_batch.recv[buf]; pf_proj_splits = tp_size; pf_proj_stride = tp_batch.slot_floats;
    pf_wait_flags = tp_batch.flags + buf * kMaxTp; pf_wait_seq = tp_bseq;
    return 0;
  }
  return tp_allreduce_sum(pf_proj, (size_t)n * d, stream);
}

int Engine::add_rmsnorm_proj(const float* w, int n, bool have_proj) {
  B200_CUDA(launch_add_rmsnorm_bf16(pf_x, have_proj ? pf_proj_src : nullptr, pf_proj_splits, pf_proj_stride, w, pf_xb, n, lc.d_model, lc.rms_eps,
                                    stream, have_proj ? pf_wait_flags : nullptr, tp_size, pf_wait_seq, tp_batch.error));
  return 0;
}

int Engine::prefill_gemm(int n_prompt, int pos0) {
  const b200_config& c = cfg;
  const int d = c.d_model;
  cudaStream_t s = stream;
  const int n = n_prompt - pos0;  // rows to compute: positions pos0 .. n_prompt-1
  int rc0 = forward_gemm(n, prompt_dev + pos0, nullptr, nullptr, nullptr, pos0);
  if (rc0) return rc0;
  // hand the last position over to the batch-1 path: residual stream, then lm_head (mode 0) + token pick
  B200_CUDA(launch_take_row(pf_x, pf_proj_src, pf_proj_splits, pf_proj_stride, d, n - 1, x, s));  // prefill: never the push path
  int n_tail = 0;
  int rc = enqueue_tail(s, false, &n_tail);
  if (rc) return rc;
  return 0;
}

// Puts the prompt through the model and produces the first generated token.
// mode GEMM: tensor-core prefill; mode GEMV (or a 1-token prompt): token by token through the decode graph.
int Engine::run_prompt(const int32_t* prompt, int n_prompt, int max_out, int greedy, int mode, const b200_sampling* sp, int reuse) {
  int rc;
  if (reuse < 0 || reuse >= n_prompt || mode != B200_MODE_GEMM) reuse = 0;
  if (mode == B200_MODE_GEMM && n_prompt > 1) {
    if ((rc = ensure_prefill())) return rc;
    if ((rc = begin_sequence(prompt, n_prompt, max_out, greedy, sp))) return rc;
    // the state kernels see "position n-1 is being processed": the token after it is the first generated one
    StepState s0 = {};
    s0.pos = n_prompt - 1; s0.cur_token = prompt[n_prompt - 1]; s0.n_prompt = n_prompt; s0.max_out = max_out; s0.eos = cfg.eos_token;
    s0.greedy = greedy;
    B200_CUDA(cudaMemcpyAsync(st, &s0, sizeof s0, cudaMemcpyHostToDevice, stream));
    B200_CUDA(cudaStreamSynchronize(stream));
    return prefill_gemm(n_prompt, reuse);
  }
  if ((rc = begin_sequence(prompt, n_prompt, max_out, greedy, sp))) return rc;
  return launch_steps(n_prompt);
}

}  // namespace b200
