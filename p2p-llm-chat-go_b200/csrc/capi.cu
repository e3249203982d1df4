// capi.cu — the extern "C" surface declared in include/b200llm.h.  Thin: argument checks, host<->device
// staging for the parity hooks, and forwarding into Engine.  No arithmetic happens on the host.
#include <cstdio>
#include <cstring>

#include "engine.h"
#include "gguf.h"

namespace b200 {
int tp_unique_id(void* out128);
const char* last_error();
}  // namespace b200

using namespace b200;

namespace {

struct DevBuf {  // RAII device scratch for the stand-alone hooks
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t n) {
    cudaError_t e = cudaMalloc(&p, n ? n : 16);
    return e == cudaSuccess ? 0 : cuda_fail(e, "cudaMalloc(hook)");
  }
  template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

int have_device() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    set_error("no CUDA device available (%s); libb200llm has no CPU fallback", cudaGetErrorString(e));
    (void)cudaGetLastError();
    return B200_ENODEV;
  }
  return 0;
}

int sm_count_now() {
  int dev = 0, n = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  return n;
}

// Upload a GGUF-layout matrix and bring it into the device layout of its type.
int upload_matrix(int type, const void* w, int64_t rows, int64_t cols, DevBuf& out) {
  const size_t bytes = (size_t)rows * gemv_row_bytes(type, cols);
  int rc = out.alloc(bytes);
  if (rc) return rc;
  if (type == T_Q6_K) {
    DevBuf tmp;
    if ((rc = tmp.alloc(bytes))) return rc;
    B200_CUDA(cudaMemcpy(tmp.p, w, bytes, cudaMemcpyHostToDevice));
    B200_CUDA(launch_q6k_repack(tmp.as<uint8_t>(), out.as<uint8_t>(), rows, cols / QK_K, 1, 0));
    B200_CUDA(cudaDeviceSynchronize());
  } else {
    B200_CUDA(cudaMemcpy(out.p, w, bytes, cudaMemcpyHostToDevice));
  }
  return 0;
}

int gemv_hook(int type, const void* w, int64_t rows, int64_t cols, const float* x, const float* norm_w, float eps, float* y) {
  int rc = have_device();
  if (rc) return rc;
  if (!w || !x || !y) { set_error("null argument"); return B200_EINVAL; }
  if (!gemv_shape_ok(type, rows, cols, EPI_STORE)) {
    set_error("gemv shape %lldx%lld unsupported for type %d", (long long)rows, (long long)cols, type);
    return B200_EINVAL;
  }
  DevBuf dw, dx, dn, dy;
  if ((rc = upload_matrix(type, w, rows, cols, dw)) || (rc = dx.alloc((size_t)cols * 4)) || (rc = dy.alloc((size_t)rows * 4))) return rc;
  B200_CUDA(cudaMemcpy(dx.p, x, (size_t)cols * 4, cudaMemcpyHostToDevice));
  if (norm_w) {
    if ((rc = dn.alloc((size_t)cols * 4))) return rc;
    B200_CUDA(cudaMemcpy(dn.p, norm_w, (size_t)cols * 4, cudaMemcpyHostToDevice));
  }
  GemvArgs a = {};
  a.W = dw.as<uint8_t>(); a.N = (int)rows; a.K = (int)cols; a.x = dx.as<float>(); a.norm_w = dn.as<float>(); a.eps = eps; a.y = dy.as<float>();
  B200_CUDA(launch_gemv(type, norm_w ? PRO_RMSNORM : PRO_COPY, EPI_STORE, a, 0, false, sm_count_now()));
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(y, dy.p, (size_t)rows * 4, cudaMemcpyDeviceToHost));
  return 0;
}

}  // namespace

extern "C" {

const char* b200_last_error(void) { return last_error(); }
const char* b200_version(void) { return "b200llm 0.1 (sm_100a)"; }

unsigned long long b200_launch_count(void) { return launch_count(); }

int b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
  return n;
}

void b200_default_sampling(b200_sampling* s) {
  if (!s) return;
  memset(s, 0, sizeof *s);
  s->temperature = 0.8f; s->top_p = 0.9f; s->repeat_penalty = 1.1f; s->top_k = 40; s->repeat_last_n = 64;
  s->seed = 0; s->num_predict = 128; s->prefill_mode = B200_MODE_GEMM;
}

static void llama31(b200_config* c, int layers, int d, int heads, int ff) {
  memset(c, 0, sizeof *c);
  c->n_layers = layers; c->d_model = d; c->n_heads = heads; c->n_kv_heads = 8; c->head_dim = 128; c->d_ffn = ff;
  c->vocab = 128256; c->max_ctx = 1024; c->rms_eps = 1e-5f; c->rope_theta = 500000.0f; c->rope_factor = 8.0f;
  c->rope_low_freq_factor = 1.0f; c->rope_high_freq_factor = 4.0f; c->rope_orig_ctx = 8192;
  c->weight_format = B200_WEIGHTS_BF16; c->device = 0; c->max_seqs = 1; c->seed = 0; c->eos_token = 128009;
}
void b200_llama31_8b(b200_config* c) { if (c) llama31(c, 32, 4096, 32, 14336); }
void b200_llama31_70b(b200_config* c) { if (c) llama31(c, 80, 8192, 64, 28672); }

int b200_init(const b200_config* cfg, b200_engine** out) {
  if (!cfg || !out) { set_error("null argument"); return B200_EINVAL; }
  *out = nullptr;
  Engine* e = new (std::nothrow) Engine();
  if (!e) return B200_ENOMEM;
  int rc = e->init(*cfg);
  if (rc) { delete e; return rc; }
  *out = reinterpret_cast<b200_engine*>(e);
  return 0;
}

int b200_tp_unique_id(void* id128) {
  if (!id128) { set_error("null argument"); return B200_EINVAL; }
  return tp_unique_id(id128);
}

int b200_init_tp(const b200_config* cfg, int tp_rank, int tp_size, const void* id128, b200_engine** out) {
  if (!cfg || !out) { set_error("null argument"); return B200_EINVAL; }
  *out = nullptr;
  Engine* e = new (std::nothrow) Engine();
  if (!e) return B200_ENOMEM;
  e->tp_rank = tp_rank;
  e->tp_size = tp_size;
  int rc = e->init(*cfg, id128);
  if (rc) { delete e; return rc; }
  *out = reinterpret_cast<b200_engine*>(e);
  return 0;
}

void b200_shutdown(b200_engine* h) { delete reinterpret_cast<Engine*>(h); }

int b200_get_config(const b200_engine* h, b200_config* out) {
  if (!h || !out) { set_error("null argument"); return B200_EINVAL; }
  *out = reinterpret_cast<const Engine*>(h)->cfg;
  return 0;
}

int b200_set_tensor(b200_engine* h, const char* name, int type, const void* data, int64_t rows, int64_t cols) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !name || !data) { set_error("null argument"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  auto it = e->slots.find(name);
  if (it == e->slots.end()) { set_error("unknown tensor '%s'", name); return B200_EINVAL; }
  Slot& s = it->second;
  if (s.type != type || s.full_rows * s.full_cols != rows * cols || (s.full_rows != 1 && (s.full_rows != rows || s.full_cols != cols))) {
    set_error("tensor '%s' expects type %d shape %lldx%lld, got type %d shape %lldx%lld", name, s.type, (long long)s.full_rows,
              (long long)s.full_cols, type, (long long)rows, (long long)cols);
    return B200_EINVAL;
  }
  for (auto& q : e->seqs)
    if (q.r) { set_error("b200_set_tensor: the engine is serving requests"); return B200_ESTATE; }
  {
    std::lock_guard<std::mutex> ql(e->qmu);
    if (!e->queue.empty()) { set_error("b200_set_tensor: the engine is serving requests"); return B200_ESTATE; }
  }
  B200_CUDA(cudaSetDevice(e->cfg.device));
  // the caller always passes the FULL tensor; a tensor-parallel rank keeps its window of it
  const size_t bytes = (size_t)s.full_rows * (s.type == T_F32 ? (size_t)s.full_cols * 4 : gemv_row_bytes(s.type, s.full_cols));
  DevBuf tmp;
  int rc = tmp.alloc(bytes);
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(tmp.p, data, bytes, cudaMemcpyHostToDevice, e->stream));
  if ((rc = e->place_tensor(s, tmp.as<uint8_t>()))) return rc;
  if ((rc = e->refresh_shadow(name))) return rc;  // k-quant models: the tensor-core path's bf16 copy follows
  e->invalidate_prefix_cache();                   // cached K/V were computed with the old weights
  B200_CUDA(cudaStreamSynchronize(e->stream));
  s.set = true;
  return 0;
}

int b200_tensor_info(const b200_engine* h, const char* name, int* type, int64_t* rows, int64_t* cols, int64_t* nbytes) {
  const Engine* e = reinterpret_cast<const Engine*>(h);
  if (!e || !name) { set_error("null argument"); return B200_EINVAL; }
  auto it = e->slots.find(name);
  if (it == e->slots.end()) { set_error("unknown tensor '%s'", name); return B200_EINVAL; }
  const Slot& s = it->second;
  if (type) *type = s.type;
  if (rows) *rows = s.full_rows;
  if (cols) *cols = s.full_cols;
  if (nbytes) *nbytes = (int64_t)s.full_rows * (int64_t)(s.type == T_F32 ? (size_t)s.full_cols * 4 : gemv_row_bytes(s.type, s.full_cols));
  return 0;
}

int b200_get_tensor(b200_engine* h, const char* name, void* buf, int64_t cap) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !name || !buf) { set_error("null argument"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  auto it = e->slots.find(name);
  if (it == e->slots.end()) { set_error("unknown tensor '%s'", name); return B200_EINVAL; }
  const Slot& s = it->second;
  if (!s.set) { set_error("tensor '%s' has not been set", name); return B200_ESTATE; }
  if (s.full_rows != s.rows || s.full_cols != s.cols) { set_error("b200_get_tensor: '%s' is sharded across tensor-parallel ranks", name); return B200_ESTATE; }
  const size_t bytes = (size_t)s.rows * (s.type == T_F32 ? (size_t)s.cols * 4 : gemv_row_bytes(s.type, s.cols));
  if ((int64_t)bytes > cap) { set_error("buffer too small: need %zu bytes", bytes); return B200_ERANGE; }
  B200_CUDA(cudaSetDevice(e->cfg.device));
  DevBuf tmp;
  int rc = tmp.alloc(bytes);
  if (rc) return rc;
  if ((rc = e->fetch_tensor(s, tmp.as<uint8_t>()))) return rc;
  B200_CUDA(cudaMemcpyAsync(buf, tmp.p, bytes, cudaMemcpyDeviceToHost, e->stream));
  B200_CUDA(cudaStreamSynchronize(e->stream));
  return 0;
}

int64_t b200_weight_bytes(const b200_engine* h) { return h ? reinterpret_cast<const Engine*>(h)->weight_bytes : 0; }

static int tok_encode(const Tokenizer& tk, int32_t vocab, const char* utf8, size_t n, int add_bos, int32_t* ids, size_t cap, size_t* n_out) {
  // n bytes cannot become fewer than n / (longest token) ids: refuse texts that cannot fit `cap` before spending
  // time on them (a front end passes cap = max_ctx, so an oversized body is rejected without being tokenised)
  if (cap > 0 && n / tk.max_token_bytes() > cap) {
    *n_out = n / tk.max_token_bytes();
    set_error("tokenize: %zu bytes of text cannot fit %zu ids", n, cap);
    return B200_ERANGE;
  }
  std::vector<int32_t> v;
  tk.encode(utf8, n, add_bos != 0, vocab, &v);
  *n_out = v.size();
  if (v.size() > cap || (v.size() && !ids)) { set_error("tokenize: need room for %zu ids", v.size()); return B200_ERANGE; }
  if (!v.empty()) memcpy(ids, v.data(), v.size() * 4);
  return 0;
}
static int tok_decode(const Tokenizer& tk, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out) {
  std::string s;
  tk.decode(ids, n, &s);
  *n_out = s.size();
  if (s.size() > cap || (s.size() && !buf)) { set_error("detokenize: need room for %zu bytes", s.size()); return B200_ERANGE; }
  if (!s.empty()) memcpy(buf, s.data(), s.size());
  return 0;
}

int b200_tokenize(b200_engine* h, const char* utf8, size_t n, int add_bos, int32_t* ids, size_t cap, size_t* n_out) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || (!utf8 && n) || !n_out) { set_error("null argument"); return B200_EINVAL; }
  return tok_encode(e->tok, e->cfg.vocab, utf8, n, add_bos, ids, cap, n_out);
}
int b200_detokenize(b200_engine* h, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || (!ids && n) || !n_out) { set_error("null argument"); return B200_EINVAL; }
  return tok_decode(e->tok, ids, n, buf, cap, n_out);
}

int b200_prefix_match(b200_engine* h, const int32_t* ids, size_t n) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !ids || n < 2 || !e->use_prefix_cache) return 0;
  const std::vector<int32_t> p(ids, ids + n);
  int best = 0;
  for (int s = 0; s < (int)e->slot_kv.size(); ++s) best = std::max(best, e->common_prefix(s, p));
  return best < Engine::kMinReuse ? 0 : best;
}

int b200_chat_template(b200_engine* h, char* buf, size_t cap, size_t* n_out) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !n_out) { set_error("null argument"); return B200_EINVAL; }
  const std::string& t = e->tok.chat_template;
  *n_out = t.size();
  if (t.size() > cap || (t.size() && !buf)) { set_error("chat template: need room for %zu bytes", t.size()); return B200_ERANGE; }
  if (!t.empty()) memcpy(buf, t.data(), t.size());
  return 0;
}

int b200_stop_ids(b200_engine* h, int32_t* ids, size_t cap, size_t* n_out) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !n_out) { set_error("null argument"); return B200_EINVAL; }
  std::vector<int32_t> v = e->tok.stop_ids;
  if (v.empty() && e->cfg.eos_token >= 0) v.push_back(e->cfg.eos_token);
  *n_out = v.size();
  if (v.size() > cap || (v.size() && !ids)) { set_error("stop ids: need room for %zu", v.size()); return B200_ERANGE; }
  if (!v.empty()) memcpy(ids, v.data(), v.size() * 4);
  return 0;
}

int b200_special_token(b200_engine* h, const char* text) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !text) return -1;
  return e->tok.special_id(text);
}

// ---- GGUF ----------------------------------------------------------------------------------------------
struct TokHandle { Tokenizer tok; int32_t vocab = 1 << 30; };

static int gguf_config(const GgufFile& f, b200_config* cfg, int64_t* n_tensors) {
  const std::string arch = f.get_s("general.architecture", "");
  if (arch != "llama") { set_error("GGUF architecture '%s' is not supported (llama only)", arch.c_str()); return B200_EINVAL; }
  b200_llama31_8b(cfg);
  cfg->n_layers = (int32_t)f.get_u("llama.block_count", 0);
  cfg->d_model = (int32_t)f.get_u("llama.embedding_length", 0);
  cfg->n_heads = (int32_t)f.get_u("llama.attention.head_count", 0);
  cfg->n_kv_heads = (int32_t)f.get_u("llama.attention.head_count_kv", (uint64_t)cfg->n_heads);
  cfg->d_ffn = (int32_t)f.get_u("llama.feed_forward_length", 0);
  cfg->head_dim = (int32_t)f.get_u("llama.attention.key_length", cfg->n_heads ? (uint64_t)(cfg->d_model / cfg->n_heads) : 0);
  cfg->rms_eps = (float)f.get_f("llama.attention.layer_norm_rms_epsilon", 1e-5);
  cfg->rope_theta = (float)f.get_f("llama.rope.freq_base", 10000.0);
  const GgufTensor* emb = f.tensor("token_embd.weight");
  const GgufValue* toks = f.kv("tokenizer.ggml.tokens");
  cfg->vocab = (int32_t)f.get_u("llama.vocab_size", toks ? toks->strs.size() : (emb && emb->ne.size() == 2 ? (uint64_t)emb->ne[1] : 0));
  // Llama-3.1 files carry the "llama3" RoPE scaling as a rope_freqs tensor; its parameters are the published ones
  if (f.tensor("rope_freqs.weight")) { cfg->rope_factor = 8.0f; cfg->rope_low_freq_factor = 1.0f; cfg->rope_high_freq_factor = 4.0f; cfg->rope_orig_ctx = 8192; }
  else cfg->rope_factor = 1.0f;
  cfg->eos_token = (int32_t)f.get_u("tokenizer.ggml.eos_token_id", 128009);
  cfg->seed = -1;
  const GgufTensor* q = f.tensor("blk.0.attn_q.weight");
  if (!q) { set_error("GGUF file has no blk.0.attn_q.weight"); return B200_EINVAL; }
  // a label only (the file dictates every tensor's storage type): 16-bit float files count as the bf16 recipe
  cfg->weight_format = (q->type == T_BF16 || q->type == T_F16) ? B200_WEIGHTS_BF16 : B200_WEIGHTS_Q4_K_M;
  if (n_tensors) *n_tensors = (int64_t)f.tensors().size();
  return 0;
}

int b200_gguf_info(const char* path, b200_config* cfg, int64_t* n_tensors, int64_t* n_vocab_tokens, int64_t* n_merges) {
  if (!path || !cfg) { set_error("null argument"); return B200_EINVAL; }
  GgufFile f;
  const std::string err = f.open(path);
  if (!err.empty()) { set_error("%s", err.c_str()); return B200_EINVAL; }
  const GgufValue* toks = f.kv("tokenizer.ggml.tokens");
  const GgufValue* mg = f.kv("tokenizer.ggml.merges");
  if (n_vocab_tokens) *n_vocab_tokens = toks ? (int64_t)toks->strs.size() : 0;
  if (n_merges) *n_merges = mg ? (int64_t)mg->strs.size() : 0;
  return gguf_config(f, cfg, n_tensors);
}

int b200_load_gguf(const char* path, int device, int max_ctx, int max_seqs, b200_engine** out) {
  if (!path || !out) { set_error("null argument"); return B200_EINVAL; }
  *out = nullptr;
  GgufFile f;
  const std::string err = f.open(path);
  if (!err.empty()) { set_error("%s", err.c_str()); return B200_EINVAL; }
  b200_config cfg;
  int rc = gguf_config(f, &cfg, nullptr);
  if (rc) return rc;
  cfg.device = device; cfg.max_ctx = max_ctx > 0 ? max_ctx : 2048; cfg.max_seqs = max_seqs > 0 ? max_seqs : 1;
  Engine* e = new (std::nothrow) Engine();
  if (!e) return B200_ENOMEM;
  for (const GgufTensor& t : f.tensors())
    if (t.ne.size() == 2) e->type_override[t.name] = (int)t.type;  // the file dictates every matrix's storage type
  if (f.kv("tokenizer.ggml.tokens")) {
    const std::string terr = e->tok.load(f);
    if (!terr.empty()) { set_error("%s", terr.c_str()); delete e; return B200_EINVAL; }
  }
  if ((rc = e->init(cfg))) { delete e; return rc; }
  {
    std::lock_guard<std::mutex> lk(e->mu);
    size_t maxb = 0;
    for (auto& kv : e->slots) maxb = std::max(maxb, (size_t)kv.second.full_rows * (kv.second.type == T_F32 ? (size_t)kv.second.full_cols * 4 : gemv_row_bytes(kv.second.type, kv.second.full_cols)));
    DevBuf tmp;
    if ((rc = tmp.alloc(maxb))) { delete e; return rc; }
    for (auto& kv : e->slots) {
      Slot& s = kv.second;
      const GgufTensor* t = f.tensor(kv.first);
      if (!t) { set_error("GGUF file lacks tensor '%s'", kv.first.c_str()); delete e; return B200_EINVAL; }
      const int64_t cols = t->ne.empty() ? 0 : t->ne[0], rows = t->ne.size() > 1 ? t->ne[1] : 1;
      if ((int)t->type != s.type || rows * cols != s.full_rows * s.full_cols || (s.full_rows != 1 && cols != s.full_cols)) {
        set_error("GGUF tensor '%s': type %u shape %lldx%lld, engine expects type %d shape %lldx%lld", kv.first.c_str(), t->type,
                  (long long)rows, (long long)cols, s.type, (long long)s.full_rows, (long long)s.full_cols);
        delete e;
        return B200_EINVAL;
      }
      cudaError_t ce = cudaMemcpyAsync(tmp.p, f.data(*t), t->nbytes, cudaMemcpyHostToDevice, e->stream);
      if (ce != cudaSuccess) { rc = cuda_fail(ce, "gguf upload"); delete e; return rc; }
      if ((rc = e->place_tensor(s, tmp.as<uint8_t>()))) { delete e; return rc; }
      ce = cudaStreamSynchronize(e->stream);
      if (ce != cudaSuccess) { rc = cuda_fail(ce, "gguf place"); delete e; return rc; }
      s.set = true;
    }
  }
  *out = reinterpret_cast<b200_engine*>(e);
  return 0;
}

int b200_tok_load_gguf(const char* path, b200_tok** out) {
  if (!path || !out) { set_error("null argument"); return B200_EINVAL; }
  GgufFile f;
  const std::string err = f.open(path);
  if (!err.empty()) { set_error("%s", err.c_str()); return B200_EINVAL; }
  TokHandle* h = new (std::nothrow) TokHandle();
  if (!h) return B200_ENOMEM;
  const std::string terr = h->tok.load(f);
  if (!terr.empty()) { set_error("%s", terr.c_str()); delete h; return B200_EINVAL; }
  h->vocab = (int32_t)h->tok.vocab_size();
  *out = reinterpret_cast<b200_tok*>(h);
  return 0;
}
void b200_tok_free(b200_tok* t) { delete reinterpret_cast<TokHandle*>(t); }
int b200_tok_encode(const b200_tok* t, const char* utf8, size_t n, int add_bos, int32_t* ids, size_t cap, size_t* n_out) {
  const TokHandle* h = reinterpret_cast<const TokHandle*>(t);
  if (!h || (!utf8 && n) || !n_out) { set_error("null argument"); return B200_EINVAL; }
  return tok_encode(h->tok, h->vocab, utf8, n, add_bos, ids, cap, n_out);
}
int b200_tok_decode(const b200_tok* t, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out) {
  const TokHandle* h = reinterpret_cast<const TokHandle*>(t);
  if (!h || (!ids && n) || !n_out) { set_error("null argument"); return B200_EINVAL; }
  return tok_decode(h->tok, ids, n, buf, cap, n_out);
}

int b200_pretokenize(const char* utf8, size_t n, int32_t* ends, size_t cap, size_t* n_out) {
  if ((!utf8 && n) || !n_out || (!ends && cap)) { set_error("null argument"); return B200_EINVAL; }
  std::vector<std::string> words;
  pretokenize(utf8, n, &words);
  *n_out = words.size();
  if (words.size() > cap) { set_error("output buffer too small (%zu pieces)", words.size()); return B200_EINVAL; }
  size_t off = 0;
  for (size_t i = 0; i < words.size(); ++i) { off += words[i].size(); ends[i] = (int32_t)off; }
  return 0;
}

int b200_submit(b200_engine* h, const int32_t* prompt_ids, size_t n, const b200_sampling* s, b200_req** out) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !prompt_ids || !out || n == 0) { set_error("null/empty argument"); return B200_EINVAL; }
  if (n > (size_t)e->cfg.max_ctx) { set_error("prompt of %zu tokens exceeds max_ctx %d", n, e->cfg.max_ctx); return B200_EINVAL; }
  for (size_t i = 0; i < n; ++i)
    if (prompt_ids[i] < 0 || prompt_ids[i] >= e->cfg.vocab) { set_error("token id %d out of range", prompt_ids[i]); return B200_EINVAL; }
  if (!e->all_set()) { set_error("not all tensors are set"); return B200_ESTATE; }
  Request* r = new (std::nothrow) Request();
  if (!r) return B200_ENOMEM;
  r->e = e;
  r->prompt.assign(prompt_ids, prompt_ids + n);  // copied: the caller's buffer is free after return
  if (s) r->sp = *s; else b200_default_sampling(&r->sp);
  r->t_submit = std::chrono::steady_clock::now();
  {
    std::lock_guard<std::mutex> lk(e->qmu);
    e->queue.push_back(r);
  }
  e->qcv.notify_one();
  *out = reinterpret_cast<b200_req*>(r);
  return 0;
}

int b200_next(b200_req* h, int32_t* id, int* done, int timeout_ms) {
  Request* r = reinterpret_cast<Request*>(h);
  if (!r || !id || !done) { set_error("null argument"); return B200_EINVAL; }
  std::unique_lock<std::mutex> lk(r->mu);
  if (r->done_delivered) { set_error("request already finished"); return B200_ESTATE; }
  auto ready = [&] { return !r->tokens.empty() || r->finished; };
  if (timeout_ms < 0) r->cv.wait(lk, ready);
  else if (!r->cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready)) { set_error("timeout"); return B200_ETIMEOUT; }
  if (r->tokens.empty()) {  // finished without (more) tokens: error or cancellation
    r->done_delivered = true;
    *done = 1;
    *id = -1;
    if (r->error) return r->error;
    set_error("request ended without a token");
    return B200_ESTATE;
  }
  *id = r->tokens.front();
  r->tokens.pop_front();
  *done = (r->complete && r->tokens.empty()) ? 1 : 0;
  if (*done) r->done_delivered = true;
  return 0;
}

int b200_stats_get(b200_req* h, b200_stats* out) {
  Request* r = reinterpret_cast<Request*>(h);
  if (!r || !out) { set_error("null argument"); return B200_EINVAL; }
  std::unique_lock<std::mutex> lk(r->mu);
  // the counters are final once the worker has drained the stream; that is moments after the last token
  if (r->complete || r->cancelled) r->cv.wait(lk, [&] { return r->finished; });
  *out = r->stats;
  return 0;
}

void b200_cancel(b200_req* h) {
  Request* r = reinterpret_cast<Request*>(h);
  if (!r) return;
  std::lock_guard<std::mutex> lk(r->mu);
  r->cancelled = true;
}

void b200_free(b200_req* h) {
  Request* r = reinterpret_cast<Request*>(h);
  if (!r) return;
  {
    std::unique_lock<std::mutex> lk(r->mu);
    r->cancelled = true;
    r->cv.wait(lk, [&] { return r->finished; });
  }
  delete r;
}

int b200_forward_logits(b200_engine* h, const int32_t* ids, size_t n, int mode, float* logits) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !ids || !logits || n == 0) { set_error("null/empty argument"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  B200_CUDA(cudaSetDevice(e->cfg.device));
  return e->forward_logits(ids, (int)n, mode, logits);
}

int b200_forward_logits_forced(b200_engine* h, const int32_t* prompt, size_t n_prompt, int prefill_mode, const int32_t* forced, size_t n_forced,
                               float* logits) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !prompt || !logits || n_prompt == 0 || (n_forced && !forced)) { set_error("null/empty argument"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  B200_CUDA(cudaSetDevice(e->cfg.device));
  return e->forward_logits_forced(prompt, (int)n_prompt, prefill_mode, forced, (int)n_forced, logits);
}

int b200_generate_greedy(b200_engine* h, const int32_t* prompt, size_t n_prompt, int32_t* ids, size_t n_new, int prefill_mode,
                         float* ttft_ms, float* decode_ms) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !prompt || !ids || n_prompt == 0 || n_new == 0) { set_error("null/empty argument"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  B200_CUDA(cudaSetDevice(e->cfg.device));
  return e->generate_greedy(prompt, (int)n_prompt, ids, (int)n_new, prefill_mode, ttft_ms, decode_ms);
}

int b200_dequant(int type, const void* blocks, int64_t n_blocks, float* out) {
  int rc = have_device();
  if (rc) return rc;
  if (type != T_Q4_K && type != T_Q6_K) { set_error("b200_dequant: type must be Q4_K or Q6_K"); return B200_EINVAL; }
  if (n_blocks < 0 || (n_blocks && (!blocks || !out))) { set_error("null argument"); return B200_EINVAL; }
  if (n_blocks == 0) return 0;
  const size_t inb = (size_t)n_blocks * (type == T_Q4_K ? Q4K_BYTES : Q6K_BYTES);
  DevBuf din, dout;
  if ((rc = din.alloc(inb)) || (rc = dout.alloc((size_t)n_blocks * QK_K * 4))) return rc;
  B200_CUDA(cudaMemcpy(din.p, blocks, inb, cudaMemcpyHostToDevice));
  B200_CUDA(launch_dequant(type, din.p, n_blocks, dout.as<float>(), nullptr, 0));
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(out, dout.p, (size_t)n_blocks * QK_K * 4, cudaMemcpyDeviceToHost));
  return 0;
}

int b200_quant_q16(const float* x, int64_t n, int16_t* q, float* d) {
  int rc = have_device();
  if (rc) return rc;
  if (n < 0 || n % 32 || (n && (!x || !q || !d))) { set_error("b200_quant_q16: n must be a multiple of 32"); return B200_EINVAL; }
  if (n == 0) return 0;
  DevBuf dx, dq, dd;
  if ((rc = dx.alloc((size_t)n * 4)) || (rc = dq.alloc((size_t)n * 2)) || (rc = dd.alloc((size_t)n / 32 * 4))) return rc;
  B200_CUDA(cudaMemcpy(dx.p, x, (size_t)n * 4, cudaMemcpyHostToDevice));
  B200_CUDA(launch_quant_q16(dx.as<float>(), n, dq.as<int16_t>(), dd.as<float>(), 0));
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(q, dq.p, (size_t)n * 2, cudaMemcpyDeviceToHost));
  B200_CUDA(cudaMemcpy(d, dd.p, (size_t)n / 32 * 4, cudaMemcpyDeviceToHost));
  return 0;
}

int b200_gemv(int type, const void* w, int64_t rows, int64_t cols, const float* x, float* y) {
  return gemv_hook(type, w, rows, cols, x, nullptr, 0.f, y);
}
int b200_gemv_rmsnorm(int type, const void* w, int64_t rows, int64_t cols, const float* x, const float* norm_w, float eps, float* y) {
  if (!norm_w) { set_error("null norm_w"); return B200_EINVAL; }
  return gemv_hook(type, w, rows, cols, x, norm_w, eps, y);
}

int b200_attn_decode(const float* q, const float* k, const float* v, int n_ctx, int n_heads, int n_kv_heads, float* out) {
  int rc = have_device();
  if (rc) return rc;
  if (!q || !k || !v || !out || n_ctx < 1 || n_kv_heads < 1 || n_heads % n_kv_heads) { set_error("bad argument"); return B200_EINVAL; }
  const int G = n_heads / n_kv_heads;
  if (G != 1 && G != 2 && G != 4 && G != 8) { set_error("GQA group %d unsupported", G); return B200_EINVAL; }
  const int kvd = n_kv_heads * kHeadDim, n_pages = (n_ctx + kPageTokens - 1) / kPageTokens, S = 20;
  // build a paged bf16 cache with a shuffled (non-identity) block table so the indirection is exercised
  std::vector<int32_t> bt(n_pages);
  for (int i = 0; i < n_pages; ++i) bt[i] = (int32_t)((i * 7 + 3) % n_pages);
  if (n_pages % 7 == 0) for (int i = 0; i < n_pages; ++i) bt[i] = n_pages - 1 - i;
  const size_t cache_elems = (size_t)n_pages * n_kv_heads * kPageTokens * kHeadDim;
  DevBuf dkf, dvf, dkc, dvc, dq, dout, dbt, dpart, dcnt;
  std::vector<float> kp(cache_elems, 0.f), vp(cache_elems, 0.f);
  for (int t = 0; t < n_ctx; ++t)
    for (int hh = 0; hh < n_kv_heads; ++hh) {
      const size_t dst = (((size_t)bt[t / kPageTokens] * n_kv_heads + hh) * kPageTokens + (t % kPageTokens)) * kHeadDim;
      memcpy(&kp[dst], k + (size_t)t * kvd + (size_t)hh * kHeadDim, kHeadDim * 4);
      memcpy(&vp[dst], v + (size_t)t * kvd + (size_t)hh * kHeadDim, kHeadDim * 4);
    }
  if ((rc = dkf.alloc(cache_elems * 4)) || (rc = dvf.alloc(cache_elems * 4)) || (rc = dkc.alloc(cache_elems * 2)) || (rc = dvc.alloc(cache_elems * 2)) ||
      (rc = dq.alloc((size_t)n_heads * kHeadDim * 4)) || (rc = dout.alloc((size_t)n_heads * kHeadDim * 4)) || (rc = dbt.alloc((size_t)n_pages * 4)) ||
      (rc = dpart.alloc(attn_partial_floats(n_heads, n_kv_heads, S) * 4)) || (rc = dcnt.alloc((size_t)n_kv_heads * 4)))
    return rc;
  B200_CUDA(cudaMemcpy(dkf.p, kp.data(), cache_elems * 4, cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemcpy(dvf.p, vp.data(), cache_elems * 4, cudaMemcpyHostToDevice));
  B200_CUDA(launch_f32_to_bf16(dkf.as<float>(), dkc.as<uint16_t>(), (int64_t)cache_elems, 0));
  B200_CUDA(launch_f32_to_bf16(dvf.as<float>(), dvc.as<uint16_t>(), (int64_t)cache_elems, 0));
  B200_CUDA(cudaMemcpy(dq.p, q, (size_t)n_heads * kHeadDim * 4, cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemcpy(dbt.p, bt.data(), (size_t)n_pages * 4, cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemset(dcnt.p, 0, (size_t)n_kv_heads * 4));
  AttnArgs a = {};
  a.q = dq.as<float>(); a.k_cache = dkc.as<uint16_t>(); a.v_cache = dvc.as<uint16_t>(); a.block_table = dbt.as<int32_t>();
  a.n_heads = n_heads; a.n_kv_heads = n_kv_heads; a.out = dout.as<float>(); a.partial = dpart.as<float>();
  a.counters = dcnt.as<unsigned int>(); a.n_splits = S; a.fixed_ctx = n_ctx;
  B200_CUDA(launch_attn_decode(a, 0, false));
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(out, dout.p, (size_t)n_heads * kHeadDim * 4, cudaMemcpyDeviceToHost));
  return 0;
}

int b200_attn_decode_noalloc(const float* q_dev, const uint16_t* k_cache_dev, const uint16_t* v_cache_dev, const int32_t* block_table_dev,
                             int n_ctx, int n_heads, int n_kv_heads, float* out_dev, void* stream) {
  int rc0 = have_device();
  if (rc0) return rc0;
  if (!q_dev || !k_cache_dev || !v_cache_dev || !block_table_dev || !out_dev) { set_error("null argument"); return B200_EINVAL; }
  if (n_heads > attn_noalloc_max_heads()) {
    set_error("b200_attn_decode_noalloc: n_heads=%d exceeds the built-in scratch (B200_NOALLOC_MAX_HEADS=%d); use b200_attn_decode",
              n_heads, attn_noalloc_max_heads());
    return B200_ERANGE;
  }
  const int G = (n_kv_heads > 0 && n_heads % n_kv_heads == 0) ? n_heads / n_kv_heads : 0;
  if (n_ctx < 1 || n_heads < 1 || (G != 1 && G != 2 && G != 4 && G != 8)) {
    set_error("b200_attn_decode_noalloc: bad shape (n_ctx=%d n_heads=%d n_kv_heads=%d)", n_ctx, n_heads, n_kv_heads);
    return B200_EINVAL;
  }
  B200_CUDA(launch_attn_decode_noalloc(q_dev, k_cache_dev, v_cache_dev, block_table_dev, n_ctx, n_heads, n_kv_heads, out_dev,
                                       reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

// y != null: Y fp32 [n][rows].  h != null: ffn gate/up with the SwiGLU epilogue, H bf16 [n][rows/2].
static int gemm_common(int type, const void* w, int64_t rows, int64_t cols, const float* x, int64_t n, float* y, uint16_t* h) {
  int rc = have_device();
  if (rc) return rc;
  if (!w || !x || (!y && !h)) { set_error("null argument"); return B200_EINVAL; }
  if (type != T_BF16 && type != T_F16 && type != T_Q4_K && type != T_Q6_K) { set_error("b200_gemm: type must be BF16, F16, Q4_K or Q6_K"); return B200_EINVAL; }
  if (!gemm_shape_ok(rows, cols, n) || (type != T_BF16 && type != T_F16 && cols % QK_K)) {
    set_error("gemm shape rows=%lld (need %%8) cols=%lld (need %%64; %%256 for k-quants) n=%lld unsupported", (long long)rows, (long long)cols, (long long)n);
    return B200_EINVAL;
  }
  DevBuf dw, dw16, dxf, dx16, dy;
  const size_t wbytes = (size_t)rows * gemv_row_bytes(type, cols);
  if (h && rows % 2) { set_error("b200_gemm_swiglu: rows must be even (gate_i, up_i interleaved)"); return B200_EINVAL; }
  const int max_splits = 4;  // what the scratch below is sized for when the SwiGLU path has to split K
  if ((rc = dw.alloc(wbytes)) || (rc = dxf.alloc((size_t)n * cols * 4)) || (rc = dx16.alloc((size_t)n * cols * 2)) ||
      (rc = dy.alloc((size_t)n * rows * 4 * (h ? max_splits : 1))))
    return rc;
  B200_CUDA(cudaMemcpy(dw.p, w, wbytes, cudaMemcpyHostToDevice));
  const uint16_t* w16 = dw.as<uint16_t>();
  if (type != T_BF16) {  // GEMM numerics: k-quant weights are dequantised and rounded to bf16
    if ((rc = dw16.alloc((size_t)rows * cols * 2))) return rc;
    if (type == T_F16) B200_CUDA(launch_f16_to_bf16(dw.as<uint16_t>(), dw16.as<uint16_t>(), rows * cols, 0));
    else B200_CUDA(launch_dequant(type, dw.p, rows * cols / QK_K, nullptr, dw16.as<uint16_t>(), 0));
    w16 = dw16.as<uint16_t>();
  }
  B200_CUDA(cudaMemcpy(dxf.p, x, (size_t)n * cols * 4, cudaMemcpyHostToDevice));
  B200_CUDA(launch_f32_to_bf16(dxf.as<float>(), dx16.as<uint16_t>(), n * cols, 0));
  // scratch for the stream-K schedule, as the engine provides it (the heuristic in gemm.cu decides)
  DevBuf wsb, flb;
  GemmWorkspace wk = {};
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, 0));
  wk.sm_count = prop.multiProcessorCount;
  wk.ws_bytes = (size_t)wk.sm_count * 256 * 256 * 4;
  if ((rc = wsb.alloc(wk.ws_bytes)) || (rc = flb.alloc((size_t)(wk.sm_count + 8) * 4))) return rc;
  B200_CUDA(cudaMemset(flb.p, 0, (size_t)(wk.sm_count + 8) * 4));
  wk.ws = wsb.as<float>(); wk.flags = flb.as<unsigned int>();
  if (h) {
    DevBuf dh;
    if ((rc = dh.alloc((size_t)n * (rows / 2) * 2))) return rc;
    B200_CUDA(launch_gemm_bf16_swiglu(w16, rows, cols, dx16.as<uint16_t>(), n, dh.as<uint16_t>(), rows / 2, dy.as<float>(), max_splits, 0, &wk));
    B200_CUDA(cudaDeviceSynchronize());
    B200_CUDA(cudaMemcpy(h, dh.p, (size_t)n * (rows / 2) * 2, cudaMemcpyDeviceToHost));
    return 0;
  }
  B200_CUDA(launch_gemm_bf16(w16, rows, cols, dx16.as<uint16_t>(), n, dy.as<float>(), rows, 0, &wk));
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(y, dy.p, (size_t)n * rows * 4, cudaMemcpyDeviceToHost));
  return 0;
}

int b200_gemm(int type, const void* w, int64_t rows, int64_t cols, const float* x, int64_t n, float* y) {
  if (!y) { set_error("null argument"); return B200_EINVAL; }
  return gemm_common(type, w, rows, cols, x, n, y, nullptr);
}

int b200_gemm_swiglu(int type, const void* w, int64_t rows, int64_t cols, const float* x, int64_t n, uint16_t* h) {
  if (!h) { set_error("null argument"); return B200_EINVAL; }
  return gemm_common(type, w, rows, cols, x, n, nullptr, h);
}

int b200_sample(const float* logits, int vocab, const b200_sampling* sp, const int32_t* history, int n_hist, uint64_t step, int32_t* id) {
  int rc = have_device();
  if (rc) return rc;
  if (!logits || !sp || !id || vocab < 1 || vocab > (1 << 18) || n_hist < 0 || (n_hist && !history)) { set_error("bad argument"); return B200_EINVAL; }
  DevBuf dl, dp, dh, dout;
  SampleParams p0 = {};
  p0.temperature = sp->temperature; p0.top_p = sp->top_p; p0.repeat_penalty = sp->repeat_penalty; p0.top_k = sp->top_k;
  p0.repeat_last_n = sp->repeat_last_n; p0.seed = sp->seed;
  if ((rc = dl.alloc((size_t)vocab * 4)) || (rc = dp.alloc(sizeof p0)) || (rc = dh.alloc((size_t)n_hist * 4)) || (rc = dout.alloc(4))) return rc;
  B200_CUDA(cudaMemcpy(dl.p, logits, (size_t)vocab * 4, cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemcpy(dp.p, &p0, sizeof p0, cudaMemcpyHostToDevice));
  if (n_hist) B200_CUDA(cudaMemcpy(dh.p, history, (size_t)n_hist * 4, cudaMemcpyHostToDevice));
  SampleArgs a = {};
  a.logits = dl.as<float>(); a.vocab = vocab; a.params = dp.as<SampleParams>(); a.history = dh.as<int32_t>(); a.st = nullptr;
  a.out = dout.as<int32_t>(); a.n_hist = n_hist; a.step = step;
  DevBuf dpart;  // the engine's two-kernel path (vocab >= 4096); B200_SAMPLE_ONE_KERNEL=1: the single-CTA kernel
  if (!getenv("B200_SAMPLE_ONE_KERNEL")) {
    if ((rc = dpart.alloc(sample_part_bytes(vocab)))) return rc;
    a.part = dpart.as<unsigned long long>();
  }
  B200_CUDA(launch_sample(a, 0, false));
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(id, dout.p, 4, cudaMemcpyDeviceToHost));
  return 0;
}

int b200_bench_decode(b200_engine* h, int ctx, int n_steps, float* ms_per_step, int* launches_per_step) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || ctx < 1 || n_steps < 1 || ctx + n_steps > e->cfg.max_ctx) { set_error("bad argument (ctx + n_steps must fit max_ctx)"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  B200_CUDA(cudaSetDevice(e->cfg.device));
  for (auto& q : e->seqs)
    if (q.r) { set_error("engine is serving requests"); return B200_ESTATE; }
  e->bs1_owner = -1;
  e->invalidate_prefix_cache();
  int rc = e->set_active_slot(0);
  if (rc) return rc;
  if ((rc = e->ensure_graph())) return rc;
  // Start from a context of `ctx` tokens already "in" the cache (its contents do not change the bytes moved).
  std::vector<int32_t> p(1, 1);
  if ((rc = e->begin_sequence(p.data(), 1, n_steps + 3, 1))) return rc;
  StepState s0 = {};
  s0.pos = ctx - 1; s0.cur_token = 1; s0.n_prompt = 1; s0.max_out = n_steps + 3; s0.eos = -1; s0.greedy = 1;
  B200_CUDA(cudaMemcpy(e->st, &s0, sizeof s0, cudaMemcpyHostToDevice));
  if ((rc = e->launch_steps(3))) return rc;  // warm-up
  cudaEvent_t e0, e1;
  B200_CUDA(cudaEventCreate(&e0));
  B200_CUDA(cudaEventCreate(&e1));
  B200_CUDA(cudaEventRecord(e0, e->stream));
  rc = e->launch_steps(n_steps);
  cudaEventRecord(e1, e->stream);
  cudaError_t ce = cudaStreamSynchronize(e->stream);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (rc) return rc;
  if (ce != cudaSuccess) return cuda_fail(ce, "bench_decode sync");
  e->dump_trace();
  if (ms_per_step) *ms_per_step = ms / n_steps;
  if (launches_per_step) *launches_per_step = e->launches_per_step;
  return 0;
}

int b200_bench_batched(b200_engine* h, int n_seqs, int ctx, int n_steps, float* ms_per_step, int32_t* last_ids) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || n_seqs < 1 || n_seqs > e->cfg.max_seqs || ctx < 1 || n_steps < 1 || ctx + n_steps + 3 > e->cfg.max_ctx) {
    set_error("bad argument (n_seqs <= max_seqs, ctx + n_steps + 3 <= max_ctx)");
    return B200_EINVAL;
  }
  std::lock_guard<std::mutex> lk(e->mu);
  B200_CUDA(cudaSetDevice(e->cfg.device));
  for (auto& q : e->seqs)
    if (q.r) { set_error("engine is serving requests"); return B200_ESTATE; }
  e->bs1_owner = -1;
  e->invalidate_prefix_cache();
  return e->bench_batched(n_seqs, ctx, n_steps, ms_per_step, last_ids);
}

int b200_bench_kernel(b200_engine* h, const char* which, int iters, float* ms, int64_t* bytes) {
  Engine* e = reinterpret_cast<Engine*>(h);
  if (!e || !which || iters < 1) { set_error("bad argument"); return B200_EINVAL; }
  std::lock_guard<std::mutex> lk(e->mu);
  B200_CUDA(cudaSetDevice(e->cfg.device));
  const b200_config& c = e->lc;
  const int d = c.d_model, qd = c.n_heads * kHeadDim, ff = c.d_ffn;
  Layer& L = e->layers[0];
  if (!strncmp(which, "gemm_", 5)) {
    // "gemm_<matrix>@<tokens>": the tensor-core GEMM of layer 0's matrix on a synthetic bf16 activation tile
    // (bf16 models only); *bytes returns the FLOP count of one launch
    const char* at = strchr(which, '@');
    const int64_t n = at ? atoll(at + 1) : 512;
    const std::string m(which + 5, at ? (size_t)(at - which - 5) : strlen(which + 5));
    const uint16_t* W = nullptr;
    int64_t rows = 0, K = 0;
    int type = T_BF16;
    if (m == "gate_up") { W = (const uint16_t*)L.gate_up; rows = 2 * ff; K = d; type = L.gu_type; }
    else if (m == "down") { W = (const uint16_t*)L.down; rows = d; K = ff; type = L.down_type; }
    else if (m == "qkv") { W = (const uint16_t*)L.qkv; rows = qd + 2 * c.n_kv_heads * kHeadDim; K = d; type = L.v ? -1 : L.qkv_type; }
    else if (m == "attn_output") { W = (const uint16_t*)L.o; rows = d; K = qd; type = L.o_type; }
    else if (m == "lm_head") { W = (const uint16_t*)e->output; rows = c.vocab; K = d; type = e->output_type; }
    if (!W || type != T_BF16 || n < 1 || n > 4096 || !gemm_shape_ok(rows, K, n)) {
      set_error("bench '%s': unknown matrix, not bf16, or unsupported shape", which);
      return B200_EINVAL;
    }
    DevBuf xb, yb;
    int rc;
    if ((rc = e->ensure_prefill())) return rc;  // the stream-K workspace lives with the prefill buffers
    if ((rc = xb.alloc((size_t)n * K * 2)) || (rc = yb.alloc((size_t)8 * n * rows * 4))) return rc;
    B200_CUDA(launch_synth(T_BF16, xb.p, n, K, 1234, 77, 1.0f, e->stream));
    cudaEvent_t e0, e1;
    B200_CUDA(cudaEventCreate(&e0));
    B200_CUDA(cudaEventCreate(&e1));
    int splits = 8;
    for (int i = -3; i < iters; ++i) {
      if (i == 0) B200_CUDA(cudaEventRecord(e0, e->stream));
      if (m == "down" || m == "attn_output" || m == "qkv") {  // as the engine calls them: K may be split
        splits = 8;
        B200_CUDA(launch_gemm_bf16_splitk(W, rows, K, xb.as<uint16_t>(), n, yb.as<float>(), rows, n * rows, &splits, e->stream,
                                          e->gemm_ws.ws ? &e->gemm_ws : nullptr));
      } else {
        B200_CUDA(launch_gemm_bf16(W, rows, K, xb.as<uint16_t>(), n, yb.as<float>(), rows, e->stream, e->gemm_ws.ws ? &e->gemm_ws : nullptr));
      }
    }
    B200_CUDA(cudaEventRecord(e1, e->stream));
    B200_CUDA(cudaStreamSynchronize(e->stream));
    float total = 0.f;
    cudaEventElapsedTime(&total, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (ms) *ms = total / iters;
    if (bytes) *bytes = 2 * n * rows * K;
    return 0;
  }
  GemvArgs a = {};
  int type, pro, epi;
  if (!strcmp(which, "gate_up")) {
    a.W = L.gate_up; a.N = 2 * ff; a.K = d; a.x = e->x; a.norm_w = L.ffn_norm; a.eps = c.rms_eps; a.y = e->h;
    type = L.gu_type; pro = PRO_RMSNORM; epi = EPI_SWIGLU;
  } else if (!strcmp(which, "down")) {
    a.W = L.down; a.N = d; a.K = ff; a.x = e->h; a.y = e->att;  // y: scratch, never the live residual
    type = L.down_type; pro = PRO_COPY; epi = EPI_STORE;
  } else if (!strcmp(which, "attn_output")) {
    a.W = L.o; a.N = d; a.K = qd; a.x = e->att; a.y = e->q;
    type = L.o_type; pro = PRO_COPY; epi = EPI_STORE;
  } else if (!strcmp(which, "lm_head")) {
    a.W = e->output; a.N = c.vocab; a.K = d; a.x = e->x; a.norm_w = e->output_norm; a.eps = c.rms_eps; a.y = e->logits;
    type = e->output_type; pro = PRO_RMSNORM; epi = EPI_STORE;
  } else {
    set_error("unknown kernel '%s' (gate_up, down, attn_output, lm_head)", which);
    return B200_EINVAL;
  }
  if (!e->flush_buf) {
    e->flush_bytes = (size_t)256 << 20;  // 2x the 126 MB L2
    int rc = e->alloc(&e->flush_buf, e->flush_bytes);
    if (rc) return rc;
  }
  cudaEvent_t e0, e1;
  B200_CUDA(cudaEventCreate(&e0));
  B200_CUDA(cudaEventCreate(&e1));
  const int64_t wbytes = (int64_t)a.N * (int64_t)gemv_row_bytes(type, a.K);
  float total = 0.f;
  if (wbytes > (int64_t)200 << 20) {
    // the matrix alone is larger than the 126 MB L2 and is streamed evict-first, so consecutive launches cannot
    // hit in L2: time `iters` launches back to back between one pair of events (launch latency amortised)
    for (int i = 0; i < 3; ++i) B200_CUDA(launch_gemv(type, pro, epi, a, e->stream, false, e->sm_count));
    B200_CUDA(cudaEventRecord(e0, e->stream));
    for (int i = 0; i < iters; ++i) B200_CUDA(launch_gemv(type, pro, epi, a, e->stream, false, e->sm_count));
    B200_CUDA(cudaEventRecord(e1, e->stream));
    B200_CUDA(cudaStreamSynchronize(e->stream));
    cudaEventElapsedTime(&total, e0, e1);
  } else {
    for (int i = -3; i < iters; ++i) {  // smaller matrices: flush L2 (write 256 MB) before every timed launch
      B200_CUDA(launch_l2_flush(e->flush_buf, e->flush_bytes, e->stream));
      B200_CUDA(cudaEventRecord(e0, e->stream));
      B200_CUDA(launch_gemv(type, pro, epi, a, e->stream, false, e->sm_count));
      B200_CUDA(cudaEventRecord(e1, e->stream));
      B200_CUDA(cudaStreamSynchronize(e->stream));
      float t = 0.f;
      cudaEventElapsedTime(&t, e0, e1);
      if (i >= 0) total += t;
    }
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (ms) *ms = total / iters;
  if (bytes) *bytes = wbytes;
  return 0;
}

}  // extern "C"
