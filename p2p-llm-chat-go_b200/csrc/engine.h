// engine.h — the host-side runtime of libb200llm: weights, paged KV pool, the captured decode-step graph,
// and the request worker.  One Engine drives one GPU (data-parallel replicas = one Engine per process/GPU).
#pragma once
#include <cuda_runtime.h>

#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200llm.h"
#include "kernels.h"
#include "tokenizer.h"

namespace b200 {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);  // records the message, returns B200_ECUDA

#define B200_CUDA(expr)                                   \
  do {                                                    \
    cudaError_t _e = (expr);                              \
    if (_e != cudaSuccess) return cuda_fail(_e, #expr);   \
  } while (0)

// Where a GGUF-named tensor lives on the device: a row range of a (possibly fused) buffer.
struct Slot {
  int type = -1;
  int64_t rows = 0, cols = 0;
  uint8_t* base = nullptr;  // device buffer holding this tensor's rows
  int64_t row0 = 0;         // first row inside the buffer
  int64_t row_step = 1;     // 2 for the interleaved gate/up buffer
  // tensor parallelism: this slot holds the window [src_row0, +rows) x [src_col0, +cols) of the full tensor
  int64_t full_rows = 0, full_cols = 0, src_row0 = 0, src_col0 = 0;
  bool set = false;
};

struct Layer {
  // fused projection buffers (device layout of their type)
  uint8_t* qkv = nullptr;  int qkv_type = -1;   // [q;k;v] rows, or [q;k] when V has another type
  uint8_t* v = nullptr;    int v_type = -1;     // only when split from qkv
  uint8_t* o = nullptr;    int o_type = -1;
  uint8_t* gate_up = nullptr; int gu_type = -1; // rows interleaved gate_i, up_i
  uint8_t* down = nullptr; int down_type = -1;
  float* attn_norm = nullptr;
  float* ffn_norm = nullptr;
  uint16_t* k_cache = nullptr;  // [n_pages][n_kv][16][128]
  uint16_t* v_cache = nullptr;
  // bf16 K-major operands of the tensor-core path: the GEMV buffers themselves for bf16 models, dequantised
  // shadow copies (same fused row order) for k-quant models
  uint16_t *qkv16 = nullptr, *o16 = nullptr, *gu16 = nullptr, *down16 = nullptr;
};

struct Request;

struct NcclApi;  // dlopen'ed libnccl entry points (tp.cu)

struct Engine {
  b200_config cfg{};   // the full model
  b200_config lc{};    // this rank's shard of it (n_heads, n_kv_heads, d_ffn, vocab divided by tp_size)
  int tp_rank = 0, tp_size = 1;
  void* nccl_comm = nullptr;
  float* tp_partial = nullptr;  // [d_model] partial sums of the row-parallel projections (NCCL fallback path)
  // fused all-reduce over peer memory (batch-1 decode): see TpLink in kernels.h
  bool tp_fused = false;
  TpLink* tp_link = nullptr;      // device copy
  TpLink tp_link_host{};
  float* x_alt = nullptr;         // second residual buffer (consumers read one, publish into the other)
  std::vector<void*> tp_ipc_opened;
  TpBatch tp_batch{};             // batched-decode all-reduce over peer memory (see kernels.h)
  unsigned int tp_bseq = 0;       // all-reduces pushed so far (every rank counts the same calls)
  bool tp_batch_push = false;     // B200_TP_PUSH=1 (experiment: slower than NCCL's in-switch reduction at TP-8)
  int tp_setup_link();
  int tp_check_error();
  int tp_init(const void* unique_id);
  int tp_allreduce_sum(float* buf, size_t n, cudaStream_t s);
  int tp_allreduce_max_u64(unsigned long long* buf, cudaStream_t s);
  int tp_allreduce_max_u64_n(unsigned long long* buf, size_t n, cudaStream_t s);
  int tp_allgather(const float* send, float* recv, size_t n_per_rank, cudaStream_t s);
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  std::mutex mu;  // serialises every device-touching entry point

  std::map<std::string, int> type_override;  // per-tensor storage types dictated by a GGUF file (else: the recipe)
  Tokenizer tok;                              // vocabulary of the GGUF file, or the byte-level fallback
  std::map<std::string, Slot> slots;
  std::vector<Layer> layers;
  uint8_t* token_embd = nullptr; int embd_type = -1;
  uint8_t* output = nullptr;     int output_type = -1;
  float* output_norm = nullptr;
  std::vector<void*> allocs;
  int64_t weight_bytes = 0;  // bytes streamed per decoded token (matrices + norms)

  // activations and per-sequence state (batch-1 path)
  float *x = nullptr, *q = nullptr, *att = nullptr, *h = nullptr, *logits = nullptr, *rope = nullptr;
  float* attn_partial = nullptr;
  unsigned int* attn_counters = nullptr;
  int n_splits = 20;
  StepState* st = nullptr;
  SampleParams* sp_dev = nullptr;
  int32_t* prompt_dev = nullptr;
  int32_t* block_table = nullptr;  // [max_seqs][pages_per_seq]
  int pages_per_seq = 0;
  int32_t* out_tokens = nullptr;       // mapped pinned host
  int32_t* out_tokens_dev = nullptr;   // device alias of the above
  volatile int32_t* host_n_out = nullptr;
  int32_t* host_n_out_dev = nullptr;

  // prefill (GEMM path) work buffers, sized for max_ctx tokens, allocated on first use
  bool pf_ready = false;
  float *pf_x = nullptr, *pf_qkv = nullptr, *pf_q = nullptr, *pf_att = nullptr, *pf_proj = nullptr, *pf_gu = nullptr;
  uint16_t *pf_xb = nullptr, *pf_hb = nullptr;  // bf16 GEMM operands: normalised rows / attention output; ffn hidden
  float* pf_partial = nullptr;
  unsigned int* pf_counters = nullptr;
  int pf_splits = 4;
  int pf_rows = 0;             // rows the work buffers were sized for
  int pf_proj_splits = 1;      // split-K partials currently sitting in pf_proj (consumers add them) ...
  int64_t pf_proj_stride = 0;  // ... this many floats apart
  const float* pf_proj_src = nullptr;        // where they are (pf_proj, or a peer-memory receive buffer under TP)
  const unsigned int* pf_wait_flags = nullptr;  // peer-memory path: flags the consumer waits on, and for which value
  unsigned int pf_wait_seq = 0;
  // residual += partials; xb = bf16(rmsnorm(residual) * w): the consumer of whatever the last projection left behind
  int add_rmsnorm_proj(const float* w, int n, bool have_proj);
  int reduce_proj(int n, bool decode);  // after a row-parallel projection wrote pf_proj: TP all-reduce (push or NCCL); no-op on one GPU
  int64_t pf_proj_floats = 0;  // capacity of pf_proj
  // K splits the qkv / gate-up GEMMs may use for n rows: their outputs are [pf_rows][width] buffers, every partial
  // is [n][width] (these two are column-parallel under TP: partials stay local, so TP does not restrict them)
  int max_wide_splits(int n) const { const int f = pf_rows / (n < 1 ? 1 : n); return f < 1 ? 1 : (f > 16 ? 16 : f); }
  unsigned long long* sample_part = nullptr;  // scratch of the two-kernel sampler (per-slice top-k keys)
  GemmWorkspace gemm_ws = {};  // stream-K scratch of the tensor-core GEMMs (gate/up, batched lm_head)
  // most K splits a projection over n rows may use: every partial is [n][d_model] and all must fit pf_proj
  int max_proj_splits(int n) const {
    const int64_t fit = pf_proj_floats / ((int64_t)n * lc.d_model);
    return tp_size > 1 ? 1 : (int)(fit < 1 ? 1 : (fit > 16 ? 16 : fit));
  }

  // sequence slots (continuous batching): slot s statically owns block-table row s
  struct Seq {
    Request* r = nullptr;
    int n_prompt = 0, pos = 0, delivered = 0, max_out = 0;  // pos = position of `token` (not yet in the cache)
    int32_t token = 0;
    bool greedy = true;
    int reused = 0;                // prompt tokens whose K/V were found in the slot's cache (prefix reuse)
    std::vector<int32_t> gen;      // tokens generated so far (host copy; with the prompt = what the slot's cache holds)
    std::chrono::steady_clock::time_point t0, t_load, t_first;
  };
  std::vector<Seq> seqs;
  // Prefix reuse (a peer's successive suggestions share the chat history).  slot_kv[s] = the tokens whose
  // K/V the cache of slot s still holds from the request it served last, positions 0..size-1.  A new request takes the
  // free slot with the longest common prefix and prefills only the rest of its prompt.
  std::vector<std::vector<int32_t>> slot_kv;
  mutable std::mutex kv_mu;         // guards slot_kv (the worker updates it; b200_prefix_match reads it from any thread)
  bool use_prefix_cache = true;     // off: cfg.reserved bit 1, B200_NO_PREFIX_CACHE, tensor-parallel engines
  static constexpr int kMinReuse = 16;  // shorter common prefixes are not worth a different code path
  int common_prefix(int slot, const std::vector<int32_t>& prompt) const;
  void invalidate_prefix_cache(int slot = -1);  // -1: every slot (weights changed)
  int32_t* active_table = nullptr;  // block-table row used by the batch-1 graph and single-sequence prefill
  int bs1_owner = -1;               // slot whose state currently lives in StepState / prompt_dev
  bool force_batched = false;       // cfg.reserved bit 0: never use the batch-1 path for decode (tests)
  bool batched_ready = false;
  bool can_batch() const { return tp_size == 1 && cfg.max_seqs > 1; }  // may the scheduler run batched steps at all
  int32_t *b_tok = nullptr, *b_pos = nullptr, *b_ctx = nullptr, *b_slot = nullptr, *b_next = nullptr;  // device [max_seqs]
  int32_t* b_next_host = nullptr;   // pinned
  int32_t* b_in_host = nullptr;     // pinned [4][max_seqs]: token / position / context / slot of the step being enqueued
  // the batched step as a CUDA graph per batch size (the launch sequence depends on nothing else); {exec, kernel nodes}
  std::map<int, std::pair<cudaGraphExec_t, int>> batch_graphs;
  static constexpr int kBatchGraphAfter = 16;
  int batch_last = 0, batch_same = 0;  // batch size of the previous step, and for how many steps it has held
  bool use_batch_graph = true;      // off: B200_NO_BATCH_GRAPH, tensor parallelism (per-step sequence numbers, NCCL calls)
  int enqueue_batched_body(int B);  // layer stack + lm_head + per-row argmax over the rows described by b_tok/b_pos/b_ctx/b_slot
  int32_t* hist_dev = nullptr;      // [max_seqs][max_ctx] token history per slot
  SampleParams* sp_arr = nullptr;   // [max_seqs]
  float* b_logits = nullptr;        // [max_seqs][vocab]
  uint16_t* output16 = nullptr;     // bf16 lm_head operand of the tensor-core path

  cudaGraphExec_t step_graph = nullptr;
  int step_graph_nodes = 0;  // kernel nodes of the captured step (counted as launches at every replay)
  bool use_pdl = true, use_graph = true, use_fa_prefill = true, use_mma_decode = true, use_attn_prefetch = true, use_attn_cluster = false;
  // n_seqs sequences decoding together from a context of ctx tokens, synchronously (no scheduler thread): the
  // batched step of scheduler.cu, tensor-parallel aware (every rank makes the same call)
  int bench_batched(int n_seqs, int ctx, int n_steps, float* ms_per_step, int32_t* last_ids);
  unsigned long long* b_keys = nullptr;  // [max_seqs] argmax keys (vocab-parallel lm_head)
  int launches_per_step = 0;
  int l2_chain_mb = -1;  // L2 prefetch chain: MB of the next kernel's weights (-1: per-format default, 0: off)
  unsigned long long* trace_dev = nullptr;   // B200_TRACE=<file>: 4 stamps per launch of the decode step (see GemvArgs::trace)
  int dump_trace();                          // the last step's stamps -> "<launch> <name> t0 t1 t2 t3" lines (ns)
  std::vector<std::string> trace_names;
  void* flush_buf = nullptr;
  size_t flush_bytes = 0;

  // request worker
  std::thread worker;
  std::mutex qmu;
  std::condition_variable qcv;
  std::deque<Request*> queue;
  bool stopping = false;

  ~Engine();
  int init(const b200_config& c, const void* nccl_id = nullptr);
  int alloc(void** p, size_t bytes);
  int place_tensor(const Slot& s, const uint8_t* dev_src);      // GGUF-layout rows on device -> slot
  int fetch_tensor(const Slot& s, uint8_t* dev_dst);            // slot -> GGUF-layout rows on device
  int synthesize();
  bool all_set() const;

  // lm_head + sampler + finish_step (with_lm_head = false: the persistent kernel already ran the lm_head)
  int enqueue_tail(cudaStream_t s, bool pdl, int* n, bool consume_tp = false, bool with_lm_head = true);
  void lm_head_args(GemvArgs* lm, int* pro, bool consume_tp);
  int enqueue_step(cudaStream_t s);          // launches the kernels of one decode step (capturable)
  int ensure_graph();
  int launch_steps(int n);                   // n graph replays (or direct launches)
  int begin_sequence(const int32_t* prompt, int n_prompt, int max_out, int greedy, const b200_sampling* sp = nullptr);
  int generate_greedy(const int32_t* prompt, int n_prompt, int32_t* ids, int n_new, int prefill_mode, float* ttft_ms,
                      float* decode_ms);
  int forward_logits(const int32_t* ids, int n, int mode, float* logits_host);
  int forward_logits_forced(const int32_t* prompt, int n_prompt, int prefill_mode, const int32_t* forced, int n_forced, float* logits_host);
  int ensure_prefill();                      // shadow weights + work buffers of the GEMM path
  int refresh_shadow(const std::string& name);  // re-derive the bf16 shadow of one tensor (after b200_set_tensor)
  bool has_shadows = false;                  // k-quant model whose bf16 shadows have been built
  uint8_t* shadow_tmp = nullptr;             // staging of refresh_shadow (largest per-layer tensor, GGUF layout)
  uint16_t* shadow_tmp16 = nullptr;
  int prefill_gemm(int n_prompt, int pos0 = 0);  // prompt_dev[pos0..n) -> KV cache (positions below pos0 are already there), first token
  // the tensor-core layer stack over n rows: one sequence at positions 0..n-1 (arrays null) or one row per sequence
  // segs != null: the rows are several prompts laid end to end (batched admission); rows [row0, row0+n) of segment i
  // are positions pos0.. of the sequence in slot `slot` (pos_arr / slot_arr carry the same per row for RoPE / KV append)
  struct PrefillSeg { int row0, n, pos0, slot; };
  int forward_gemm(int n, const int32_t* tokens_dev, const int32_t* pos_arr, const int32_t* ctx_arr, const int32_t* slot_arr, int pos0 = 0,
                   const std::vector<PrefillSeg>* segs = nullptr);
  int32_t *pfb_tok = nullptr, *pfb_pos = nullptr, *pfb_slot = nullptr;  // device [pf_rows]: row -> token / position / slot
  int prefill_rows_cap() const;              // rows one batched prefill may hold
  int set_active_slot(int slot);
  int ensure_batched();
  int batched_step(const std::vector<int>& slots, std::vector<int32_t>* next);
  int admit(Request* r, int slot);           // prefill a new request into a slot, deliver its first token
  // several waiting requests prefilled in ONE pass of the layer stack (their prompts are the rows of the same GEMMs)
  void admit_batch(const std::vector<Request*>& rs, const std::vector<int>& slots);
  int decode_bs1_window(int slot);           // batch-1 graph steps for the only active sequence
  void finish_request(int slot, int rc, int reason);
  void deliver(int slot, const int32_t* toks, int n);
  // either prefill flavour; reuse > 0 (GEMM mode): the cache already holds positions [0, reuse) of this prompt
  int run_prompt(const int32_t* prompt, int n_prompt, int max_out, int greedy, int mode, const b200_sampling* sp = nullptr, int reuse = 0);
  void worker_loop();
};

struct Request {
  Engine* e = nullptr;
  std::vector<int32_t> prompt;
  b200_sampling sp{};
  std::mutex mu;
  std::condition_variable cv;
  std::deque<int32_t> tokens;
  bool finished = false, cancelled = false, done_delivered = false;
  bool complete = false;  // the final token (EOS / length) has been queued
  std::chrono::steady_clock::time_point t_submit;  // b200_submit: durations count from here, so they include queueing
  int error = 0;
  b200_stats stats{};
};

}  // namespace b200
