// kernels.h — host-callable launchers of the sm_100a kernels (decode path + helpers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// Kernels this library has put on a GPU in this process: every launcher below counts its launch, a CUDA-graph replay
// counts the kernel nodes of the graph; launches recorded during stream capture are not counted (they run at replay).
void count_launch(unsigned long long n = 1);
unsigned long long launch_count();
void count_launches_pause(bool on);  // capture in progress on this thread

constexpr int kHeadDim = 128;    // Llama-3.1 8B and 70B
constexpr int kPageTokens = 16;  // KV page = 16 tokens of one kv head = 4 KB (bf16)
constexpr int QK_K = 256, Q4K_BYTES = 144, Q6K_BYTES = 210;  // GGUF k-quant super-block

// ggml type ids (== B200_TYPE_*)
enum { T_F32 = 0, T_F16 = 1, T_Q4_K = 12, T_Q6_K = 14, T_BF16 = 30 };

enum { PRO_COPY = 0, PRO_RMSNORM = 1, PRO_TP_RMSNORM = 2 };
enum { EPI_STORE = 0, EPI_RESID = 1, EPI_SWIGLU = 2, EPI_QKV = 3, EPI_LOGITS = 4, EPI_TP_PUSH = 5 };

// Tensor-parallel link of one rank (device-resident): peers' receive buffers, mapped over NVLink.
// The row-parallel GEMVs (o, down) push their partial sums straight into every peer's buffer from their
// epilogue (EPI_TP_PUSH); the next GEMV's prologue (PRO_TP_RMSNORM) adds the partials in rank order to the residual
// stream and normalises — the all-reduce costs no launch of its own.  Every element is an 8-byte pair
// {value, sequence number of the exchange}, written with one 8-byte store (which arrives whole), so the consumer
// needs no separate flag: it polls the elements themselves until they carry its exchange's number.  (The first
// version pushed plain floats, then system-scope fence -> CTA ticket -> fence -> one release store per peer: ~11 us
// from the last partial sum to the consumer at TP-2, and a release store per peer on top at TP-8.)
constexpr int kMaxTp = 8;
struct TpLink {
  float* peer_recv[2][kMaxTp];      // [buffer][peer] -> that peer's recv buffer ([size][d] {float, seq} pairs, slot = source rank)
  unsigned int* peer_flags[kMaxTp]; // -> that peer's flags [2][kMaxTp] (sequence number per source rank)
  float* recv[2];                   // this rank's own buffers
  float* xsum[2];                   // local: the reduced residual stream as {value, seq} pairs, [buffer][d] (tp_prologue, phase 2)
  unsigned int* flags;              // this rank's own flags
  unsigned int* epoch;              // decode steps completed (advanced by finish_step)
  unsigned int* ticket;             // CTA ticket of the pushing kernel (self-resetting)
  int* error;                       // set if a wait timed out
  int rank, size, d, per_step;      // per_step = all-reduces per decode step (2 * n_layers)
};

// Batched decode (<= kTpBatchRows sequences) under tensor parallelism: the all-reduce of the [rows][d] partial
// products is a push over NVLink peer memory too.  tp_push_kernel copies this rank's partial into slot `rank` of
// every peer's receive buffer and raises a sequence flag there; the consumer (add_rmsnorm) waits for the `size`
// flags and adds the slots in rank order, which is the deterministic reduction it already does for split-K.
constexpr int kTpBatchRows = 32;
struct TpBatch {
  float* peer_recv[2][kMaxTp];       // [buffer][peer] -> that peer's [size][kTpBatchRows * d] floats
  unsigned int* peer_flags[kMaxTp];  // -> that peer's flags [2][kMaxTp]
  float* recv[2];                    // own buffers
  unsigned int* flags;               // own flags
  unsigned int* ticket;              // CTA ticket of the push kernel (self-resetting)
  int* error;
  int rank, size;
  long long slot_floats;             // kTpBatchRows * d
};
cudaError_t launch_tp_push(const TpBatch& b, const float* src, int64_t count, int buf, unsigned int seq, cudaStream_t s);

// Device-resident state of the sequence being decoded; every kernel of a step reads it, the last kernel
// of the step (finish_step) advances it, so one captured CUDA graph serves every position.
struct StepState {
  int32_t pos;        // position of the token being processed (== tokens already in the KV cache)
  int32_t cur_token;  // token id at `pos`
  int32_t n_prompt;   // prompt length; while pos+1 < n_prompt the next token is forced from the prompt
  int32_t n_out;      // tokens generated so far
  int32_t max_out;    // stop after this many generated tokens
  int32_t eos;        // stop token (<0: none)
  int32_t done;       // 0 running, 1 stop (EOS), 2 length
  int32_t greedy;     // 1: token = argmax key; 0: token comes from sample_kernel via `sampled`
  int32_t sampled;
  int32_t pad_;
  unsigned long long argmax_key;  // packed (ordered logit, ~index); reset every step
};

struct GemvArgs {
  const uint8_t* W;  // weights, device layout of the tensor's type
  int N, K;          // rows, cols
  const float* x;    // input vector [K], fp32
  const float* norm_w;  // PRO_RMSNORM: weight [K]
  float eps;
  float* y;  // STORE: y[N]; RESID: y[N] += W.x; SWIGLU: y[N/2]; QKV: q[q_rows]; LOGITS: logits[N]
  // EPI_QKV
  int row_base;  // logical row of W's row 0 in the fused [q;k;v] row space
  int q_rows, kv_rows, n_kv_heads;
  const float* rope;      // [max_ctx][64][2] cos,sin
  uint16_t* k_cache;      // this layer: [n_pages][n_kv][16][128] bf16 bits
  uint16_t* v_cache;
  const int32_t* block_table;  // physical page of each logical page of the active sequence
  const StepState* st;         // QKV: position; LOGITS: skip while prefilling
  // EPI_LOGITS
  int store_logits;
  unsigned long long* argmax_key;
  // L2 prefetch chaining: when a warp has streamed its last unit it asks L2 for a slice of the weights the
  // NEXT kernel of the step will stream, so that kernel starts on L2 hits while HBM would otherwise idle
  const uint8_t* next_W;
  unsigned long long next_bytes;
  // tensor parallelism over peer memory (null without it)
  const TpLink* tp;
  int tp_index;        // which all-reduce of the step this kernel produces (EPI_TP_PUSH) / consumes (PRO_TP_RMSNORM)
  float* x_out;        // PRO_TP_RMSNORM: where CTA 0 leaves the updated residual stream (x + sum of partials)
  // debug (B200_TRACE): 4 x %globaltimer of this launch — CTA 0 at entry, after the dependency wait, after the
  // prologue; and the latest exit over all CTAs (atomicMax).  null: no stamps
  unsigned long long* trace;
};

struct GemvPlan {  // launch geometry chosen on the host
  int grid, block;
  size_t smem;
};

// pdl: launch with programmatic stream serialization (kernel N+1 prefetches weights under kernel N's tail)
cudaError_t launch_gemv(int wtype, int pro, int epi, const GemvArgs& a, cudaStream_t s, bool pdl, int sm_count);
// two projections of the same input in one launch (Q4_K q,k + Q6_K v); cudaErrorNotSupported for other type pairs
cudaError_t launch_gemv_dual_qkv(int type_a, const GemvArgs& a, int type_b, const GemvArgs& b, cudaStream_t s, bool pdl, int sm_count);
size_t gemv_row_bytes(int wtype, int64_t K);
bool gemv_shape_ok(int wtype, int64_t N, int64_t K, int epi);

struct AttnArgs {
  const float* q;   // [n_heads*128] (already rotated)
  const uint16_t* k_cache;
  const uint16_t* v_cache;
  const int32_t* block_table;
  const StepState* st;  // ctx = st->pos + 1
  int n_heads, n_kv_heads;
  float* out;           // [n_heads*128]
  float* partial;       // scratch [n_kv][n_splits][group][130]
  unsigned int* counters;  // scratch [n_kv], must be zero before the first launch; self-resetting
  int n_splits;
  int fixed_ctx;        // >0: use this context length instead of st->pos+1 (test hook / prefill)
  // several query positions in one launch (prefill): query z has ctx = fixed_ctx + z, q/out advance by
  // q_stride floats, scratch by one query's worth.  n_queries <= 1: the single-query decode case.
  int n_queries;
  int q_stride;
  // batched decode: query z belongs to the sequence in slot q_slot[z] (block-table row) and sees q_ctx[z] tokens
  const int32_t* q_ctx;
  const int32_t* q_slot;
  int pages_per_seq;
  // batch-1 decode step: cache rows [0, prefetch_tokens) of the active sequence may be read before the preceding
  // kernel has finished (they exist and are not written by it, except the row of the current position, which is
  // re-read).  0: the plain kernel.
  int prefetch_tokens;
  unsigned long long* trace;  // debug (B200_TRACE), as in GemvArgs
  int use_cluster;            // batch-1 decode step: the 16-CTA cluster kernel (merge over distributed shared memory)
};
cudaError_t launch_attn_decode(const AttnArgs& a, cudaStream_t s, bool pdl);
size_t attn_partial_floats(int n_heads, int n_kv_heads, int n_splits);
bool attn_cluster_supported(int group);  // can the device schedule the 16-CTA cluster kernel for this GQA group size?
// allocation-free variant: scratch in a static __device__ array inside attn.cu (see there)
cudaError_t launch_attn_decode_noalloc(const float* q, const uint16_t* k_cache, const uint16_t* v_cache,
                                       const int32_t* block_table, int n_ctx, int n_heads, int n_kv_heads, float* out,
                                       cudaStream_t s);
int attn_noalloc_max_heads();

struct SampleParams {  // device copy of the request's sampling options
  float temperature, top_p, repeat_penalty;
  int32_t top_k, repeat_last_n;
  unsigned long long seed;
};
struct SampleArgs {
  const float* logits;
  int vocab;
  const SampleParams* params;
  const int32_t* history;  // every token of the sequence so far (prompt, then generated)
  const StepState* st;     // engine mode: n_hist = st->pos+1, step = st->n_out, no-op when st->greedy
  int32_t* out;
  int n_hist;              // >= 0: stand-alone mode (test hook) with explicit history length and step
  unsigned long long step;
  unsigned long long* part;  // scratch of sample_part_bytes(vocab): two-kernel path (null: one CTA does everything)
};
cudaError_t launch_sample(const SampleArgs& a, cudaStream_t s, bool pdl);  // 1 launch, or 2 when a.part is set and vocab >= 4096
size_t sample_part_bytes(int vocab);

struct FinishArgs {
  StepState* st;
  int32_t* prompt;           // device [max_ctx]: the prompt, extended with every generated token (history)
  int32_t* out_tokens;       // device-visible (mapped pinned host) [max_out]
  volatile int32_t* host_n_out;  // mapped pinned host counter
  const uint8_t* embd;       // token_embd in GGUF layout
  int embd_type;
  int d_model;
  float* x;                  // residual stream for the next step
  int max_ctx;
  unsigned int* tp_epoch;    // tensor parallelism over peer memory: step counter to advance (or null)
};
cudaError_t launch_finish_step(const FinishArgs& a, cudaStream_t s, bool pdl);
cudaError_t launch_embed(const uint8_t* embd, int embd_type, int d_model, const int32_t* token_dev, float* x,
                         cudaStream_t s);

// dequantisation of GGUF-layout blocks (bit-exact w.r.t. ggml / gguf-py)
cudaError_t launch_dequant(int type, const void* blocks, int64_t n_blocks, float* out_f32, uint16_t* out_bf16,
                           cudaStream_t s);
cudaError_t launch_quant_q16(const float* x, int64_t n, int16_t* q, float* d, cudaStream_t s);
// Q6_K rows: GGUF block order <-> device section order ([ql][qh][scales][d] per row)
cudaError_t launch_q6k_repack(const uint8_t* src, uint8_t* dst, int64_t rows, int64_t blocks_per_row, int to_device,
                              cudaStream_t s);
// deterministic random-init tensors (see synth.cu for the recipe)
cudaError_t launch_synth(int type, void* dst, int64_t rows, int64_t cols, uint64_t seed, uint32_t tensor_id,
                         float sigma, cudaStream_t s);
cudaError_t launch_f32_to_bf16(const float* x, uint16_t* y, int64_t n, cudaStream_t s);
cudaError_t launch_f16_to_bf16(const uint16_t* x, uint16_t* y, int64_t n, cudaStream_t s);  // fp16 bits -> bf16 bits (RNE): GEMM operand of F16 tensors
cudaError_t launch_l2_flush(void* buf, size_t bytes, cudaStream_t s);
cudaError_t launch_add_inplace(float* x, const float* y, int n, cudaStream_t s);

// tcgen05/TMEM bf16 GEMM (gemm.cu): Y[t*ldy + r] = sum_k X[t][k] W[r][k], fp32 out
// optional scratch for the stream-K schedule (gemm.cu): sm_count x 256 KB of partial tiles + one flag per SM
// (flags zeroed once by the owner of the workspace; the kernel re-arms them itself)
struct GemmWorkspace {
  float* ws;
  size_t ws_bytes;
  unsigned int* flags;
  int sm_count;
};
cudaError_t launch_gemm_bf16(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, float* Y,
                             int64_t ldy, cudaStream_t s, const GemmWorkspace* wk = nullptr);
cudaError_t launch_gemm_bf16_splitk(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, float* Y,
                                    int64_t ldy, int64_t split_stride, int* splits, cudaStream_t s, const GemmWorkspace* wk = nullptr);
// ffn gate/up (rows interleaved gate_i, up_i) with SwiGLU: H[t*ldh + i] bf16.  scratch: fp32 [max_splits][n_tokens][rows],
// used only when the schedule splits K (33..128 tokens); H must not overlap X.
cudaError_t launch_gemm_bf16_swiglu(const uint16_t* W, int64_t rows, int64_t K, const uint16_t* X, int64_t n_tokens, uint16_t* H, int64_t ldh,
                                    float* scratch, int max_splits, cudaStream_t s, const GemmWorkspace* wk = nullptr);
bool gemm_shape_ok(int64_t rows, int64_t K, int64_t n_tokens);

// causal attention over a whole prompt on the tensor cores (attn_prefill.cu); out is bf16 [T][n_heads*128]
cudaError_t launch_attn_prefill(const float* q, const uint16_t* k_cache, const uint16_t* v_cache, const int32_t* block_table, uint16_t* out,
                                int T, int pos0, int n_heads, int n_kv_heads, cudaStream_t s);
// several prompts in one launch: segment = (first row in q/out, rows, first position, cache slot)
struct AttnPrefillSegs {
  static constexpr int kMax = 16;
  int n, pages_per_seq;
  int4 seg[kMax];
};
cudaError_t launch_attn_prefill_segs(const float* q, const uint16_t* k_cache, const uint16_t* v_cache, const int32_t* block_table,
                                     int pages_per_seq, uint16_t* out, const int4* segs, int n_segs, int n_heads, int n_kv_heads, cudaStream_t s);

// batched decode (one query per sequence) on the tensor cores; bf16 output when out_bf16 is given, else a.out
cudaError_t launch_attn_decode_batched(const AttnArgs& a, uint16_t* out_bf16, cudaStream_t s);

// element-wise kernels of the prefill / batched pass (prefill.cu)
cudaError_t launch_embed_tokens(const uint8_t* embd, int embd_type, int d_model, const int32_t* tokens_dev, int n, float* x, cudaStream_t s);
// `add`: n_parts split-K partials of [tokens][d], part_stride floats apart, summed in ascending order
// wait_flags != null: every CTA first waits until wait_flags[0..n_wait) have reached wait_seq (tp_push_kernel)
cudaError_t launch_add_rmsnorm_bf16(float* x, const float* add, int n_parts, int64_t part_stride, const float* w, uint16_t* xb, int tokens,
                                    int d, float eps, cudaStream_t s, const unsigned int* wait_flags = nullptr, int n_wait = 0,
                                    unsigned int wait_seq = 0, int* wait_error = nullptr);
cudaError_t launch_rope_kv(const float* qkv, float* q, uint16_t* k_cache, uint16_t* v_cache, const int32_t* block_table, const float* rope,
                           int tokens, int q_rows, int kv_rows, int n_kv_heads, int pos0, const int32_t* pos_arr, const int32_t* slot_arr,
                           int pages_per_seq, int n_parts, int64_t part_stride, cudaStream_t s);  // qkv: n_parts split-K partials
cudaError_t launch_argmax_rows(const float* logits, int rows, int vocab, int32_t* next, cudaStream_t s);
// vocab-parallel: per-row key of the local maximum (global id = id_base + index), max-all-reduced, then decoded
cudaError_t launch_argmax_rows_keys(const float* logits, int rows, int vocab, int id_base, unsigned long long* keys, cudaStream_t s);
cudaError_t launch_keys_to_ids(const unsigned long long* keys, int32_t* next, int rows, cudaStream_t s);
cudaError_t launch_append_history(const int32_t* next, const int32_t* slot_arr, const int32_t* pos_arr, int32_t* history, int max_ctx, int rows,
                                  cudaStream_t s);
cudaError_t launch_swiglu_bf16(const float* gu, uint16_t* h, int64_t n_pairs, int n_parts, int64_t part_stride, cudaStream_t s);
cudaError_t launch_take_row(const float* x, const float* add, int n_parts, int64_t part_stride, int d, int row, float* out, cudaStream_t s);

}  // namespace b200
