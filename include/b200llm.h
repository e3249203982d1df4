/* b200llm.h — C ABI of libb200llm.so: the B200-native (sm_100a) inference engine behind the
 * "Suggest a Reply" call of NajyFannoun/P2P-LLM-Chat-Go.
 *
 * What this boundary replaces.  The reference has exactly one inference call site,
 * web/streamlit_app.py:89-101 (`ai_suggest`): an HTTP POST of {"model","prompt","stream":false} to
 * {OLLAMA_URL}/api/generate whose JSON answer's "response" field is shown to the user.  The server behind
 * that URL (Ollama -> llama.cpp, README.md:62-70) is external to the reference.  go/cmd/llmd (and its
 * runnable C++ twin, csrc/llmd.cpp) serve that same HTTP contract and call *this* header through cgo;
 * every entry point below is what such an FFI binds.  INTEGRATION.md shows the cgo and ctypes stubs.
 *
 * Conventions: plain C types only; return 0 on success, a negative B200_E* code on failure (message via
 * b200_last_error(), thread-local); never aborts, never throws across the boundary; input buffers are
 * copied before the call returns (cgo pointer rules); the library owns b200_engine / b200_req handles until
 * the matching free; no callbacks into the host language (pull model).  All entry points are thread-safe;
 * calls on one engine are serialised internally.  There is NO CPU fallback: without a CUDA device every
 * compute entry point fails with B200_ENODEV.
 */
#ifndef B200LLM_H
#define B200LLM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

/* error codes */
#define B200_OK 0
#define B200_EINVAL (-1)  /* bad argument / unsupported shape            */
#define B200_ENODEV (-2)  /* no CUDA device / driver                     */
#define B200_ECUDA (-3)   /* CUDA runtime error (see b200_last_error)    */
#define B200_ENOMEM (-4)  /* device or KV-pool exhaustion                */
#define B200_ESTATE (-5)  /* call not valid in the current state         */
#define B200_ETIMEOUT (-6)
#define B200_ERANGE (-7)  /* output buffer too small / size beyond a stated maximum */

/* tensor storage types: ggml / GGUF type ids, so a GGUF loader passes them through unchanged */
#define B200_TYPE_F32 0
#define B200_TYPE_F16 1   /* as stored by F16 GGUF files; the tensor-core path rounds it to bf16 */
#define B200_TYPE_Q4_K 12
#define B200_TYPE_Q6_K 14
#define B200_TYPE_BF16 30

/* weight recipes for synthetic / whole-model set-up */
#define B200_WEIGHTS_BF16 0   /* every matrix bf16 (BASELINE config 2)                                    */
#define B200_WEIGHTS_Q4_K_M 1 /* Ollama's default mix: Q4_K, with Q6_K for output, and attn_v / ffn_down   */
                              /* in layers i<n/8 || i>=7n/8 || (i-n/8)%3==2  (BASELINE config 3)           */

/* numerics modes of a forward pass (DESIGN.md "Numerics contract"; the CPU oracle implements both) */
#define B200_MODE_GEMV 0 /* batch-1 path: fp32 activations, fused-dequant GEMV kernels        */
#define B200_MODE_GEMM 1 /* prefill / batched path: bf16 GEMM inputs on tcgen05, fp32 accum   */

typedef struct b200_engine b200_engine;
typedef struct b200_req b200_req;

/* Model + engine description.  Replaces what `ollama pull llama3.1` + the Ollama model loader establish
 * for the reference (README.md:70, web/streamlit_app.py:28 LLM_MODEL). */
typedef struct b200_config {
  int32_t n_layers, d_model, n_heads, n_kv_heads, head_dim, d_ffn, vocab, max_ctx;
  float rms_eps, rope_theta;
  float rope_factor, rope_low_freq_factor, rope_high_freq_factor; /* "llama3" scaling; factor<=1: off */
  int32_t rope_orig_ctx;
  int32_t weight_format; /* B200_WEIGHTS_*                                                        */
  int32_t device;        /* CUDA ordinal                                                          */
  int32_t max_seqs;      /* concurrently resident sequences (sizes the paged KV pool); >=1        */
  int64_t seed;          /* >=0: synthesise deterministic random-init weights on the device;      */
                         /* <0 : tensors arrive through b200_set_tensor                           */
  int32_t eos_token;     /* generation stops on this id (<0: never)                               */
  int32_t reserved;      /* bit 0: decode every step on the batched tensor-core path, even for a single active   */
                         /* sequence (makes the numerics of concurrent requests independent of arrival timing)    */
                         /* bit 1: no prefix reuse (by default a request whose prompt starts with the tokens a   */
                         /* free slot's KV cache still holds is prefilled from the first differing token only)    */
} b200_config;

/* Sampling options = Ollama's `options` object (the reference sends none, web/streamlit_app.py:91-95, so
 * Ollama's defaults apply: temperature 0.8, top_k 40, top_p 0.9, repeat_penalty 1.1, repeat_last_n 64). */
typedef struct b200_sampling {
  float temperature; /* <=0: greedy */
  float top_p;
  float repeat_penalty;
  int32_t top_k;         /* <=0: whole vocabulary (capped at B200_MAX_TOP_K) */
  int32_t repeat_last_n; /* 0: off, <0: whole context */
  uint64_t seed;
  int32_t num_predict;   /* max new tokens (>0) */
  int32_t prefill_mode;  /* B200_MODE_GEMM (default, tensor cores) or B200_MODE_GEMV (token by token) */
} b200_sampling;
#define B200_MAX_TOP_K 256

/* Per-request counters = the *_duration / *_count fields of Ollama's final JSON object (nanoseconds).
 * Time runs from b200_submit: load_ns = wait for a sequence slot + set-up, prompt_eval_ns = prefill up to the first
 * token (so load_ns + prompt_eval_ns is the time to first token the caller sees), eval_ns = first to last token.
 * prompt_eval_count = prompt tokens actually evaluated (less than the prompt when a cached prefix was reused). */
typedef struct b200_stats {
  int64_t total_ns, load_ns, prompt_eval_ns, eval_ns;
  int32_t prompt_eval_count, eval_count;
  int32_t done_reason; /* 0 running, 1 "stop" (EOS), 2 "length", 3 cancelled */
  int32_t reserved;
} b200_stats;

/* ---- library ---------------------------------------------------------------------------------------- */
B200_API const char* b200_last_error(void);
B200_API const char* b200_version(void);
B200_API int b200_device_count(void);
/* Kernels of this library launched on a GPU by this process so far (graph replays count their kernel nodes):
 * what bench.py reports as gpu_launches, and a cheap way for a host to see that work really ran on the device. */
B200_API unsigned long long b200_launch_count(void);
B200_API void b200_default_sampling(b200_sampling* s);          /* Ollama defaults */
B200_API void b200_llama31_8b(b200_config* cfg);                /* shapes of Llama-3.1-8B   */
B200_API void b200_llama31_70b(b200_config* cfg);               /* shapes of Llama-3.1-70B  */

/* ---- engine life cycle (replaces: Ollama model load, README.md:62-70) ------------------------------- */
B200_API int b200_init(const b200_config* cfg, b200_engine** out);
/* Tensor parallelism (BASELINE config 5: Llama-3.1-70B over 8xB200).  One process/thread per GPU calls
 * b200_init_tp with the same full-model cfg, its rank, the degree, and the 128-byte NCCL id that rank 0 got
 * from b200_tp_unique_id (the host distributes it: torch.distributed, a file, ...).  Weights are sharded
 * column-/row-/vocab-parallel; o and down end in an NCCL all-reduce over NVLink.  Every rank must then issue
 * the same calls in the same order (SPMD): b200_set_tensor (full tensors; each rank keeps its window),
 * b200_forward_logits, b200_generate_greedy, b200_bench_decode.  bf16 weights, greedy decoding, max_seqs 1. */
B200_API int b200_tp_unique_id(void* id128);
B200_API int b200_init_tp(const b200_config* cfg, int tp_rank, int tp_size, const void* id128, b200_engine** out);
B200_API void b200_shutdown(b200_engine* e);
B200_API int b200_get_config(const b200_engine* e, b200_config* out);

/* Tensors in GGUF naming: token_embd.weight, blk.N.{attn_norm,attn_q,attn_k,attn_v,attn_output,ffn_norm,
 * ffn_gate,ffn_up,ffn_down}.weight, output_norm.weight, output.weight.  `data` is host memory in GGUF
 * row-major block layout; it is copied (and re-laid-out for the GPU) before the call returns. */
B200_API int b200_set_tensor(b200_engine* e, const char* name, int type, const void* data, int64_t rows,
                             int64_t cols);
/* Inverse of b200_set_tensor: the tensor back in GGUF layout (bit-identical to what was set/synthesised). */
B200_API int b200_tensor_info(const b200_engine* e, const char* name, int* type, int64_t* rows, int64_t* cols,
                              int64_t* nbytes);
B200_API int b200_get_tensor(b200_engine* e, const char* name, void* buf, int64_t cap);
B200_API int64_t b200_weight_bytes(const b200_engine* e); /* bytes streamed per decoded token (roofline) */

/* ---- text <-> ids (replaces: Ollama's tokenizer + template step of /api/generate) ------------------- */
B200_API int b200_tokenize(b200_engine* e, const char* utf8, size_t n, int add_bos, int32_t* ids, size_t cap,
                           size_t* n_out);
B200_API int b200_detokenize(b200_engine* e, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out);

/* What the model file says about chat framing (replaces: Ollama's per-model template and stop parameters, which it
 * derives from the same GGUF keys — the reference sends a bare prompt, web/streamlit_app.py:93, and relies on them).
 * b200_chat_template: the raw `tokenizer.chat_template` string ("" / n_out 0 when the engine has none).
 * b200_stop_ids: every token id that ends a reply (eos, eot, eom, end_of_text, im_end as present in the vocabulary).
 * b200_special_token: id of a control token by its text ("<|start_header_id|>"), -1 if the vocabulary has none. */
B200_API int b200_chat_template(b200_engine* e, char* buf, size_t cap, size_t* n_out);
B200_API int b200_stop_ids(b200_engine* e, int32_t* ids, size_t cap, size_t* n_out);
B200_API int b200_special_token(b200_engine* e, const char* text);
/* Longest prefix of `ids` whose K/V a sequence slot of this engine still holds from an earlier request (0 if none
 * worth reusing).  A front end that spreads sessions over several engines (one per GPU; the reference has a single
 * OLLAMA_URL, web/streamlit_app.py:27) uses it to send a peer's next suggestion where its chat history is cached. */
B200_API int b200_prefix_match(b200_engine* e, const int32_t* ids, size_t n);

/* ---- GGUF model files (what `ollama pull llama3.1` leaves on disk, README.md:70) ---------------------- */
/* Reads shapes, storage types and the tokenizer vocabulary from a GGUF v2/v3 file of architecture "llama"
 * and uploads every tensor (BF16, F16, Q4_K, Q6_K matrices; F32 norms; the file dictates each tensor's type).
 * b200_gguf_info needs no GPU.  A stand-alone tokenizer handle is offered for hosts that tokenise elsewhere. */
typedef struct b200_tok b200_tok;
B200_API int b200_gguf_info(const char* path, b200_config* cfg, int64_t* n_tensors, int64_t* n_vocab_tokens, int64_t* n_merges);
B200_API int b200_load_gguf(const char* path, int device, int max_ctx, int max_seqs, b200_engine** out);
B200_API int b200_tok_load_gguf(const char* path, b200_tok** out);
B200_API void b200_tok_free(b200_tok* t);
B200_API int b200_tok_encode(const b200_tok* t, const char* utf8, size_t n, int add_bos, int32_t* ids, size_t cap, size_t* n_out);
B200_API int b200_tok_decode(const b200_tok* t, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out);
/* The pre-tokenisation step on its own (needs no vocabulary): the Llama-3 ("llama-bpe") regex evaluated on Unicode code
 * points; ends[i] = byte offset one past piece i, pieces are consecutive and cover the text.  Replaces the regex split
 * Ollama / llama.cpp apply before BPE; exposed so tests can compare it with a regex engine. */
B200_API int b200_pretokenize(const char* utf8, size_t n, int32_t* ends, size_t cap, size_t* n_out);

/* ---- generation (replaces: the body of Ollama's /api/generate that web/streamlit_app.py:91 waits on) - */
B200_API int b200_submit(b200_engine* e, const int32_t* prompt_ids, size_t n, const b200_sampling* s,
                         b200_req** out);
/* Blocks the calling OS thread until the request's next token (or timeout_ms; <0 = forever).
 * *done is set with the last token; after done, further calls return B200_ESTATE. */
B200_API int b200_next(b200_req* r, int32_t* id, int* done, int timeout_ms);
B200_API int b200_stats_get(b200_req* r, b200_stats* out);
B200_API void b200_cancel(b200_req* r);
B200_API void b200_free(b200_req* r);

/* ---- test / parity hooks (host buffers in, host buffers out; each runs the production kernel) ------- */
/* Teacher-forced logits of the LAST position after running `n` tokens from an empty context. */
B200_API int b200_forward_logits(b200_engine* e, const int32_t* ids, size_t n, int mode, float* logits);
/* Teacher-forced logits along a given continuation: the prompt runs in `prefill_mode`, then forced[k] is fed at
 * position n_prompt + k through the batch-1 decode step (GEMV numerics) regardless of the model's own pick.
 * logits: (n_forced + 1) rows of `vocab` floats — row 0 after the prompt, row k + 1 after forced[k].  This is how the
 * parity test follows the CPU reference's 128 greedy tokens (BASELINE.json north_star) step by step. */
B200_API int b200_forward_logits_forced(b200_engine* e, const int32_t* prompt, size_t n_prompt, int prefill_mode,
                                        const int32_t* forced, size_t n_forced, float* logits);
/* Greedy generation with device-timed phases; fills ids[n_new]; ttft_ms/decode_ms may be NULL. */
B200_API int b200_generate_greedy(b200_engine* e, const int32_t* prompt, size_t n_prompt, int32_t* ids,
                                  size_t n_new, int prefill_mode, float* ttft_ms, float* decode_ms);

B200_API int b200_dequant(int type, const void* blocks, int64_t n_blocks, float* out); /* Q4_K / Q6_K */
B200_API int b200_quant_q16(const float* x, int64_t n, int16_t* q, float* d);          /* activations */
/* y[rows] = W[rows,cols] . x[cols] through the decode GEMV kernel of `type` (GEMV numerics). */
B200_API int b200_gemv(int type, const void* w, int64_t rows, int64_t cols, const float* x, float* y);
/* y = W . rmsnorm(x; norm_w, eps): the fused prologue variant. */
B200_API int b200_gemv_rmsnorm(int type, const void* w, int64_t rows, int64_t cols, const float* x,
                               const float* norm_w, float eps, float* y);
/* Y[n,rows] = X[n,cols] . W^T on the tcgen05/TMEM GEMM (GEMM numerics; W is BF16 or dequantised to it). */
B200_API int b200_gemm(int type, const void* w, int64_t rows, int64_t cols, const float* x, int64_t n, float* y);
/* The ffn gate/up GEMM with SwiGLU in its epilogue: rows of W interleaved gate_i, up_i (rows even);
   H[n, rows/2] = bf16( silu(X.W[2i]) * (X.W[2i+1]) ), bf16 bit patterns. */
B200_API int b200_gemm_swiglu(int type, const void* w, int64_t rows, int64_t cols, const float* x, int64_t n, uint16_t* h);
/* One query position of paged GQA decode attention; K,V given as fp32 [n_ctx][n_kv*128], stored as bf16. */
B200_API int b200_attn_decode(const float* q, const float* k, const float* v, int n_ctx, int n_heads,
                              int n_kv_heads, float* out);
/* The sampler kernel on one logits vector (history = previous tokens for the repeat penalty). */
B200_API int b200_sample(const float* logits, int vocab, const b200_sampling* s, const int32_t* history,
                         int n_hist, uint64_t step, int32_t* id);

/* ---- allocation-free entry point -------------------------------------------------------------------- */
/* No counterpart in the reference (the reference has no GPU code at all).
 * The same paged GQA decode-attention kernel as b200_attn_decode, for hosts that cannot let this library
 * allocate device memory: a process that has already given all of HBM to its own pool (an llmd replica whose
 * KV pool fills the GPU, a PyTorch caching allocator) would see an in-library cudaMalloc fail, and even a
 * successful cudaMalloc stalls every stream of the device.  Here the caller passes DEVICE pointers for all
 * inputs and the output, and the kernel's scratch lives in a fixed ~0.64 MB array inside the library image
 * (held by every GPU that loads the library until the process exits).  The call allocates and frees nothing.
 *   q_dev           [n_heads*128] fp32, already rotated
 *   k/v_cache_dev   paged bf16 cache [page][n_kv_heads][16][128]
 *   block_table_dev physical page of each logical 16-token page, int32[ceil(n_ctx/16)]
 *   out_dev         [n_heads*128] fp32
 *   stream          a cudaStream_t (NULL = default stream); the call is asynchronous on it
 * Limit: n_heads <= B200_NOALLOC_MAX_HEADS, larger shapes are refused with B200_ERANGE.
 * The first call on a device looks up the scratch address (with that device current) — make that first
 * call before starting any stream capture.  All calls on one GPU share the scratch: issue them in stream
 * order; calls that overlap on the GPU race. */
#define B200_NOALLOC_MAX_HEADS 64
B200_API int b200_attn_decode_noalloc(const float* q_dev, const uint16_t* k_cache_dev, const uint16_t* v_cache_dev,
                                      const int32_t* block_table_dev, int n_ctx, int n_heads, int n_kv_heads,
                                      float* out_dev, void* stream);

/* ---- bench hooks: device-resident steps timed with CUDA events on the engine's stream --------------- */
/* Runs `n_steps` decode steps (graph replays) starting from the current context; returns the mean step
 * time in milliseconds and the number of kernels launched per step. */
B200_API int b200_bench_decode(b200_engine* e, int ctx, int n_steps, float* ms_per_step, int* launches_per_step);
/* `n_seqs` sequences decoding together (the continuous-batching step on the tensor cores) from a context of
 * `ctx` tokens each, driven synchronously; mean step time in ms; last_ids (may be NULL) receives the n_seqs token
 * ids of the final step.  Tensor-parallel engines: every rank makes the same call.  No reference counterpart. */
B200_API int b200_bench_batched(b200_engine* e, int n_seqs, int ctx, int n_steps, float* ms_per_step, int32_t* last_ids);
/* Times only the dominant kernel (the fused gate/up GEMV of layer 0) `iters` times, L2 flushed between. */
B200_API int b200_bench_kernel(b200_engine* e, const char* which, int iters, float* ms, int64_t* bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200LLM_H */
