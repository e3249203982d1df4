// scheduler.cu — the request worker: continuous batching over sequence slots.
//
// Every peer's suggestion is an independent request (web/streamlit_app.py:163-165), so the unit of
// scheduling is a sequence slot (its own run of KV pages).  One worker thread per engine owns the stream:
//   admit    : a waiting request gets a free slot, its prompt goes through the tensor-core prefill (or the
//              token-by-token GEMV path if it asked for it) and its first token is delivered;
//   decode   : exactly one active sequence and nobody waiting -> the batch-1 CUDA-graph path (GEMV kernels,
//              numerics mode 0); otherwise ONE batched step for all active sequences on the tensor cores
//              (weights stream once per step for the whole batch; numerics mode 1), then new arrivals are
//              admitted before the next step (iteration-level scheduling);
//   finish   : EOS / length / cancel frees the slot at once.
// A sequence moves between the two decode paths without copying anything: both use the same paged KV cache;
// only the small per-sequence state (position, last token, history) is handed over.
#include <algorithm>
#include <cstring>

#include "engine.h"

namespace b200 {

static int64_t ns_between(std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
  return (int64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(b - a).count();
}

int Engine::set_active_slot(int slot) {
  B200_CUDA(cudaMemcpyAsync(active_table, block_table + (size_t)slot * pages_per_seq, (size_t)pages_per_seq * 4, cudaMemcpyDeviceToDevice, stream));
  return 0;
}

int Engine::common_prefix(int slot, const std::vector<int32_t>& prompt) const {
  std::lock_guard<std::mutex> lk(kv_mu);
  const std::vector<int32_t>& have = slot_kv[slot];
  const size_t lim = std::min(have.size(), prompt.size() - 1);  // at least one prompt token is always evaluated
  size_t n = 0;
  while (n < lim && have[n] == prompt[n]) ++n;
  return (int)n;
}

void Engine::invalidate_prefix_cache(int slot) {
  std::lock_guard<std::mutex> lk(kv_mu);
  if (slot >= 0) { if (slot < (int)slot_kv.size()) slot_kv[slot].clear(); return; }
  for (auto& v : slot_kv) v.clear();
}

int Engine::ensure_batched() {
  if (batched_ready) return 0;
  int rc = ensure_prefill();
  if (rc) return rc;
  const b200_config& c = cfg;
  const size_t S = (size_t)c.max_seqs;
  if (!gemm_shape_ok(c.vocab, c.d_model, 1)) { set_error("vocab must be a multiple of 128 for the batched lm_head"); return B200_EINVAL; }
  if ((rc = alloc((void**)&b_tok, S * 4)) || (rc = alloc((void**)&b_pos, S * 4)) || (rc = alloc((void**)&b_ctx, S * 4)) ||
      (rc = alloc((void**)&b_slot, S * 4)) || (rc = alloc((void**)&b_next, S * 4)) || (rc = alloc((void**)&hist_dev, S * c.max_ctx * 4)) ||
      (rc = alloc((void**)&sp_arr, S * sizeof(SampleParams))) || (rc = alloc((void**)&b_logits, S * (size_t)c.vocab * 4)))
    return rc;
  B200_CUDA(cudaHostAlloc((void**)&b_next_host, S * 4, cudaHostAllocDefault));
  B200_CUDA(cudaHostAlloc((void**)&b_in_host, 4 * S * 4, cudaHostAllocDefault));
  if (output_type != T_BF16) {  // bf16 shadow of a k-quant lm_head
    const Slot& s = slots["output.weight"];
    if ((rc = alloc((void**)&output16, (size_t)s.rows * s.cols * 2))) return rc;
    has_shadows = true;
    if ((rc = refresh_shadow("output.weight"))) return rc;
  } else {
    output16 = (uint16_t*)output;
  }
  batched_ready = true;
  return 0;
}

int Engine::enqueue_batched_body(int B) {
  int rc = forward_gemm(B, b_tok, b_pos, b_ctx, b_slot);
  if (rc) return rc;
  if ((rc = add_rmsnorm_proj(output_norm, B, true))) return rc;
  B200_CUDA(launch_gemm_bf16(output16, cfg.vocab, cfg.d_model, pf_xb, B, b_logits, cfg.vocab, stream, &gemm_ws));
  B200_CUDA(launch_argmax_rows(b_logits, B, cfg.vocab, b_next, stream));
  return 0;
}

// One decode step for the sequences in `slots` (one row each): token seqs[s].token at position seqs[s].pos.
// The ~330 launches of the step depend only on the number of rows, so they are captured once per batch size and
// replayed; what changes from step to step (tokens, positions, context lengths, slots) sits in device arrays that are
// refreshed from pinned memory in front of the graph.
int Engine::batched_step(const std::vector<int>& sl, std::vector<int32_t>* next) {
  const b200_config& c = cfg;
  const int B = (int)sl.size();
  const size_t S = (size_t)c.max_seqs;
  int32_t *tok = b_in_host, *pos = b_in_host + S, *ctx = b_in_host + 2 * S, *slot = b_in_host + 3 * S;
  for (int i = 0; i < B; ++i) {
    const Seq& q = seqs[sl[i]];
    tok[i] = q.token; pos[i] = q.pos; ctx[i] = q.pos + 1; slot[i] = sl[i];
  }
  B200_CUDA(cudaMemcpyAsync(b_tok, tok, (size_t)B * 4, cudaMemcpyHostToDevice, stream));
  B200_CUDA(cudaMemcpyAsync(b_pos, pos, (size_t)B * 4, cudaMemcpyHostToDevice, stream));
  B200_CUDA(cudaMemcpyAsync(b_ctx, ctx, (size_t)B * 4, cudaMemcpyHostToDevice, stream));
  B200_CUDA(cudaMemcpyAsync(b_slot, slot, (size_t)B * 4, cudaMemcpyHostToDevice, stream));
  int rc = 0;
  // capturing and instantiating the ~330 nodes costs ~25 ms, five steps' worth: only for a batch size that has held for
  // kBatchGraphAfter steps in a row (a burst ramps through many sizes; a graph then saves ~3.5 % of each later step)
  batch_same = (B == batch_last) ? batch_same + 1 : 0;
  batch_last = B;
  const bool graph_now = use_batch_graph && tp_size == 1 && (batch_same >= kBatchGraphAfter || batch_graphs.count(B));
  if (graph_now) {
    auto it = batch_graphs.find(B);
    if (it == batch_graphs.end()) {
      cudaGraph_t g = nullptr;
      cudaGraphExec_t ex = nullptr;
      cudaError_t e = cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal);
      if (e != cudaSuccess) return cuda_fail(e, "cudaStreamBeginCapture (batched step)");
      count_launches_pause(true);
      rc = enqueue_batched_body(B);
      count_launches_pause(false);
      e = cudaStreamEndCapture(stream, &g);
      size_t n_nodes = 0;
      if (rc == 0 && e == cudaSuccess && cudaGraphGetNodes(g, nullptr, &n_nodes) == cudaSuccess) e = cudaGraphInstantiate(&ex, g, 0);
      if (g) cudaGraphDestroy(g);
      if (rc) return rc;
      if (e != cudaSuccess || !ex) {  // no graph for this engine then: plain launches from here on
        (void)cudaGetLastError();
        use_batch_graph = false;
      } else {
        if (batch_graphs.size() >= 64) {  // batch sizes come and go: bounded
          for (auto& old : batch_graphs) cudaGraphExecDestroy(old.second.first);
          batch_graphs.clear();
        }
        it = batch_graphs.emplace(B, std::make_pair(ex, (int)n_nodes)).first;
      }
    }
    if (use_batch_graph) {
      B200_CUDA(cudaGraphLaunch(it->second.first, stream));
      count_launch((unsigned long long)it->second.second);
    }
  }
  if (!(graph_now && use_batch_graph) && (rc = enqueue_batched_body(B))) return rc;
  for (int i = 0; i < B; ++i) {  // sampled requests: the sampler kernel on that row, writing over the argmax
    const Seq& q = seqs[sl[i]];
    if (q.greedy) continue;
    SampleArgs sa = {};
    sa.logits = b_logits + (size_t)i * c.vocab; sa.vocab = c.vocab; sa.params = sp_arr + sl[i];
    sa.history = hist_dev + (size_t)sl[i] * c.max_ctx; sa.st = nullptr; sa.out = b_next + i; sa.n_hist = q.pos + 1;
    sa.step = (unsigned long long)q.delivered;
    sa.part = sample_part;  // shared scratch: the rows are sampled one after the other on this stream
    B200_CUDA(launch_sample(sa, stream, false));
  }
  B200_CUDA(launch_append_history(b_next, b_slot, b_pos, hist_dev, c.max_ctx, B, stream));
  B200_CUDA(cudaMemcpyAsync(b_next_host, b_next, (size_t)B * 4, cudaMemcpyDeviceToHost, stream));
  B200_CUDA(cudaStreamSynchronize(stream));
  next->assign(b_next_host, b_next_host + B);
  return 0;
}

int Engine::bench_batched(int B, int ctx, int n_steps, float* ms_per_step, int32_t* last_ids) {
  int rc = ensure_batched();
  if (rc) return rc;
  const b200_config& c = lc;  // per-rank shapes
  if (!b_keys && (rc = alloc((void**)&b_keys, (size_t)cfg.max_seqs * 8))) return rc;
  std::vector<int32_t> tok(B, 1), pos(B), cx(B), slot(B);
  for (int i = 0; i < B; ++i) slot[i] = i;
  B200_CUDA(cudaMemcpyAsync(b_tok, tok.data(), (size_t)B * 4, cudaMemcpyHostToDevice, stream));
  B200_CUDA(cudaMemcpyAsync(b_slot, slot.data(), (size_t)B * 4, cudaMemcpyHostToDevice, stream));
  cudaEvent_t e0, e1;
  B200_CUDA(cudaEventCreate(&e0));
  B200_CUDA(cudaEventCreate(&e1));
  const int warm = 3;
  for (int step = 0; step < warm + n_steps && !rc; ++step) {
    if (step == warm) B200_CUDA(cudaEventRecord(e0, stream));
    for (int i = 0; i < B; ++i) { pos[i] = ctx - 1 + step; cx[i] = ctx + step; }
    B200_CUDA(cudaMemcpyAsync(b_pos, pos.data(), (size_t)B * 4, cudaMemcpyHostToDevice, stream));
    B200_CUDA(cudaMemcpyAsync(b_ctx, cx.data(), (size_t)B * 4, cudaMemcpyHostToDevice, stream));
    if ((rc = forward_gemm(B, b_tok, b_pos, b_ctx, b_slot))) break;
    if ((rc = add_rmsnorm_proj(output_norm, B, true))) return rc;
    B200_CUDA(launch_gemm_bf16(output16, c.vocab, c.d_model, pf_xb, B, b_logits, c.vocab, stream, &gemm_ws));
    if (tp_size > 1) {  // vocab-parallel lm_head: per-row (logit, id) keys, max over the ranks
      B200_CUDA(launch_argmax_rows_keys(b_logits, B, c.vocab, tp_rank * c.vocab, b_keys, stream));
      if ((rc = tp_allreduce_max_u64_n(b_keys, (size_t)B, stream))) break;
      B200_CUDA(launch_keys_to_ids(b_keys, b_next, B, stream));
    } else {
      B200_CUDA(launch_argmax_rows(b_logits, B, c.vocab, b_next, stream));
    }
    B200_CUDA(cudaMemcpyAsync(b_tok, b_next, (size_t)B * 4, cudaMemcpyDeviceToDevice, stream));  // feed the next step
    B200_CUDA(cudaStreamSynchronize(stream));  // pos / cx are rewritten next iteration (and the scheduler syncs per step too)
  }
  cudaEventRecord(e1, stream);
  cudaError_t ce = cudaStreamSynchronize(stream);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (rc) return rc;
  if (ce != cudaSuccess) return cuda_fail(ce, "bench_batched sync");
  if (last_ids) B200_CUDA(cudaMemcpy(last_ids, b_next, (size_t)B * 4, cudaMemcpyDeviceToHost));
  if (ms_per_step) *ms_per_step = ms / n_steps;
  return 0;
}

void Engine::deliver(int slot, const int32_t* toks, int n) {
  Seq& q = seqs[slot];
  Request* r = q.r;
  if (q.delivered == 0 && n > 0) q.t_first = std::chrono::steady_clock::now();
  std::lock_guard<std::mutex> lk(r->mu);
  for (int i = 0; i < n; ++i) {
    r->tokens.push_back(toks[i]);
    q.gen.push_back(toks[i]);
    ++q.delivered;
    if (toks[i] == cfg.eos_token || q.delivered >= q.max_out) r->complete = true;
  }
  r->cv.notify_all();
}

void Engine::finish_request(int slot, int rc, int reason) {
  Seq& q = seqs[slot];
  Request* r = q.r;
  const auto t_end = std::chrono::steady_clock::now();
  // what this slot's cache holds now: the prompt and every generated token that was fed back (all but the last).
  // Before `finished` is raised: from then on the client may free the request
  {
    std::lock_guard<std::mutex> lk(kv_mu);
    slot_kv[slot].clear();
    if (use_prefix_cache && rc == 0 && q.delivered > 0) {
      slot_kv[slot] = r->prompt;
      slot_kv[slot].insert(slot_kv[slot].end(), q.gen.begin(), q.gen.end() - 1);
    }
  }
  {
    std::lock_guard<std::mutex> lk(r->mu);
    r->stats.load_ns = ns_between(q.t0, q.t_load);
    r->stats.prompt_eval_ns = ns_between(q.t_load, q.delivered ? q.t_first : t_end);
    r->stats.eval_ns = q.delivered ? ns_between(q.t_first, t_end) : 0;
    r->stats.total_ns = ns_between(q.t0, t_end);
    r->stats.prompt_eval_count = q.n_prompt - q.reused;  // tokens actually evaluated (as Ollama reports with a warm cache)
    r->stats.eval_count = q.delivered;
    r->stats.done_reason = rc ? 3 : reason;
    r->error = rc;
    r->finished = true;
    r->cv.notify_all();
  }
  q.r = nullptr;
  if (bs1_owner == slot) bs1_owner = -1;
}

// Prefill a new request into `slot` and deliver its first token.  Uses the single-sequence state
// (StepState / prompt_dev / active_table), which afterwards describes this sequence (bs1_owner = slot).
int Engine::admit(Request* r, int slot) {
  Seq& q = seqs[slot];
  q = Seq();
  q.r = r;
  q.t0 = r->t_submit;  // the wait for a free slot counts (it is part of what the peer waits for): it lands in load_ns
  q.n_prompt = (int)r->prompt.size();
  int max_out = r->sp.num_predict > 0 ? r->sp.num_predict : cfg.max_ctx;
  if (q.n_prompt + max_out - 1 > cfg.max_ctx) max_out = cfg.max_ctx - q.n_prompt + 1;
  q.max_out = max_out;
  q.greedy = r->sp.temperature <= 0.f;
  if (max_out < 1) { set_error("prompt fills the whole context"); return B200_EINVAL; }
  int rc;
  const bool gemm = r->sp.prefill_mode == B200_MODE_GEMM && q.n_prompt > 1;
  if (gemm && (rc = ensure_prefill())) return rc;
  if ((rc = ensure_graph())) return rc;
  if (cfg.max_seqs > 1 && (rc = ensure_batched())) return rc;
  if ((rc = set_active_slot(slot))) return rc;
  q.t_load = std::chrono::steady_clock::now();
  int reuse = (use_prefix_cache && gemm) ? common_prefix(slot, r->prompt) : 0;
  if (reuse < kMinReuse) reuse = 0;
  invalidate_prefix_cache(slot);  // being overwritten from position `reuse` on; re-recorded when the request finishes
  q.reused = reuse;
  if ((rc = run_prompt(r->prompt.data(), q.n_prompt, max_out, q.greedy ? 1 : 0, r->sp.prefill_mode, &r->sp, reuse))) return rc;
  bs1_owner = slot;
  if (batched_ready) {  // per-slot history and sampling options for the batched path
    SampleParams sp0 = {};
    sp0.temperature = r->sp.temperature; sp0.top_p = r->sp.top_p; sp0.repeat_penalty = r->sp.repeat_penalty;
    sp0.top_k = r->sp.top_k; sp0.repeat_last_n = r->sp.repeat_last_n; sp0.seed = r->sp.seed;
    B200_CUDA(cudaMemcpyAsync(sp_arr + slot, &sp0, sizeof sp0, cudaMemcpyHostToDevice, stream));
  }
  B200_CUDA(cudaStreamSynchronize(stream));
  if (*host_n_out < 1) { set_error("prefill produced no token"); return B200_ECUDA; }
  const int32_t first = out_tokens[0];
  if (batched_ready)  // prompt + first token -> this slot's history row (prompt_dev holds exactly that now)
    B200_CUDA(cudaMemcpyAsync(hist_dev + (size_t)slot * cfg.max_ctx, prompt_dev, (size_t)std::min(q.n_prompt + 1, cfg.max_ctx) * 4,
                              cudaMemcpyDeviceToDevice, stream));
  q.token = first;
  q.pos = q.n_prompt;
  deliver(slot, &first, 1);
  return 0;
}

// Batched admission: the prompts of several waiting requests go through the layer stack together (one set of GEMMs
// over all their rows, causal attention per prompt), then each gets its first token from the batch-1 tail (mode-0
// lm_head, sampler) exactly as a lone admission would.  Failures end the affected requests, never the worker.
void Engine::admit_batch(const std::vector<Request*>& rs, const std::vector<int>& sl) {
  const int B = (int)rs.size();
  std::vector<PrefillSeg> segs;
  std::vector<int32_t> tok, pos, slot;
  int rc = ensure_prefill();
  if (!rc) rc = ensure_graph();
  if (!rc) rc = ensure_batched();
  const auto t_load = std::chrono::steady_clock::now();
  for (int i = 0; i < B; ++i) {  // every request owns its slot from here on, so a failure below can end it properly
    Request* r = rs[i];
    Seq& q = seqs[sl[i]];
    q = Seq();
    q.r = r;
    q.t0 = r->t_submit;
    q.t_load = t_load;
    q.n_prompt = (int)r->prompt.size();
    int max_out = r->sp.num_predict > 0 ? r->sp.num_predict : cfg.max_ctx;
    if (q.n_prompt + max_out - 1 > cfg.max_ctx) max_out = cfg.max_ctx - q.n_prompt + 1;
    q.max_out = max_out;
    q.greedy = r->sp.temperature <= 0.f;
    int reuse = use_prefix_cache ? common_prefix(sl[i], r->prompt) : 0;
    if (reuse < kMinReuse) reuse = 0;
    invalidate_prefix_cache(sl[i]);
    q.reused = reuse;
    PrefillSeg g = {(int)tok.size(), q.n_prompt - reuse, reuse, sl[i]};
    segs.push_back(g);
    for (int t = reuse; t < q.n_prompt; ++t) { tok.push_back(r->prompt[t]); pos.push_back(t); slot.push_back(sl[i]); }
  }
  const int N = (int)tok.size();
  auto cu = [&](cudaError_t e, const char* what) { if (!rc && e != cudaSuccess) rc = cuda_fail(e, what); };
  if (!rc) {
    cu(cudaMemcpyAsync(pfb_tok, tok.data(), (size_t)N * 4, cudaMemcpyHostToDevice, stream), "batched prefill: tokens");
    cu(cudaMemcpyAsync(pfb_pos, pos.data(), (size_t)N * 4, cudaMemcpyHostToDevice, stream), "batched prefill: positions");
    cu(cudaMemcpyAsync(pfb_slot, slot.data(), (size_t)N * 4, cudaMemcpyHostToDevice, stream), "batched prefill: slots");
  }
  const bool tr_ = getenv("B200_SCHED_TRACE") != nullptr;  // debug: host-side timeline of a batched admission on stderr (adds a stream sync)
  auto now_ = [] { return std::chrono::steady_clock::now(); };
  auto us_ = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
  const auto t_a = now_();
  if (!rc) rc = forward_gemm(N, pfb_tok, pfb_pos, nullptr, pfb_slot, 0, &segs);
  const auto t_b = now_();
  if (tr_) cudaStreamSynchronize(stream);
  const auto t_c = now_();
  // the residual stream of every row now sits in pf_x (+ the last down projection still to be added): hand each
  // prompt's last row to the batch-1 tail
  const float* proj_src = pf_proj_src;
  const int proj_splits = pf_proj_splits;
  const int64_t proj_stride = pf_proj_stride;
  for (int i = 0; i < B; ++i) {
    Request* r = rs[i];
    const int s = sl[i];
    Seq& q = seqs[s];
    int rq = rc;
    if (!rq) rq = set_active_slot(s);
    if (!rq) rq = begin_sequence(r->prompt.data(), q.n_prompt, q.max_out, q.greedy ? 1 : 0, &r->sp);
    if (!rq) {
      StepState s0 = {};
      s0.pos = q.n_prompt - 1; s0.cur_token = r->prompt[q.n_prompt - 1]; s0.n_prompt = q.n_prompt; s0.max_out = q.max_out; s0.eos = cfg.eos_token;
      s0.greedy = q.greedy ? 1 : 0;
      cudaError_t e = cudaMemcpyAsync(st, &s0, sizeof s0, cudaMemcpyHostToDevice, stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
      if (e == cudaSuccess) e = launch_take_row(pf_x, proj_src, proj_splits, proj_stride, cfg.d_model, segs[i].row0 + segs[i].n - 1, x, stream);
      if (e != cudaSuccess) rq = cuda_fail(e, "batched prefill: hand-over");
    }
    int n_tail = 0;
    if (!rq) rq = enqueue_tail(stream, false, &n_tail);
    if (!rq) {
      SampleParams sp0 = {};
      sp0.temperature = r->sp.temperature; sp0.top_p = r->sp.top_p; sp0.repeat_penalty = r->sp.repeat_penalty;
      sp0.top_k = r->sp.top_k; sp0.repeat_last_n = r->sp.repeat_last_n; sp0.seed = r->sp.seed;
      cudaError_t e = cudaMemcpyAsync(sp_arr + s, &sp0, sizeof sp0, cudaMemcpyHostToDevice, stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
      if (e != cudaSuccess) rq = cuda_fail(e, "batched prefill: first token");
      else if (*host_n_out < 1) { set_error("prefill produced no token"); rq = B200_ECUDA; }
    }
    if (rq) { finish_request(s, rq, 3); continue; }
    const int32_t first = out_tokens[0];
    cudaMemcpyAsync(hist_dev + (size_t)s * cfg.max_ctx, prompt_dev, (size_t)std::min(q.n_prompt + 1, cfg.max_ctx) * 4, cudaMemcpyDeviceToDevice, stream);
    bs1_owner = s;
    q.token = first;
    q.pos = q.n_prompt;
    deliver(s, &first, 1);
    if (q.token == cfg.eos_token) finish_request(s, 0, 1);
    else if (q.delivered >= q.max_out) finish_request(s, 0, 2);
  }
  if (tr_) fprintf(stderr, "SCHED admit_batch B=%d N=%d prep %.0f us, enqueue %.0f us, gpu-wait %.0f us, tails %.0f us\n", B, N, us_(t_load, t_a), us_(t_a, t_b), us_(t_b, t_c), us_(t_c, now_()));
}

// The only active sequence decodes on the batch-1 graph.  Returns when it finishes, is cancelled, or another
// request shows up (the caller then carries on with batched steps).
int Engine::decode_bs1_window(int slot) {
  Seq& q = seqs[slot];
  int rc = 0;
  if (bs1_owner != slot) {  // hand the sequence over to the single-sequence state
    if ((rc = set_active_slot(slot))) return rc;
    StepState s0 = {};
    s0.pos = q.pos; s0.cur_token = q.token; s0.n_prompt = std::min(q.n_prompt, q.pos); s0.n_out = q.delivered; s0.max_out = q.max_out;
    s0.eos = cfg.eos_token; s0.greedy = q.greedy ? 1 : 0;
    B200_CUDA(cudaMemcpyAsync(st, &s0, sizeof s0, cudaMemcpyHostToDevice, stream));
    B200_CUDA(cudaMemcpyAsync(prompt_dev, hist_dev + (size_t)slot * cfg.max_ctx, (size_t)cfg.max_ctx * 4, cudaMemcpyDeviceToDevice, stream));
    SampleParams sp0 = {};
    sp0.temperature = q.r->sp.temperature; sp0.top_p = q.r->sp.top_p; sp0.repeat_penalty = q.r->sp.repeat_penalty;
    sp0.top_k = q.r->sp.top_k; sp0.repeat_last_n = q.r->sp.repeat_last_n; sp0.seed = q.r->sp.seed;
    B200_CUDA(cudaMemcpyAsync(sp_dev, &sp0, sizeof sp0, cudaMemcpyHostToDevice, stream));
    B200_CUDA(cudaStreamSynchronize(stream));
    *host_n_out = q.delivered;
    B200_CUDA(launch_embed(token_embd, embd_type, cfg.d_model, &st->cur_token, x, stream));
    bs1_owner = slot;
  }
  int launched = q.delivered;  // token-producing steps issued so far (== tokens that will exist)
  const int window = 4;
  bool interrupted = false;
  for (;;) {
    bool cancelled;
    {
      std::lock_guard<std::mutex> lk(q.r->mu);
      cancelled = q.r->cancelled;
    }
    const int avail = *host_n_out;
    if (avail > q.delivered) {
      const int n = avail - q.delivered, from = q.delivered;
      deliver(slot, out_tokens + from, n);
      q.token = out_tokens[avail - 1];
      q.pos = q.n_prompt + avail - 1;
    }
    int reason = 0;
    if (cancelled) reason = 3;
    else if (q.delivered > 0 && q.token == cfg.eos_token) reason = 1;
    else if (q.delivered >= q.max_out) reason = 2;
    if (reason) {
      cudaStreamSynchronize(stream);
      finish_request(slot, 0, reason);
      return 0;
    }
    if (!interrupted) {
      // a waiting request interrupts the window only if the batched path can take both (can_batch); otherwise
      // this sequence runs to its end on the batch-1 graph and the queue is served afterwards
      std::lock_guard<std::mutex> lk(qmu);
      interrupted = (can_batch() && !queue.empty()) || stopping;
    }
    if (interrupted) {  // drain what is in flight, then let the scheduler batch
      if (launched == q.delivered) {
        // history of this slot for the batched path = what the finish kernel kept in prompt_dev
        if (batched_ready)
          B200_CUDA(cudaMemcpyAsync(hist_dev + (size_t)slot * cfg.max_ctx, prompt_dev, (size_t)cfg.max_ctx * 4, cudaMemcpyDeviceToDevice, stream));
        B200_CUDA(cudaStreamSynchronize(stream));
        return 0;
      }
      std::this_thread::sleep_for(std::chrono::microseconds(20));
      continue;
    }
    if (launched < q.max_out && launched - q.delivered < window) {
      if ((rc = launch_steps(1))) return rc;
      ++launched;
    } else {
      std::this_thread::sleep_for(std::chrono::microseconds(20));
    }
  }
}

void Engine::worker_loop() {
  cudaSetDevice(cfg.device);
  force_batched = (cfg.reserved & 1) != 0;
  for (;;) {
    // ---- wait for work
    std::vector<int> active;
    for (int s = 0; s < (int)seqs.size(); ++s)
      if (seqs[s].r) active.push_back(s);
    {
      std::unique_lock<std::mutex> lk(qmu);
      if (active.empty()) qcv.wait(lk, [&] { return stopping || !queue.empty(); });
      if (stopping) {
        for (Request* p : queue) {
          std::lock_guard<std::mutex> l2(p->mu);
          p->finished = true; p->error = B200_ESTATE; p->cv.notify_all();
        }
        queue.clear();
        lk.unlock();
        for (int s : active) finish_request(s, B200_ESTATE, 3);
        return;
      }
    }
    std::lock_guard<std::mutex> dev(mu);
    // ---- admit waiting requests into free slots; while sequences are decoding, at most a few prefills per
    // iteration so that running requests keep producing tokens (bounded inter-token stall)
    int budget = active.empty() ? (int)seqs.size() : 4;
    if (!active.empty()) {
      // a long queue (>= 16 waiting, B200_ADMIT_DEEP): full prefill batches of 8 prompts, so half as many decode steps
      // are interleaved while the batch is still small and a step still costs nearly what it costs at full batch
      // (256 peers: 9,050 -> 9,770 tok/s; the running sequences' inter-token stall goes from ~32 to ~62 ms meanwhile)
      std::lock_guard<std::mutex> lk(qmu);
      static const int deep = getenv("B200_ADMIT_DEEP") ? atoi(getenv("B200_ADMIT_DEEP")) : 16;
      if ((int)queue.size() >= deep) budget = 8;
    }
    // tensor-parallel engines are driven SPMD (every rank makes the same calls): admission timing would differ
    // between ranks, so they serve one request at a time; several slots exist only for b200_bench_batched
    if (tp_size > 1) budget = active.empty() ? 1 : 0;
    std::vector<Request*> batch_r;   // tensor-core prefills gathered this iteration: they share one pass (admit_batch)
    std::vector<int> batch_s;
    std::vector<char> taken(seqs.size(), 0);
    int batch_rows = 0;
    const int rows_cap = prefill_rows_cap();
    auto flush_batch = [&] {
      if (batch_r.size() == 1) {
        const int s = batch_s[0];
        const int rc = admit(batch_r[0], s);
        if (rc) { if (!seqs[s].r) seqs[s].r = batch_r[0]; finish_request(s, rc, 3); }
        else {
          Seq& q = seqs[s];
          if (q.token == cfg.eos_token) finish_request(s, 0, 1);
          else if (q.delivered >= q.max_out) finish_request(s, 0, 2);
        }
      } else if (batch_r.size() > 1) {
        admit_batch(batch_r, batch_s);
      }
      batch_r.clear(); batch_s.clear(); batch_rows = 0;
    };
    while (budget > 0) {
      int n_free = 0;
      for (size_t i = 0; i < seqs.size(); ++i) n_free += (seqs[i].r || taken[i]) ? 0 : 1;
      if (n_free == 0) break;
      --budget;
      Request* r = nullptr;
      {
        std::lock_guard<std::mutex> lk(qmu);
        if (queue.empty()) break;
        r = queue.front();
        queue.pop_front();
      }
      // the free slot whose cache shares the longest prefix with this prompt; without a usable prefix, a slot that
      // holds nothing (so other peers' histories stay cached), else the first free one
      int s = -1, best = -1;
      for (int i = 0; i < (int)seqs.size(); ++i) {
        if (seqs[i].r || taken[i]) continue;
        int score = use_prefix_cache ? common_prefix(i, r->prompt) : 0;
        if (score < kMinReuse) {
          std::lock_guard<std::mutex> lk(kv_mu);
          score = slot_kv[i].empty() ? 1 : 0;
        }
        if (score > best) { best = score; s = i; }
      }
      const int n_prompt = (int)r->prompt.size();
      int max_out = r->sp.num_predict > 0 ? r->sp.num_predict : cfg.max_ctx;
      if (n_prompt + max_out - 1 > cfg.max_ctx) max_out = cfg.max_ctx - n_prompt + 1;
      const bool batchable = rows_cap > 0 && r->sp.prefill_mode == B200_MODE_GEMM && n_prompt > 1 && max_out >= 1 && n_prompt <= rows_cap;
      if (!batchable) {  // token-by-token prefill, or a request admit() will refuse: on its own
        flush_batch();
        batch_r.push_back(r); batch_s.push_back(s);
        flush_batch();
        continue;
      }
      if (batch_rows + n_prompt > rows_cap) flush_batch();
      batch_r.push_back(r); batch_s.push_back(s);
      taken[s] = 1;
      batch_rows += n_prompt;
    }
    flush_batch();
    active.clear();
    for (int s = 0; s < (int)seqs.size(); ++s)
      if (seqs[s].r) active.push_back(s);
    if (active.empty()) continue;
    // ---- decode
    bool waiting;
    {
      std::lock_guard<std::mutex> lk(qmu);
      waiting = !queue.empty();
    }
    // The batched step needs its own state (ensure_batched: per-slot history and sampling options) and is not
    // tensor-parallel aware: engines with one slot, and TP engines (SPMD, one request at a time), stay on the
    // batch-1 graph whatever is waiting.
    if (active.size() == 1 && (!can_batch() || (!waiting && !force_batched))) {
      const int rc = decode_bs1_window(active[0]);
      if (rc && seqs[active[0]].r) finish_request(active[0], rc, 3);
      continue;
    }
    // cancelled sequences leave before the step
    for (size_t i = 0; i < active.size();) {
      bool c;
      {
        std::lock_guard<std::mutex> lk(seqs[active[i]].r->mu);
        c = seqs[active[i]].r->cancelled;
      }
      if (c) { finish_request(active[i], 0, 3); active.erase(active.begin() + i); }
      else ++i;
    }
    if (active.empty()) continue;
    std::vector<int32_t> next;
    int rc = ensure_batched();
    if (!rc) {
      if (bs1_owner >= 0) bs1_owner = -1;  // the single-sequence state goes stale once a batched step runs
      rc = batched_step(active, &next);
    }
    for (size_t i = 0; i < active.size(); ++i) {
      const int s = active[i];
      if (rc) { finish_request(s, rc, 3); continue; }
      Seq& q = seqs[s];
      q.pos += 1;
      q.token = next[i];
      deliver(s, &next[i], 1);
      if (q.token == cfg.eos_token) finish_request(s, 0, 1);
      else if (q.delivered >= q.max_out) finish_request(s, 0, 2);
    }
  }
}

}  // namespace b200
