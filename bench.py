#!/usr/bin/env python
"""bench.py — headline benchmark of the "Suggest a Reply" inference path (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA engine through the C ABI
    python bench.py --impl reference --steps K --warmup W    # the CPU restatement (oracle) on host cores

A *step* is one suggestion request on the configuration BASELINE.json quotes the metric on (configs[1]):
Llama-3.1-8B architecture, random-init bf16 weights, batch 1, 512-token synthetic chat-history prompt,
128 greedy tokens.  `value` = decode tokens/s summed over all ranks (device-timed with CUDA events inside
the C ABI, inputs resident in HBM); `ttft_ms` = p50 time to first token; `e2e` = the same metric measured
by a caller of b200_submit / b200_next with HOST buffers (prompt ids in, token ids out).  With N > 1 (launched
by torchrun) every rank serves its own request stream on its own GPU — independent replicas, no collective
on the data path — and the timing is the max over ranks.

Extra objects on the same JSON line: `roofline` (dominant kernel = the fused gate/up GEMV, timed alone with CUDA
events: 20 launches back to back — its 235 MB matrix is larger than the 126 MB L2 and streamed evict-first, so no
launch can hit what the previous one read; against MEASURED_PEAKS.json; `traffic` is read from the ncu capture
under profiles/), `step_roofline` (bytes streamed per token / step time), `cpu_baseline` (the oracle timed on this
box's host cores: the real 512-token context, a bounded number of decode tokens), `clocks` (nvidia-smi samples
during the timed region), `q4_k_m` (BASELINE configs[2] measured the same way), `peers256` (configs[3] through the
C ABI) and `http_peers256` (configs[3] through ONE llmd URL owning every GPU of the run), `http` (configs[1] over
HTTP), and with N >= 2 `tp_parity` / `tp` (configs[4]: tensor parallel over the N GPUs; 70B at N = 8).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# torchrun exports OMP_NUM_THREADS=1; the CPU legs (oracle) are meant to use every host thread
os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)

N_PROMPT, N_NEW = 512, 128
METRIC, UNIT = "decode_tok_per_s_llama3.1_8b_bs1", "tok/s"


def synthetic_prompt(vocab, n, seed=1):
    """512 ids framed as alternating chat turns: deterministic, < 128000 so no special tokens."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, min(vocab, 128000), n).astype(np.int32)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi samples during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------------------
def cpu_weights_8b(orc, cfg, max_ctx=64):
    """Random bf16 tensors of the 8B shapes built on the host (values only matter for timing): one random
    pool rolled per tensor; exponents confined to 2^-7..2^-4 so activations stay finite."""
    from tests.util import make_cfg
    oc = make_cfg(cfg.n_layers, cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.d_ffn, cfg.vocab, max_ctx,
                  cfg.rms_eps, cfg.rope_theta, cfg.rope_factor, cfg.rope_low_freq_factor, cfg.rope_high_freq_factor,
                  cfg.rope_orig_ctx, 1)
    m = orc.Model(oc)
    rng = np.random.default_rng(0)
    pool = rng.integers(0, 65536, 1 << 26, dtype=np.uint16)
    pool = (pool & np.uint16(0x807F)) | ((np.uint16(120) + (pool >> np.uint16(8) & np.uint16(3))) << np.uint16(7))
    d, qd, kvd, ff = cfg.d_model, cfg.n_heads * cfg.head_dim, cfg.n_kv_heads * cfg.head_dim, cfg.d_ffn

    def mat(rows, cols, k):
        n = rows * cols
        reps = -(-n // pool.size)
        a = np.empty(n, np.uint16)
        for i in range(reps):
            seg = a[i * pool.size:(i + 1) * pool.size]
            seg[:] = np.roll(pool, 977 * (k + i))[:seg.size] if i == 0 else pool[:seg.size]
        return a

    k = 0
    for name, rows, cols in [("token_embd.weight", cfg.vocab, d), ("output.weight", cfg.vocab, d)]:
        m.set_tensor(name, orc.BF16, mat(rows, cols, k), rows, cols); k += 1
    m.set_tensor("output_norm.weight", orc.F32, np.ones(d, np.float32), 1, d)
    for l in range(cfg.n_layers):
        for s, rows, cols in (("attn_q", qd, d), ("attn_k", kvd, d), ("attn_v", kvd, d), ("attn_output", d, qd),
                              ("ffn_gate", ff, d), ("ffn_up", ff, d), ("ffn_down", d, ff)):
            m.set_tensor(f"blk.{l}.{s}.weight", orc.BF16, mat(rows, cols, k), rows, cols); k += 1
        for s in ("attn_norm", "ffn_norm"):
            m.set_tensor(f"blk.{l}.{s}.weight", orc.F32, np.ones(d, np.float32), 1, d)
    return m


CPU_CTX_EXTRA = 16  # decode tokens the CPU legs may append to the 512-token context


class CpuRoute:
    """BASELINE configs[0] as SURVEY 8(d) states it: the CPU path "reported via the same /api/generate route".  The
    oracle is served by oracle/serve.py (test infrastructure; `llmd` has no CPU path); the conversation travels as
    token ids in Ollama's `context` field and the timings are read from Ollama's own reply fields."""

    def __init__(self, model, vocab):
        import requests
        from oracle import serve
        self.requests, self.srv = requests, serve.OracleServer(model, vocab)
        self.prompt = [int(t) for t in synthetic_prompt(vocab, N_PROMPT)]

    def post(self, num_predict):
        r = self.requests.post(self.srv.url + "/api/generate", timeout=3600, json={
            "model": "llama3.1", "prompt": "", "context": self.prompt, "stream": False,
            "options": {"temperature": 0, "num_predict": num_predict, "seed": 0}})
        if r.status_code != 200:
            raise RuntimeError(f"oracle route: HTTP {r.status_code} {r.text[:200]}")
        return r.json()

    def prefill(self):
        """The 512-token prompt (its prefill numerics, all host threads), once; returns seconds.  Later requests with the
        same ids find it cached, as with Ollama's prompt cache."""
        d = self.post(1)
        return d["prompt_eval_duration"] / 1e9

    def decode_sample(self, n_new):
        """n_new decode steps from the 512-token context: (tok/s, seconds) from eval_count / eval_duration (the first of
        the num_predict tokens comes out of the prompt evaluation, so n_new + 1 are asked for)."""
        d = self.post(n_new + 1)
        return (d["eval_count"] - 1) / (d["eval_duration"] / 1e9), d["eval_duration"] / 1e9

    def close(self):
        self.srv.close()


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The reference repo holds none
    (Ollama is external), so this is the in-tree CPU restatement on all host threads.
    Same configuration as the GPU arm — 8B bf16, the 512-token prompt in context — with each step a bounded
    sample of the 128 decode tokens (the full 512 + 128 request takes ~2 minutes on these cores)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import orc
    from types import SimpleNamespace
    # Llama-3.1-8B shapes, stated here so that this arm never loads libb200llm.so
    cfg = SimpleNamespace(n_layers=32, d_model=4096, n_heads=32, n_kv_heads=8, head_dim=128, d_ffn=14336, vocab=128256,
                          rms_eps=1e-5, rope_theta=500000.0, rope_factor=8.0, rope_low_freq_factor=1.0,
                          rope_high_freq_factor=4.0, rope_orig_ctx=8192)
    cores = os.cpu_count() or 1
    model = cpu_weights_8b(orc, cfg, max_ctx=N_PROMPT + CPU_CTX_EXTRA)
    route = CpuRoute(model, cfg.vocab)
    prefill_s = route.prefill()   # ~36 s, once, outside the timed steps
    n_n = 6
    for _ in range(min(args.warmup, 2)):
        route.decode_sample(2)
    vals, t_all = [], 0.0
    for _ in range(args.steps):
        v, t = route.decode_sample(n_n)
        vals.append(v); t_all += t
    route.close()
    value = statistics.mean(vals)
    sample = (f"8B bf16 random-init through POST /api/generate served by the CPU oracle (oracle/serve.py): the 512-token "
              f"prompt in context (CPU prefill {prefill_s:.0f} s, once), {n_n} of the 128 greedy decode tokens per step")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "ttft_ms": prefill_s * 1e3,
        "config": {"workload": "llama3.1-8b bf16 bs1 suggestion: 512-token synthetic prompt, 128 greedy tokens",
                   "weights": "random-init", "n_prompt": N_PROMPT, "n_new": N_NEW, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# --------------------------------------------------------------------------------------------------------
def e2e_request(eng, pkg, prompt, n_new):
    """What llmd does per request: host prompt ids -> b200_submit -> b200_next per token."""
    sp = pkg.default_sampling()
    sp.temperature, sp.num_predict, sp.prefill_mode = 0.0, n_new, pkg.MODE_GEMM
    t0 = time.perf_counter()
    req = eng.submit(prompt, sp)
    t_first, n = None, 0
    while True:
        _, done = req.next()
        n += 1
        if t_first is None:
            t_first = time.perf_counter()
        if done:
            break
    t1 = time.perf_counter()
    req.free()
    return (n - 1) / (t1 - t_first), (t_first - t0) * 1e3


def max_over_ranks(dist, values, device):
    """Every rank contributes its timings; the job is as slow as its slowest rank."""
    if dist is None:
        return list(values)
    import torch
    t = torch.tensor(values, device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def aggregate_tok_per_s(world, steps, n_new, max_decode_ms):
    """Whole-job decode throughput: every rank decoded steps*(n_new-1) tokens within the slowest rank's time."""
    return world * steps * (n_new - 1) / (max_decode_ms / 1e3)


def run_peers(eng, pkg, cfg, n_peers, n_prompt, n_new):
    """BASELINE configs[3]: n_peers UI sessions click 'Suggest a reply' at once (closed loop, one request each);
    continuous batching inside one engine.  Returns aggregate tok/s and server-side TTFT percentiles."""
    sp = pkg.default_sampling()
    sp.temperature, sp.num_predict, sp.prefill_mode = 0.0, n_new, pkg.MODE_GEMM
    prompts = [synthetic_prompt(cfg.vocab, n_prompt, seed=100 + i) for i in range(n_peers)]
    t0 = time.perf_counter()
    reqs = [eng.submit(p, sp) for p in prompts]
    n_tok = 0
    for r in reqs:
        n_tok += sum(1 for _ in r.tokens())
    wall = time.perf_counter() - t0
    st = [r.stats() for r in reqs]
    ttft = sorted((x.load_ns + x.prompt_eval_ns) / 1e6 for x in st)      # from b200_submit: includes the wait for a slot
    prefill = sorted(x.prompt_eval_ns / 1e6 for x in st)
    totals = sorted(x.total_ns / 1e6 for x in st)
    [r.free() for r in reqs]
    pct = lambda v, q: v[min(len(v) - 1, int(q * len(v)))]
    return {"peers": n_peers, "n_prompt": n_prompt, "n_new": n_new, "aggregate_tok_per_s": n_tok / wall, "wall_s": wall,
            "ttft_ms_p50": pct(ttft, 0.5), "ttft_ms_p99": pct(ttft, 0.99), "prefill_ms_p50": pct(prefill, 0.5),
            "request_ms_p50": pct(totals, 0.5),
            "note": "all 256 requests submitted at t=0 (closed loop, one each); TTFT from submission, queueing included"}


def run_http(n_runs=20):
    """SURVEY 8(d) config 2 as the reference would see it: llmd (the Ollama drop-in) serving the random-init 8B bf16
    model over HTTP; the web/streamlit_app.py:91-95 request with a 512-token prompt, streamed.  TTFT = request start ->
    first NDJSON chunk at the client; decode tok/s = eval_count / eval_duration of the final object (server side).
    Every run sends a different prompt, so the server's prefix cache never shortens the prefill being timed."""
    import requests
    proc, url = start_llmd([0], 4, 1024)
    try:
        ttft_c, ttft_s, tps, n_prompt = [], [], [], 0
        for i in range(n_runs + 2):
            # no vocabulary in a random-init engine: ids 0..255 are bytes, so 511 ASCII characters + BOS = 512 tokens
            rng = np.random.default_rng(7000 + i)
            text = "".join(chr(c) for c in rng.integers(97, 123, N_PROMPT - 1))
            body = {"model": "llama3.1", "prompt": text, "raw": True, "stream": True,
                    "options": {"temperature": 0, "num_predict": N_NEW, "seed": 0}}
            t0 = time.perf_counter()
            first, last = None, None
            with requests.post(url + "/api/generate", json=body, stream=True, timeout=120) as r:
                r.raise_for_status()
                for ln in r.iter_lines():
                    if not ln:
                        continue
                    if first is None:
                        first = time.perf_counter() - t0
                    last = ln
            fin = json.loads(last)
            if i < 2:
                continue  # warm-up: graph capture, prefill buffers
            n_prompt = fin["prompt_eval_count"]
            ttft_c.append(first * 1e3)
            ttft_s.append((fin["load_duration"] + fin["prompt_eval_duration"]) / 1e6)
            tps.append(fin["eval_count"] / (fin["eval_duration"] / 1e9))
        return {"runs": n_runs, "prompt_tokens": n_prompt, "ttft_ms_client_p50": statistics.median(ttft_c),
                "ttft_ms_server_p50": statistics.median(ttft_s), "decode_tok_per_s_p50": statistics.median(tps),
                "api": "POST /api/generate (stream) on llmd, 127.0.0.1, requests client"}
    finally:
        stop_llmd(proc)


def start_llmd(gpus, max_seqs, max_ctx):
    import socket
    import requests
    import __graft_entry__ as ge
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, LLMD_ADDR=f"127.0.0.1:{port}", LLMD_MAX_CTX=str(max_ctx), LLMD_SEED="0", LLMD_MODEL="llama3.1", LLMD_EOS="-1",
               LLMD_MAX_SEQS=str(max_seqs), LLMD_GPUS=",".join(str(g) for g in gpus), LLMD_MAX_CONNS="2048")
    proc = subprocess.Popen([os.path.join(ge.PKG_DIR, "llmd")], env=env, stderr=subprocess.DEVNULL)
    url = f"http://127.0.0.1:{port}"
    for _ in range(3000):
        if proc.poll() is not None:
            raise RuntimeError("llmd exited")
        try:
            if requests.get(url + "/", timeout=1).status_code == 200:
                return proc, url
        except requests.RequestException:
            time.sleep(0.1)
    proc.kill()
    raise RuntimeError("llmd did not come up")


def stop_llmd(proc):
    proc.terminate()
    try:
        proc.wait(timeout=30)
    except Exception:
        proc.kill()


def run_http_peers(n_gpus, n_peers=256):
    """BASELINE configs[3] as the reference would drive it: n_peers UI sessions post the web/streamlit_app.py:91-95
    request (512-token prompt, 128 new tokens) to ONE URL at the same moment; llmd owns all n_gpus GPUs behind it
    (one engine each, least-outstanding-tokens routing).  Streamed, so the client sees its own time to first token."""
    import requests
    per = -(-n_peers // n_gpus)
    proc, url = start_llmd(range(n_gpus), per, N_PROMPT + N_NEW)
    try:
        def body(i):
            # no vocabulary in a random-init engine: ids 0..255 are bytes; a distinct 512-byte history per peer
            rng = np.random.default_rng(1000 + i)
            text = "".join(chr(c) for c in rng.integers(97, 123, N_PROMPT - 1))
            return {"model": "llama3.1", "prompt": text, "raw": True, "stream": True,
                    "options": {"temperature": 0, "num_predict": N_NEW, "seed": 0}}
        out = [None] * n_peers

        def peer(i, bodies, t_start):
            first = None
            n = 0
            try:
                with requests.post(url + "/api/generate", json=bodies[i], stream=True, timeout=300) as r:
                    for ln in r.iter_lines():
                        if not ln:
                            continue
                        if first is None:
                            first = time.perf_counter() - t_start[0]
                        last = ln
                fin = json.loads(last)
                out[i] = (first, time.perf_counter() - t_start[0], fin.get("eval_count", 0), fin.get("prompt_eval_count", 0))
            except Exception as ex:
                out[i] = ex

        def burst(n, seed0):
            bodies = [body(seed0 + i) for i in range(n)]
            t_start = [time.perf_counter()]
            ts = [threading.Thread(target=peer, args=(i, bodies, t_start)) for i in range(n)]
            [t.start() for t in ts]
            [t.join() for t in ts]
            return time.perf_counter() - t_start[0]
        burst(min(8 * n_gpus, n_peers), 5000)          # warm-up: graph capture, prefill buffers on every engine
        wall = burst(n_peers, 0)
        bad = [o for o in out if not isinstance(o, tuple)]
        if bad:
            raise RuntimeError(f"{len(bad)} of {n_peers} requests failed: {bad[0]!r}")
        ttft = sorted(o[0] * 1e3 for o in out)
        n_tok = sum(o[2] for o in out)
        met = {}
        for ln in requests.get(url + "/metrics", timeout=10).text.splitlines():
            if ln.startswith("llmd_gpu_requests_total{"):
                met[ln.split("{")[1].split("}")[0]] = float(ln.rsplit(" ", 1)[1])
        pct = lambda v, q: v[min(len(v) - 1, int(q * len(v)))]
        return {"peers": n_peers, "gpus": n_gpus, "n_prompt": N_PROMPT, "n_new": N_NEW, "aggregate_tok_per_s": n_tok / wall, "wall_s": wall,
                "ttft_ms_client_p50": pct(ttft, 0.5), "ttft_ms_client_p99": pct(ttft, 0.99), "requests_per_gpu": met,
                "api": "256 x POST /api/generate (stream) to one llmd URL; llmd routes over its engines (LLMD_GPUS)"}
    finally:
        stop_llmd(proc)


def tp_parity(pkg, dist, rank, world, local):
    """Tensor parallelism checked in this run (the test box that runs tests/test_gpu_tp.py may have one GPU): a small
    Llama-architecture model tensor-parallel over all `world` ranks must reproduce the single-GPU engine and the CPU
    oracle — logits in both numerics modes, and every rank the same 24 greedy tokens."""
    from tests.util import tiny_cfg, oracle_from_engine
    mk = lambda: tiny_cfg(pkg, pkg.WEIGHTS_BF16, n_layers=2, seed=81, max_ctx=128, n_heads=16, n_kv_heads=8 if world > 4 else 4)
    prompt = np.random.default_rng(3).integers(0, 1024, 21).astype(np.int32)
    ref = None
    if rank == 0:
        c1 = mk()
        c1.device = local
        with pkg.Engine(c1) as e1:
            ora = oracle_from_engine(e1)
            l_gemv, l_gemm = e1.forward_logits(prompt, pkg.MODE_GEMV), e1.forward_logits(prompt, pkg.MODE_GEMM)
            ids1, _, _ = e1.generate_greedy(prompt, 24, pkg.MODE_GEMM)
        ora.reset()
        for t in prompt:
            want = ora.forward([t], mode=0)
        want_ids, margins = ora.generate_greedy(prompt, 24, prefill_mode=1)
        ref = (l_gemv, l_gemm, ids1, want, want_ids, margins)
    box = [pkg.tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    cfg = mk()
    cfg.device = local
    with pkg.Engine(cfg, tp_rank=rank, tp_size=world, nccl_id=box[0]) as eng:
        a = eng.forward_logits(prompt, pkg.MODE_GEMV)
        b = eng.forward_logits(prompt, pkg.MODE_GEMM)
        ids, _, _ = eng.generate_greedy(prompt, 24, pkg.MODE_GEMM)
    import torch
    all_ids = [None] * world
    dist.all_gather_object(all_ids, ids.tolist())
    if rank != 0:
        return None
    l_gemv, l_gemm, ids1, want, want_ids, margins = ref
    rel = lambda x, y: float(np.abs(x - y).max() / max(1.0, np.abs(y).max()))
    res = {"tp": world, "model": "2-layer d=2048 random-init", "logit_err_vs_tp1_gemv": rel(a, l_gemv), "logit_err_vs_tp1_gemm": rel(b, l_gemm),
           "logit_err_vs_oracle": rel(a, want), "ranks_agree_on_24_tokens": all(x == all_ids[0] for x in all_ids),
           "tokens_equal_tp1": all_ids[0] == ids1.tolist(), "tokens_equal_oracle": all_ids[0] == want_ids.tolist(),
           "oracle_min_margin": float(margins.min()),
           "tolerances": {"gemv_vs_tp1": 1e-4, "gemm_vs_tp1": 3e-3, "vs_oracle": 1e-3}}
    res["ok"] = bool(res["logit_err_vs_tp1_gemv"] <= 1e-4 and res["logit_err_vs_tp1_gemm"] <= 3e-3 and res["logit_err_vs_oracle"] <= 1e-3 and
                     res["ranks_agree_on_24_tokens"] and (res["tokens_equal_oracle"] or res["oracle_min_margin"] <= 1.2e-2))
    return res


def tp_section(pkg, dist, rank, world, local, model, steps):
    """BASELINE configs[4]: one model tensor-parallel over all ranks; returns the numbers of the `--workload tp` line."""
    import torch
    box = [pkg.tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    cfg = pkg.llama31_70b() if model == "70b" else pkg.llama31_8b()
    cfg.device, cfg.max_ctx, cfg.seed, cfg.eos_token = local, 1024, 0, -1
    cfg.max_seqs = 32  # slots for the bs32 line (b200_bench_batched); requests themselves are served one at a time
    eng = pkg.Engine(cfg, tp_rank=rank, tp_size=world, nccl_id=box[0])
    prompt = synthetic_prompt(cfg.vocab, N_PROMPT)
    for _ in range(3):
        eng.generate_greedy(prompt, 16)
    torch.cuda.synchronize(local); dist.barrier()
    ttfts, dec_ms = [], 0.0
    for _ in range(steps):
        _, ttft, dec = eng.generate_greedy(prompt, N_NEW)
        ttfts.append(ttft); dec_ms += dec
    torch.cuda.synchronize(local); dist.barrier()
    step_ms, launches = eng.bench_decode(N_PROMPT + N_NEW // 2, 32)
    bs = 32
    b_ms, _ = eng.bench_batched(bs, N_PROMPT + N_NEW // 2, 16)
    t = torch.zeros(bs * cfg.d_model, dtype=torch.float32, device=f"cuda:{local}")
    for _ in range(20):
        dist.all_reduce(t)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(local); dist.barrier()
    ev0.record()
    for _ in range(2 * cfg.n_layers):
        dist.all_reduce(t)
    ev1.record()
    torch.cuda.synchronize(local)
    ar_ms = ev0.elapsed_time(ev1)
    dec_ms, ttft_p50, step_ms, b_ms, ar_ms = max_over_ranks(dist, [dec_ms, statistics.median(ttfts), step_ms, b_ms, ar_ms], f"cuda:{local}")
    wb = eng.weight_bytes()
    eng.close()
    dist.barrier()
    if rank != 0:
        return None
    peaks, peak_src = measured_peaks()
    hbm = float(peaks["hbm_gbs"])
    kv = 2 * cfg.n_layers * (cfg.n_kv_heads // world) * cfg.head_dim * 2 * (N_PROMPT + N_NEW // 2)
    return {"model": f"llama3.1-{model}", "tp": world, "bs1_decode_tok_per_s": steps * (N_NEW - 1) / (dec_ms / 1e3), "ttft_ms": ttft_p50,
            "allreduce_bs1": "fused into the o/down GEMVs over NVLink peer memory" if not os.environ.get("B200_TP_NCCL") else "NCCL",
            "step_roofline": {"bound": "hbm", "bytes_per_token_per_gpu": wb + kv, "ms_per_token": step_ms,
                              "achieved": (wb + kv) / (step_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s per GPU",
                              "frac": (wb + kv) / (step_ms * 1e-3) / 1e9 / hbm, "launches_per_token": launches, "peak_source": peak_src},
            "bs32": {"decode_tok_per_s": bs / (b_ms * 1e-3), "ms_per_step": b_ms, "ctx": N_PROMPT + N_NEW // 2,
                     "allreduce_ms_per_step": ar_ms, "allreduce_share": ar_ms / b_ms, "allreduce": "NCCL, fp32 [32, d_model] x 2 per layer",
                     "note": "allreduce_* = the same 2*n_layers all-reduces timed alone (torch.distributed NCCL on the same fabric)"}}


def profile_traffic():
    """dram__bytes_read + dram__bytes_write of the dominant kernel (gate/up GEMV), one launch, from the newest
    `ncu --set full` capture committed under profiles/ (wide CSV of `ncu --page raw --csv`).  None if absent."""
    import csv
    import glob
    unit = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "prof_gemv_bf16_r*_raw.csv")), reverse=True):
        try:
            rows = list(csv.reader(open(path)))
            h, u = rows[0], rows[1]
            ir, iw, ik = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("Kernel Name")
            for r in rows[2:]:
                if "gemv_kernel" in r[ik] and ", 1, 2>" in r[ik]:   # <format, PRO_RMSNORM, EPI_SWIGLU> = gate/up
                    return int(float(r[ir]) * unit[u[ir]] + float(r[iw]) * unit[u[iw]]), os.path.relpath(path, ROOT)
        except (OSError, ValueError, KeyError, IndexError):
            continue
    return None, None


def run_tp(args):
    """--workload tp: BASELINE configs[4] on its own (the default workload reports the same object under `tp` when
    launched on N >= 2 GPUs).  Every rank drives its shard with the same calls (SPMD); timing is the max over ranks."""
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = ge.load_package()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = pkg.launch_count()
    res = tp_section(pkg, dist, rank, world, local, args.model, args.steps)
    clocks = sampler.stop()
    if rank == 0:
        print(json.dumps({
            "metric": f"decode_tok_per_s_llama3.1_{args.model}_tp{world}_bs1", "value": res["bs1_decode_tok_per_s"], "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": 3, "ms_per_step": None, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "ttft_ms": res["ttft_ms"],
            "config": {"workload": f"llama3.1-{args.model} bf16 bs1: 512-token synthetic prompt, 128 greedy tokens", "parallelism": f"tp{world}"},
            "step_roofline": res["step_roofline"], "bs32": res["bs32"], "gpu_launches": pkg.launch_count() - n0, "clocks": clocks}))
    dist.barrier()
    dist.destroy_process_group()


def profile_traffic():
    """dram__bytes_read + dram__bytes_write of the dominant kernel (gate/up GEMV), one launch, from the newest
    `ncu --set full` capture committed under profiles/ (wide CSV of `ncu --page raw --csv`).  None if absent."""
    import csv
    import glob
    unit = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "prof_gemv_bf16_r*_raw.csv")), reverse=True):
        try:
            rows = list(csv.reader(open(path)))
            h, u = rows[0], rows[1]
            ir, iw, ik = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("Kernel Name")
            for r in rows[2:]:
                if "gemv_kernel" in r[ik] and ", 1, 2>" in r[ik]:   # <format, PRO_RMSNORM, EPI_SWIGLU> = gate/up
                    return int(float(r[ir]) * unit[u[ir]] + float(r[iw]) * unit[u[iw]]), os.path.relpath(path, ROOT)
        except (OSError, ValueError, KeyError, IndexError):
            continue
    return None, None


def run_tp(args):
    """BASELINE configs[4]: one model tensor-parallel over all ranks (torchrun), NCCL all-reduce over NVLink.
    Every rank drives its shard with the same calls (SPMD); timing is the max over ranks."""
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = ge.load_package()
    box = [pkg.tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    cfg = pkg.llama31_70b() if args.model == "70b" else pkg.llama31_8b()
    cfg.device, cfg.max_ctx, cfg.seed, cfg.eos_token = local, 1024, 0, -1
    cfg.max_seqs = 32  # slots for the bs32 line (b200_bench_batched); requests themselves are served one at a time
    eng = pkg.Engine(cfg, tp_rank=rank, tp_size=world, nccl_id=box[0])
    prompt = synthetic_prompt(cfg.vocab, N_PROMPT)
    for _ in range(max(args.warmup, 3)):
        eng.generate_greedy(prompt, 16)
    torch.cuda.synchronize(local); dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ttfts, dec_ms, t0 = [], 0.0, time.perf_counter()
    for _ in range(args.steps):
        _, ttft, dec = eng.generate_greedy(prompt, N_NEW)
        ttfts.append(ttft); dec_ms += dec
    torch.cuda.synchronize(local); dist.barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    step_ms, launches = eng.bench_decode(N_PROMPT + N_NEW // 2, 32)
    # bs32: 32 sequences decoding together on the tensor-core path (NCCL all-reduce of [32, d_model] after o and down)
    bs = 32
    b_ms, b_ids = eng.bench_batched(bs, N_PROMPT + N_NEW // 2, 16)
    # the same all-reduces alone (2 per layer, fp32 [32, d_model]) on the same NVLink fabric, device-timed
    t = torch.zeros(bs * cfg.d_model, dtype=torch.float32, device=f"cuda:{local}")
    for _ in range(20):
        dist.all_reduce(t)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(local); dist.barrier()
    ev0.record()
    for _ in range(2 * cfg.n_layers):
        dist.all_reduce(t)
    ev1.record()
    torch.cuda.synchronize(local)
    ar_ms = ev0.elapsed_time(ev1)
    dec_ms, wall, ttft_p50, step_ms, b_ms, ar_ms = max_over_ranks(
        dist, [dec_ms, wall, statistics.median(ttfts), step_ms, b_ms, ar_ms], f"cuda:{local}")
    if rank == 0:
        peaks, peak_src = measured_peaks()
        hbm = float(peaks["hbm_gbs"])
        wb = eng.weight_bytes()  # per rank
        kv = 2 * cfg.n_layers * (cfg.n_kv_heads // world) * cfg.head_dim * 2 * (N_PROMPT + N_NEW // 2)
        print(json.dumps({
            "metric": f"decode_tok_per_s_llama3.1_{args.model}_tp{world}_bs1", "value": args.steps * (N_NEW - 1) / (dec_ms / 1e3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * wall / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "ttft_ms": ttft_p50,
            "config": {"workload": f"llama3.1-{args.model} bf16 bs1: 512-token synthetic prompt, 128 greedy tokens",
                       "parallelism": f"tp{world} (o/down all-reduce fused into the GEMVs over NVLink peer memory; NCCL for prefill and the argmax key)" if not os.environ.get("B200_TP_NCCL") else f"tp{world} (NCCL all-reduce after o and down)"},
            "step_roofline": {"bound": "hbm", "bytes_per_token_per_gpu": wb + kv, "ms_per_token": step_ms,
                              "achieved": (wb + kv) / (step_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s per GPU",
                              "frac": (wb + kv) / (step_ms * 1e-3) / 1e9 / hbm, "launches_per_token": launches, "peak_source": peak_src},
            "bs32": {"decode_tok_per_s": bs / (b_ms * 1e-3), "ms_per_step": b_ms, "ctx": N_PROMPT + N_NEW // 2,
                     "allreduce_ms_per_step": ar_ms, "allreduce_share": ar_ms / b_ms,
                     "allreduce": "push over NVLink peer memory (tp_push_kernel), reduced by the consumer" if os.environ.get("B200_TP_PUSH") else "NCCL",
                     "note": "b200_bench_batched: 32 sequences per step, tensor-core GEMMs, all-reduce of fp32 [32, d_model] x 2 per layer; "
                             "allreduce_* = the same 2*n_layers all-reduces timed alone (torch.distributed NCCL on the same fabric)"},
            "gpu_launches": args.steps * (N_NEW - 1) * launches, "clocks": clocks}))
    eng.close()
    dist.barrier()
    dist.destroy_process_group()


def run_ours(args):
    import torch
    import __graft_entry__ as ge
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist, host_group = None, None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")   # waits that must not park a kernel on an idle GPU
    fast = bool(os.environ.get("B200_BENCH_FAST"))
    pkg = ge.load_package()
    pkg.lib()
    cfg = pkg.llama31_8b()
    cfg.device, cfg.max_ctx, cfg.seed, cfg.eos_token = local, 1024, 0, -1
    cfg.reserved = 2          # no prefix reuse: every timed request pays its whole 512-token prefill
    eng = pkg.Engine(cfg)
    prompt = synthetic_prompt(cfg.vocab, N_PROMPT)

    def barrier():
        torch.cuda.synchronize(local)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(local)

    for _ in range(max(args.warmup, 3)):
        eng.generate_greedy(prompt, N_NEW)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    launches0 = pkg.launch_count()
    ttfts, dec_ms, t0 = [], 0.0, time.perf_counter()
    for _ in range(args.steps):
        _, ttft, dec = eng.generate_greedy(prompt, N_NEW)
        ttfts.append(ttft); dec_ms += dec
    barrier()
    wall = time.perf_counter() - t0
    n_launched = pkg.launch_count() - launches0   # counted inside the library: direct launches + kernel nodes of graph replays
    clocks = sampler.stop()
    e2e_vals, e2e_ttft = [], []
    for _ in range(args.steps):
        v, t = e2e_request(eng, pkg, prompt, N_NEW)
        e2e_vals.append(v); e2e_ttft.append(t)
    step_ms, launches = eng.bench_decode(N_PROMPT + N_NEW // 2, 32)
    k_ms, k_bytes = eng.bench_kernel("gate_up", 20)
    wbytes = eng.weight_bytes()
    eng.close()

    dec_ms, wall, ttft_p50, inv_e2e = max_over_ranks(dist, [dec_ms, wall, statistics.median(ttfts), 1.0 / statistics.mean(e2e_vals)],
                                                     f"cuda:{local}")
    value, e2e_val = aggregate_tok_per_s(world, args.steps, N_NEW, dec_ms), world / inv_e2e
    if dist is not None:
        tl = torch.tensor([float(n_launched)], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(tl)
        n_launched = int(tl.item())

    extra = {}
    # BASELINE configs[3] through the C ABI: the 256 sessions are dealt to the replicas (256/N each, one engine per GPU,
    # no collective); aggregate = all tokens / slowest replica's wall time.
    if not fast:
        try:
            per = 256 // world
            cfg.weight_format, cfg.max_seqs, cfg.max_ctx = pkg.WEIGHTS_BF16, per, N_PROMPT + N_NEW
            with pkg.Engine(cfg) as pe:
                run_peers(pe, pkg, cfg, min(8, per), N_PROMPT, 8)   # warm-up: graph capture, shadow buffers
                barrier()
                pr = run_peers(pe, pkg, cfg, per, N_PROMPT, N_NEW)
            cfg.max_seqs, cfg.max_ctx = 1, 1024
            wall_p, ttft50, ttft99 = max_over_ranks(dist, [pr["wall_s"], pr["ttft_ms_p50"], pr["ttft_ms_p99"]], f"cuda:{local}")
            pr.update({"peers": per * world, "replicas": world, "aggregate_tok_per_s": per * world * N_NEW / wall_p, "wall_s": wall_p,
                       "ttft_ms_p50": ttft50, "ttft_ms_p99": ttft99})
            extra["peers256"] = pr
        except Exception as ex:
            extra["peers256"] = {"error": str(ex)}
    # BASELINE configs[4]: tensor parallelism over the N GPUs of this run — parity first, then the numbers
    if world > 1 and not fast:
        try:
            extra["tp_parity"] = tp_parity(pkg, dist, rank, world, local)
        except Exception as ex:
            extra["tp_parity"] = {"error": str(ex)}
        try:
            tp = {"8b": tp_section(pkg, dist, rank, world, local, "8b", 3)}
            if world == 8:
                tp["70b"] = tp_section(pkg, dist, rank, world, local, "70b", 3)
            extra["tp"] = tp
        except Exception as ex:
            extra["tp"] = {"error": str(ex)}
    if rank == 0:
        peaks, peak_src = measured_peaks()
        hbm = float(peaks["hbm_gbs"])
        kv_bytes = 2 * cfg.n_layers * cfg.n_kv_heads * cfg.head_dim * 2 * (N_PROMPT + N_NEW // 2)
        ach = k_bytes / (k_ms * 1e-3) / 1e9
        traffic, traffic_src = profile_traffic()
        extra["roofline"] = {"kernel": "gemv_kernel<bf16, rmsnorm, swiglu> (ffn gate/up, 28672x4096)", "bound": "hbm",
                             "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
                             "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                             "timing": "CUDA events on the engine stream, 20 launches back to back; the 235 MB matrix exceeds the 126 MB L2",
                             "bytes_per_launch": k_bytes, "ms_per_launch": k_ms}
        sach = (wbytes + kv_bytes) / (step_ms * 1e-3) / 1e9
        extra["step_roofline"] = {"bound": "hbm", "bytes_per_token": wbytes + kv_bytes, "ms_per_token": step_ms,
                                  "achieved": sach, "peak": hbm, "unit": "GB/s", "frac": sach / hbm,
                                  "ceiling_tok_per_s": hbm * 1e9 / (wbytes + kv_bytes), "launches_per_token": launches}
        # BASELINE configs[2]: the same model in Ollama's default Q4_K_M mix (N=1 only, like the CPU baseline)
        if world == 1:
            try:
                cfg.weight_format = pkg.WEIGHTS_Q4_K_M
                q = pkg.Engine(cfg)
                for _ in range(2):
                    q.generate_greedy(prompt, N_NEW)
                _, qt, qd = q.generate_greedy(prompt, N_NEW)
                qs, _ = q.bench_decode(N_PROMPT + N_NEW // 2, 32)
                qb = q.weight_bytes() + kv_bytes
                extra["q4_k_m"] = {"decode_tok_per_s": (N_NEW - 1) / (qd / 1e3), "ttft_ms": qt, "ms_per_token": qs,
                                   "bytes_per_token": qb, "frac_of_hbm_peak": qb / (qs * 1e-3) / 1e9 / hbm}
                q.close()
            except Exception as ex:  # reported, never silently dropped
                extra["q4_k_m"] = {"error": str(ex)}
            cfg.weight_format = pkg.WEIGHTS_BF16
        # the same workloads through the HTTP boundary the reference actually calls: llmd is its own process and owns
        # every GPU of this run behind ONE url (the other ranks have closed their engines and wait on the host)
        if not fast:
            if world == 1:
                try:
                    extra["http"] = run_http()
                except Exception as ex:
                    extra["http"] = {"error": str(ex)}
            try:
                extra["http_peers256"] = run_http_peers(world)
            except Exception as ex:
                extra["http_peers256"] = {"error": str(ex)}
        # CPU baseline: the oracle on this box's host cores, the 512-token context, a bounded number of decode tokens
        if world == 1:
            try:
                from oracle import orc
                cores = os.cpu_count() or 1
                m = cpu_weights_8b(orc, cfg, max_ctx=N_PROMPT + CPU_CTX_EXTRA)
                if fast:  # development runs: two decode tokens at a 2-token context
                    m.reset()
                    m.forward(synthetic_prompt(cfg.vocab, 2), mode=1)
                    t1 = time.perf_counter()
                    m.forward([1], mode=0)
                    m.forward([1], mode=0)
                    extra["cpu_baseline"] = {"value": 2 / (time.perf_counter() - t1), "unit": UNIT, "cores": cores, "kind": "port",
                                             "sample": "8B bf16 random-init, 2 decode tokens at a 2-token context (B200_BENCH_FAST)"}
                else:
                    route = CpuRoute(m, cfg.vocab)
                    try:
                        prefill_s = route.prefill()
                        route.decode_sample(1)
                        v, _ = route.decode_sample(12)
                    finally:
                        route.close()
                    extra["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "ttft_ms": prefill_s * 1e3,
                                             "route": "POST /api/generate served by the CPU oracle (oracle/serve.py), temperature 0",
                                             "sample": f"8B bf16 random-init, the 512-token prompt in context (CPU prefill {prefill_s:.0f} s), "
                                                       "12 of the 128 greedy decode tokens"}
            except Exception as ex:
                extra["cpu_baseline"] = {"error": str(ex)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "ttft_ms": ttft_p50,
            "config": {"workload": "llama3.1-8b bf16 bs1 suggestion: 512-token synthetic prompt, 128 greedy tokens",
                       "weights": "random-init (device-side counter hash, seed 0)", "n_prompt": N_PROMPT, "n_new": N_NEW,
                       "parallelism": f"dp{world} (independent replicas, no collective)",
                       "l2": "inputs larger than L2: 15 GB of weights stream per token"},
            "e2e": {"value": e2e_val, "unit": UNIT, "ttft_ms": statistics.median(e2e_ttft), "h2d_bytes_per_step": N_PROMPT * 4,
                    "d2h_bytes_per_step": N_NEW * 4, "api": "b200_submit + b200_next (host buffers)"},
            "gpu_launches": n_launched,
            "gpu_launches_note": "b200_launch_count() across the timed region, summed over ranks: every launcher of the library counts "
                                 "itself, a CUDA-graph replay counts its kernel nodes",
            "clocks": clocks,
        }
        line.update(extra)
        print(json.dumps(line))
    if dist is not None:
        dist.barrier(group=host_group)   # the other ranks wait here, on the host, while rank 0 drives llmd over HTTP
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="suggest", choices=["suggest", "tp"],
                    help="suggest: BASELINE configs[1] on independent replicas (default); tp: one model tensor-parallel over all ranks")
    ap.add_argument("--model", default="70b", choices=["8b", "70b"], help="--workload tp only")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "tp":
        run_tp(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
