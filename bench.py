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

Extra objects on the same JSON line: `roofline` (dominant kernel = the fused gate/up GEMV, timed alone with
L2 flushed, against MEASURED_PEAKS.json), `step_roofline` (bytes streamed per token / step time),
`cpu_baseline` (the oracle timed on this box's host cores on a bounded sample), `clocks` (nvidia-smi samples
during the timed region), `q4_k_m` (BASELINE configs[2] measured the same way).
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


def cpu_decode_sample(model, vocab, n_prompt, n_new):
    """Token-by-token prompt then greedy decode on the host cores; returns decode tok/s and ttft."""
    prompt = synthetic_prompt(vocab, n_prompt)
    t0 = time.perf_counter()
    model.reset()
    for t in prompt:
        lg = model.forward([t], mode=0)
    t1 = time.perf_counter()
    tok = int(np.argmax(lg))
    for _ in range(n_new - 1):
        lg = model.forward([tok], mode=0)
        tok = int(np.argmax(lg))
    t2 = time.perf_counter()
    return (n_new - 1) / (t2 - t1), (t1 - t0) * 1e3, t2 - t0


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The reference repo holds none
    (Ollama is external), so this is the in-tree CPU restatement on all host threads."""
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
    model = cpu_weights_8b(orc, cfg)
    n_p, n_n = 2, 5   # bounded sample of the 512+128 workload: ~7 forward passes of the full 8B model per step
    for _ in range(max(args.warmup, 1) if args.warmup else 0):
        cpu_decode_sample(model, cfg.vocab, 1, 2)
    vals, t_all = [], 0.0
    for _ in range(args.steps):
        v, _, t = cpu_decode_sample(model, cfg.vocab, n_p, n_n)
        vals.append(v); t_all += t
    value = statistics.mean(vals)
    sample = f"8B bf16 random-init, {n_p}-token prompt + {n_n} greedy tokens per step (bounded sample of 512+128)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "llama3.1-8b bf16 bs1 suggestion: 512-token synthetic prompt, 128 greedy tokens",
                   "sample": sample},
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
    ttft = sorted((r.stats().load_ns + r.stats().prompt_eval_ns) / 1e6 for r in reqs)
    totals = sorted(r.stats().total_ns / 1e6 for r in reqs)
    [r.free() for r in reqs]
    return {"peers": n_peers, "n_prompt": n_prompt, "n_new": n_new, "aggregate_tok_per_s": n_tok / wall, "wall_s": wall,
            "ttft_ms_p50_in_engine": ttft[len(ttft) // 2], "ttft_ms_p99_in_engine": ttft[min(len(ttft) - 1, int(0.99 * len(ttft)))],
            "request_ms_p50": totals[len(totals) // 2], "note": "TTFT counted from admission into a slot; queueing before it is in request_ms"}


def run_http(n_runs=20):
    """SURVEY 8(d) config 2 as the reference would see it: llmd (the Ollama drop-in) serving the random-init 8B bf16
    model over HTTP; the web/streamlit_app.py:91-95 request with a 512-token prompt, streamed.  TTFT = request start ->
    first NDJSON chunk at the client; decode tok/s = eval_count / eval_duration of the final object (server side)."""
    import socket
    import subprocess
    import requests
    import __graft_entry__ as ge
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, LLMD_ADDR=f"127.0.0.1:{port}", LLMD_MAX_CTX="1024", LLMD_SEED="0", LLMD_MODEL="llama3.1", LLMD_EOS="-1",
               LLMD_MAX_SEQS="4")
    proc = subprocess.Popen([os.path.join(ge.PKG_DIR, "llmd")], env=env, stderr=subprocess.DEVNULL)
    url = f"http://127.0.0.1:{port}"
    try:
        for _ in range(1200):
            if proc.poll() is not None:
                raise RuntimeError("llmd exited")
            try:
                if requests.get(url + "/", timeout=1).status_code == 200:
                    break
            except requests.RequestException:
                time.sleep(0.1)
        else:
            raise RuntimeError("llmd did not come up")
        # no vocabulary in a random-init engine: ids 0..255 are bytes, so 511 ASCII characters + BOS = 512 tokens
        text = ("userA: are we still on for friday? userB: yes, 7pm at the usual place. " * 8)[:N_PROMPT - 1]
        body = {"model": "llama3.1", "prompt": text, "raw": True, "stream": True,
                "options": {"temperature": 0, "num_predict": N_NEW, "seed": 0}}
        ttft_c, ttft_s, tps, n_prompt = [], [], [], 0
        for i in range(n_runs + 2):
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
        proc.terminate()
        try:
            proc.wait(timeout=20)
        except Exception:
            proc.kill()


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
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = ge.load_package()
    pkg.lib()
    cfg = pkg.llama31_8b()
    cfg.device, cfg.max_ctx, cfg.seed, cfg.eos_token = local, 1024, 0, -1
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
    ttfts, dec_ms, t0 = [], 0.0, time.perf_counter()
    for _ in range(args.steps):
        _, ttft, dec = eng.generate_greedy(prompt, N_NEW)
        ttfts.append(ttft); dec_ms += dec
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    e2e_vals, e2e_ttft = [], []
    for _ in range(args.steps):
        v, t = e2e_request(eng, pkg, prompt, N_NEW)
        e2e_vals.append(v); e2e_ttft.append(t)
    step_ms, launches = eng.bench_decode(N_PROMPT + N_NEW // 2, 32)
    k_ms, k_bytes = eng.bench_kernel("gate_up", 20)
    wbytes = eng.weight_bytes()

    dec_ms, wall, ttft_p50, inv_e2e = max_over_ranks(dist, [dec_ms, wall, statistics.median(ttfts), 1.0 / statistics.mean(e2e_vals)],
                                                     f"cuda:{local}")
    value, e2e_val = aggregate_tok_per_s(world, args.steps, N_NEW, dec_ms), world / inv_e2e

    extra = {}
    # BASELINE configs[3] on several GPUs: the 256 sessions are dealt to the replicas (256/N each, one engine per GPU, no
    # collective); aggregate = all tokens / slowest replica's wall time.  (N = 1 runs it further down, after the Q4 line.)
    if world > 1 and not os.environ.get("B200_BENCH_FAST"):
        try:
            eng.close()
            eng = None
            per = 256 // world
            cfg.weight_format, cfg.max_seqs, cfg.max_ctx = pkg.WEIGHTS_BF16, per, N_PROMPT + N_NEW
            with pkg.Engine(cfg) as pe:
                run_peers(pe, pkg, cfg, min(8, per), N_PROMPT, 8)
                barrier()
                pr = run_peers(pe, pkg, cfg, per, N_PROMPT, N_NEW)
            cfg.max_seqs, cfg.max_ctx = 1, 1024
            wall_p, ttft50, ttft99 = max_over_ranks(dist, [pr["wall_s"], pr["ttft_ms_p50_in_engine"], pr["ttft_ms_p99_in_engine"]], f"cuda:{local}")
            extra["peers256"] = {"peers": per * world, "replicas": world, "n_prompt": N_PROMPT, "n_new": N_NEW,
                                 "aggregate_tok_per_s": per * world * N_NEW / wall_p, "wall_s": wall_p,
                                 "ttft_ms_p50_in_engine": ttft50, "ttft_ms_p99_in_engine": ttft99,
                                 "note": "max over replicas of wall time and of the per-replica TTFT percentiles"}
        except Exception as ex:
            extra["peers256"] = {"error": str(ex)}
    if rank == 0:
        peaks, peak_src = measured_peaks()
        hbm = float(peaks["hbm_gbs"])
        kv_bytes = 2 * cfg.n_layers * cfg.n_kv_heads * cfg.head_dim * 2 * (N_PROMPT + N_NEW // 2)
        ach = k_bytes / (k_ms * 1e-3) / 1e9
        extra["roofline"] = {"kernel": "gemv_kernel<bf16, rmsnorm, swiglu> (ffn gate/up, 28672x4096)", "bound": "hbm",
                             "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
                             # dram__bytes_read + dram__bytes_write of this kernel, one launch, ncu --set full
                             # (profiles/prof_gemv_bf16_r1_raw.csv): 234.94 MB + 3.62 MB vs 234.88 MB algorithmic
                             "traffic": 238562816,
                             "peak_source": peak_src,
                             "timing": "CUDA events on the engine stream, 20 launches back to back; the 235 MB matrix exceeds the 126 MB L2",
                             "bytes_per_launch": k_bytes, "ms_per_launch": k_ms}
        sach = (wbytes + kv_bytes) / (step_ms * 1e-3) / 1e9
        extra["step_roofline"] = {"bound": "hbm", "bytes_per_token": wbytes + kv_bytes, "ms_per_token": step_ms,
                                  "achieved": sach, "peak": hbm, "unit": "GB/s", "frac": sach / hbm,
                                  "ceiling_tok_per_s": hbm * 1e9 / (wbytes + kv_bytes), "launches_per_token": launches}
        # BASELINE configs[2]: the same model in Ollama's default Q4_K_M mix (N=1 only, like the CPU baseline)
        if eng is not None:
            eng.close()
        try:
            if world > 1:
                raise RuntimeError("measured at N=1 only")
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
        # BASELINE configs[3]: 256 concurrent peers, continuous batching, one engine (N=1 only)
        try:
            if world > 1:
                raise KeyError("done above")
            if os.environ.get("B200_BENCH_FAST"):
                raise RuntimeError("skipped (B200_BENCH_FAST)")
            cfg.weight_format, cfg.max_seqs, cfg.max_ctx = pkg.WEIGHTS_BF16, 256, N_PROMPT + N_NEW
            with pkg.Engine(cfg) as pe:
                run_peers(pe, pkg, cfg, 8, N_PROMPT, 8)   # warm-up: graph capture, shadow buffers
                extra["peers256"] = run_peers(pe, pkg, cfg, 256, N_PROMPT, N_NEW)
            cfg.max_seqs, cfg.max_ctx = 1, 1024
        except KeyError:
            pass
        except Exception as ex:
            extra["peers256"] = {"error": str(ex)}
        # the same workload through the HTTP boundary the reference actually calls (N=1 only)
        try:
            if world > 1 or os.environ.get("B200_BENCH_FAST"):
                raise RuntimeError("measured at N=1 only")
            extra["http"] = run_http()  # llmd is its own process with its own engine (ours are closed by now)
        except Exception as ex:
            extra["http"] = {"error": str(ex)}
        # CPU baseline: the oracle on this box's host cores, bounded sample of the same workload
        try:
            if world > 1:
                raise RuntimeError("measured at N=1 only")
            from oracle import orc
            cores = os.cpu_count() or 1
            m = cpu_weights_8b(orc, cfg)
            cpu_decode_sample(m, cfg.vocab, 1, 2)
            v, _, _ = cpu_decode_sample(m, cfg.vocab, 2, 12)
            extra["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                     "sample": "8B bf16 random-init, 2-token prompt + 12 greedy tokens (bounded sample of 512+128)"}
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
            # per request: begin (1) + tensor-core prefill (1 + 10/layer + 1 + lm_head/sampler/finish 3) + decode steps
            "gpu_launches": args.steps * (1 + 1 + 10 * cfg.n_layers + 1 + 3 + (N_NEW - 1) * launches),
            "clocks": clocks,
        }
        line.update(extra)
        print(json.dumps(line))
    elif eng is not None:
        eng.close()
    if dist is not None:
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
