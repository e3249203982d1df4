"""Decode tok/s with the reference's actual request (no options -> Ollama's sampler defaults) vs greedy."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
for fmt in (pkg.WEIGHTS_BF16, pkg.WEIGHTS_Q4_K_M):
    cfg = pkg.llama31_8b()
    cfg.weight_format, cfg.max_ctx, cfg.eos_token = fmt, 1024, -1
    prompt = np.random.default_rng(1).integers(0, 128000, 512).astype(np.int32)
    with pkg.Engine(cfg) as eng:
        for temp in (0.0, 0.8):
            for it in range(3):
                sp = pkg.default_sampling()
                sp.temperature, sp.num_predict, sp.prefill_mode = temp, 128, pkg.MODE_GEMM
                t0 = time.perf_counter(); req = eng.submit(prompt, sp); n = 0; tf = None
                while True:
                    _, done = req.next(); n += 1
                    if tf is None: tf = time.perf_counter()
                    if done: break
                t1 = time.perf_counter(); req.free()
            print(f"fmt={fmt} temperature={temp}: {(n - 1) / (t1 - tf):.1f} tok/s, ttft {(tf - t0) * 1e3:.2f} ms")
