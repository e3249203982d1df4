"""Profiling target: tensor-core prefill of a 512-token prompt on Llama-3.1-8B (bf16 or q4) + 1 token."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.weight_format = pkg.WEIGHTS_Q4_K_M if (len(sys.argv) > 1 and sys.argv[1] == "q4") else pkg.WEIGHTS_BF16
cfg.max_ctx, cfg.eos_token = 1024, -1
prompt = np.random.default_rng(1).integers(0, 128000, 512).astype(np.int32)
with pkg.Engine(cfg) as eng:
    for i in range(3):
        ids, ttft, dec = eng.generate_greedy(prompt, 2, pkg.MODE_GEMM)
        print(f"{sys.argv[1:]} run {i}: ttft {ttft:.3f} ms, decode {dec:.3f} ms, ids {ids.tolist()}")
