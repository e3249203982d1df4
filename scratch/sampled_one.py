import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
cfg = pkg.llama31_8b(); cfg.max_ctx, cfg.eos_token = 1024, -1
prompt = np.random.default_rng(1).integers(0, 128000, 512).astype(np.int32)
with pkg.Engine(cfg) as eng:
    sp = pkg.default_sampling(); sp.temperature, sp.num_predict, sp.prefill_mode = 0.8, 24, pkg.MODE_GEMM
    req = eng.submit(prompt, sp); print(len(list(req.tokens()))); req.free()
