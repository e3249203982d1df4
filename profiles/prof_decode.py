"""Profiling target: a few decode steps of Llama-3.1-8B (bf16 or q4) at ctx 576 through the C ABI.
   python profiles/prof_decode.py bf16|q4     (run plain first, then under ncu; see profiles/README.md)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.weight_format = pkg.WEIGHTS_Q4_K_M if (len(sys.argv) > 1 and sys.argv[1] == "q4") else pkg.WEIGHTS_BF16
cfg.max_ctx, cfg.eos_token = 1024, -1
with pkg.Engine(cfg) as eng:
    ms, n = eng.bench_decode(576, 2)
    print(f"{sys.argv[1:]}: {ms:.3f} ms/step, {n} launches/step, {eng.weight_bytes()} weight bytes/step")
