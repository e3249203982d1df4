"""Batched decode step time vs number of sequences (8B bf16, ctx 576): python scratch/bench_batched.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.max_ctx, cfg.max_seqs, cfg.eos_token = 640, 256, -1
with pkg.Engine(cfg) as eng:
    for b in ([int(v) for v in sys.argv[1:]] or (1, 2, 4, 8, 16, 32, 64, 128, 256)):
        ms, ids = eng.bench_batched(b, 576, 8)
        print(f"B={b:4d}  {ms:7.3f} ms/step  {b / ms * 1e3:9.1f} tok/s")
