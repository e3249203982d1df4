"""A/B of the tensor-core GEMMs of layer 0 (8B bf16): python scratch/ab_gemm.py [matrix@tokens ...]; env switches per process."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.weight_format = pkg.WEIGHTS_BF16
cfg.max_ctx, cfg.n_layers = 1024, 2
cfg.max_seqs = 8
which = sys.argv[1:] or ["gate_up@512", "down@512", "qkv@512", "attn_output@512"]
env = {k: v for k, v in os.environ.items() if k.startswith("B200_")}
with pkg.Engine(cfg) as eng:
    for w in which:
        ms, fl = eng.bench_kernel("gemm_" + w, 30)
        print(f"ABG {env} {w}: {ms * 1e3:.1f} us, {fl / ms / 1e9:.0f} TFLOP/s")
