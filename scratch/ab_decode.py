"""A/B of decode-step variants: python scratch/ab_decode.py bf16|q4 [n_steps]  -> one line with ms/step.
Environment switches (B200_GEMV_SHAPE*, B200_NO_ATTN_PREFETCH, ...) are read once per process: run one process per variant."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

pkg = ge.load_package()
cfg = pkg.llama31_8b()
fmt = sys.argv[1] if len(sys.argv) > 1 else "bf16"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cfg.weight_format = pkg.WEIGHTS_Q4_K_M if fmt == "q4" else pkg.WEIGHTS_BF16
cfg.max_ctx, cfg.eos_token = 1024, -1
env = {k: v for k, v in os.environ.items() if k.startswith("B200_")}
with pkg.Engine(cfg) as eng:
    best = min(eng.bench_decode(576, steps)[0] for _ in range(3))
    wb = eng.weight_bytes()
    print(f"AB {fmt} {env}: {best:.4f} ms/step = {1000 / best:.1f} tok/s, {wb / best / 1e6 / 6540.5:.3f} of 6540.5 GB/s")
