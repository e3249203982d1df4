"""Times the tensor-core GEMM of layer 0's matrices (bf16 8B shapes) through b200_bench_kernel.
   B200_GEMM_M256=0|1 python scratch/bench_gemm.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.n_layers = 1
cfg.max_ctx = 1024
with pkg.Engine(cfg) as eng:
    for which in ["gemm_gate_up@512", "gemm_gate_up@256", "gemm_gate_up@200", "gemm_qkv@512", "gemm_down@512", "gemm_down@256",
                  "gemm_attn_output@512", "gemm_lm_head@256", "gemm_lm_head@512"]:
        ms, flop = eng.bench_kernel(which, 20)
        print(f"M256={os.environ.get('B200_GEMM_M256','auto')} {which:24s} {ms*1e3:8.1f} us  {flop/ms/1e9:8.1f} TFLOP/s")
