"""Profiling target: 256 concurrent requests (512-token prompts) decoding together — continuous batching."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
import bench

pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.max_seqs, cfg.max_ctx, cfg.eos_token = 256, 640, -1
with pkg.Engine(cfg) as e:
    print(bench.run_peers(e, pkg, cfg, 256, 512, int(sys.argv[1]) if len(sys.argv) > 1 else 6))
