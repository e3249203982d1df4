"""A/B of the concurrent-request path: python scratch/ab_peers.py [peers] [n_new]; one warm-up burst, one timed burst."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
import bench

pkg = ge.load_package()
peers = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n_new = int(sys.argv[2]) if len(sys.argv) > 2 else 128
cfg = pkg.llama31_8b()
cfg.max_seqs, cfg.max_ctx, cfg.eos_token = peers, 640, -1
env = {k: v for k, v in os.environ.items() if k.startswith("B200_")}
with pkg.Engine(cfg) as e:
    bench.run_peers(e, pkg, cfg, peers, 512, 4)
    r = bench.run_peers(e, pkg, cfg, peers, 512, n_new)
    print(f"ABP {env} peers={peers}: {r['aggregate_tok_per_s']:.0f} tok/s, wall {r['wall_s']:.3f} s, ttft p50 {r['ttft_ms_p50']:.0f} p99 {r['ttft_ms_p99']:.0f} ms")
