import sys, os, json
sys.path.insert(0, os.getcwd())
import __graft_entry__ as ge, bench
pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.max_seqs, cfg.max_ctx, cfg.eos_token = 256, 640, -1
with pkg.Engine(cfg) as e:
    print(json.dumps(bench.run_peers(e, pkg, cfg, 8, 512, 8)))
    print(json.dumps(bench.run_peers(e, pkg, cfg, 64, 512, 128)))
    print(json.dumps(bench.run_peers(e, pkg, cfg, 256, 512, 128)))
