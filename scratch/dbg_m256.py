import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
from util import engine_and_oracle
for n in (257, 300, 511):
    for ov in ("0", "1"):
        os.environ["B200_GEMM_M256"] = ov
        eng, ora, cfg = engine_and_oracle(pkg, 0, n_layers=2, seed=53, source="synth", max_ctx=512)
        with eng:
            rng = np.random.default_rng(9)
            toks = rng.integers(0, cfg.vocab, n).astype(np.int32)
            got = eng.forward_logits(toks, pkg.MODE_GEMM)
            ora.reset()
            want = ora.forward(toks, mode=1)
            print(n, ov, "err", np.abs(got - want).max(), "max", np.abs(want).max(), flush=True)
