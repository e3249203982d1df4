import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
import __graft_entry__ as ge
from oracle import orc
from tests.util import oracle_from_engine
pkg = ge.load_package()
cfg = pkg.llama31_8b()
cfg.n_layers = int(sys.argv[1]) if len(sys.argv) > 1 else 32
cfg.weight_format, cfg.seed, cfg.max_ctx, cfg.eos_token = 0, 0, 64, -1
with pkg.Engine(cfg) as eng:
    ora = oracle_from_engine(eng)
    for prompt in ([128000], [128000, 9906], [5, 6, 7]):
        prompt = np.array(prompt, np.int32)
        got = eng.forward_logits(prompt, 0)
        ora.reset()
        for t in prompt: want = ora.forward([t], mode=0)
        err = np.abs(got - want)
        print(len(prompt), "maxerr", err.max(), "at", err.argmax(), "scale", np.abs(want).max(), "n_bad", (err > 1e-3).sum(),
              "first bad", np.nonzero(err > 1e-3)[0][:8], "argmax", got.argmax(), want.argmax())
