import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
import __graft_entry__ as ge
from oracle import orc
from tests.util import random_tensor
pkg = ge.load_package()
rng = np.random.default_rng(0)
for wtype, name in ((orc.BF16, "bf16"), (orc.Q4_K, "q4k"), (orc.Q6_K, "q6k")):
    for rows, cols in ((4096, 4096), (4096, 14336), (28672, 4096), (6144, 4096)):
        W = random_tensor(rng, wtype, rows, cols)
        x = rng.standard_normal(cols).astype(np.float32)
        want = orc.matmul(wtype, W, rows, cols, x, mode=0)[0].astype(np.float64)
        gots = [pkg.gemv(wtype, W, rows, cols, x) for _ in range(3)]
        rms = np.sqrt((want ** 2).mean())
        errs = [np.abs(g - want).max() / rms for g in gots]
        same = all(np.array_equal(gots[0], g) for g in gots[1:])
        # fp32 reference in numpy (pairwise) to see what plain fp32 gives
        print(name, rows, cols, "max|err|/rms:", ["%.2e" % e for e in errs], "deterministic", same)
