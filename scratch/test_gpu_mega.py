"""The persistent decode kernel (csrc/decode_mega.cu: one cooperative launch per token) against the
per-projection kernels it replaces: same arithmetic in the same order, so logits and tokens must be
bit-identical; and against the oracle like every other path (tests/test_gpu_model.py runs through it by
default).  The persistent kernel is opt-in (B200_MEGA=1): it is correct but measured slower (see its header)."""
import numpy as np
import pytest

import __graft_entry__ as ge
from tests.util import engine_and_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    p = ge.load_package()
    p.lib()
    assert p.device_count() >= 1
    return p


def _run(pkg, fmt, monkeypatch, no_mega, prompt, n_new, seed):
    if no_mega:
        monkeypatch.delenv("B200_MEGA", raising=False)
    else:
        monkeypatch.setenv("B200_MEGA", "1")
    eng, _, cfg = engine_and_oracle(pkg, fmt, n_layers=3, seed=seed, source="synth")
    with eng:
        logits = eng.forward_logits(prompt, pkg.MODE_GEMV)
        ids, _, _ = eng.generate_greedy(prompt, n_new, pkg.MODE_GEMV)
        ms, launches = eng.bench_decode(len(prompt), 2)
    return logits, ids, launches


@pytest.mark.parametrize("fmt", [0, 1])
def test_mega_is_bit_identical_to_the_per_projection_kernels(pkg, fmt, monkeypatch):
    rng = np.random.default_rng(5)
    prompt = rng.integers(0, 1024, 37).astype(np.int32)
    la, ia, na = _run(pkg, fmt, monkeypatch, False, prompt, 48, 61)
    lb, ib, nb = _run(pkg, fmt, monkeypatch, True, prompt, 48, 61)
    assert na == 3, na          # persistent kernel + sampler + finish_step (parity model: vocab 1024 -> one sampler kernel)
    assert nb > 3 * 5, nb       # ~5 launches per layer
    assert np.array_equal(la, lb)
    assert ia.tolist() == ib.tolist()


def test_mega_matches_oracle_long_context(pkg, monkeypatch):
    """More KV tiles than attention splits (ctx > 32 * splits is not reachable at max_ctx 256, but several
    tiles per split are): logits after a 200-token prompt against the oracle."""
    monkeypatch.setenv("B200_MEGA", "1")
    eng, ora, cfg = engine_and_oracle(pkg, 0, n_layers=2, seed=62, source="synth")
    with eng:
        rng = np.random.default_rng(6)
        toks = rng.integers(0, cfg.vocab, 200).astype(np.int32)
        got = eng.forward_logits(toks, pkg.MODE_GEMV)
        ora.reset()
        want = ora.forward(toks, mode=0)
        assert np.abs(got - want).max() <= 1e-3 * max(1.0, np.abs(want).max())
