"""Tensor parallelism (BASELINE config 5 plumbing) on 2 GPUs of one box: TP-2 must reproduce TP-1 and the oracle.
Skipped on single-GPU boxes.  One thread per rank (ctypes releases the GIL inside the library)."""
import threading

import numpy as np
import pytest

import __graft_entry__ as ge
from oracle import orc
from tests.util import oracle_from_engine, tiny_cfg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    p = ge.load_package()
    p.lib()
    if p.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    return p


def run_ranks(pkg, cfg_fn, tp, fn):
    nccl_id = pkg.tp_unique_id()
    out, errs = [None] * tp, []

    def rank(r):
        try:
            cfg = cfg_fn()
            cfg.device = r
            with pkg.Engine(cfg, tp_rank=r, tp_size=tp, nccl_id=nccl_id) as eng:
                out[r] = fn(eng)
        except Exception as e:  # surfaced below
            errs.append((r, repr(e)))
    ts = [threading.Thread(target=rank, args=(r,)) for r in range(tp)]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    assert not errs, errs
    return out


def test_tp2_matches_tp1_and_oracle(pkg):
    mk = lambda: tiny_cfg(pkg, pkg.WEIGHTS_BF16, n_layers=2, seed=81, max_ctx=128)
    prompt = np.random.default_rng(3).integers(0, 1024, 21).astype(np.int32)
    with pkg.Engine(mk()) as e1:
        ora = oracle_from_engine(e1)
        l1_gemv = e1.forward_logits(prompt, pkg.MODE_GEMV)
        l1_gemm = e1.forward_logits(prompt, pkg.MODE_GEMM)
        ids1, _, _ = e1.generate_greedy(prompt, 24, pkg.MODE_GEMM)

    def work(eng):
        a = eng.forward_logits(prompt, pkg.MODE_GEMV)
        b = eng.forward_logits(prompt, pkg.MODE_GEMM)
        ids, ttft, dec = eng.generate_greedy(prompt, 24, pkg.MODE_GEMM)
        return a, b, ids
    res = run_ranks(pkg, mk, 2, work)
    for r in range(2):
        a, b, ids = res[r]
        assert np.abs(a - l1_gemv).max() <= 1e-4 * max(1.0, np.abs(l1_gemv).max()), r     # fp32 sums regrouped across ranks
        assert np.abs(b - l1_gemm).max() <= 3e-3 * max(1.0, np.abs(l1_gemm).max()), r     # + bf16 input rounding flips
    assert res[0][2].tolist() == res[1][2].tolist()                                        # ranks agree on every token
    ora.reset()
    for t in prompt:
        want = ora.forward([t], mode=0)
    assert np.abs(res[0][0] - want).max() <= 1e-3 * max(1.0, np.abs(want).max())
    want_ids, margins = ora.generate_greedy(prompt, 24, prefill_mode=1)
    if margins.min() > 1.2e-2:
        assert res[0][2].tolist() == want_ids.tolist() == ids1.tolist()


@pytest.mark.parametrize("push", [False, True])
def test_tp2_batched_step_ranks_agree(pkg, push, monkeypatch):
    """b200_bench_batched under tensor parallelism (vocab-parallel lm_head GEMM, per-row argmax keys max-reduced over
    the ranks): both ranks must pick the same ids, and the ids of TP-1.  The hook feeds its own tokens back for seven
    steps and a random-init model has near-ties, where regrouped fp32 sums may pick the other token and the chain
    then diverges: so several models, and most of them must agree to the last token."""
    if push:  # the all-reduce as a push over peer memory (opt-in experiment) instead of NCCL
        monkeypatch.setenv("B200_TP_PUSH", "1")
    else:
        monkeypatch.delenv("B200_TP_PUSH", raising=False)
    agree = 0
    for seed in range(82, 87):
        def mk():
            c = tiny_cfg(pkg, pkg.WEIGHTS_BF16, n_layers=2, seed=seed, max_ctx=128)
            c.max_seqs = 8
            return c
        with pkg.Engine(mk()) as e1:
            _, ids1 = e1.bench_batched(8, 30, 4)
        res = run_ranks(pkg, mk, 2, lambda eng: eng.bench_batched(8, 30, 4)[1])
        assert res[0].tolist() == res[1].tolist()
        assert ((0 <= res[0]) & (res[0] < 1024)).all()
        agree += int(res[0].tolist() == ids1.tolist())
    assert agree >= 3, agree


def test_tp_rejects_unsupported(pkg):
    cfg = tiny_cfg(pkg, pkg.WEIGHTS_Q4_K_M, n_layers=2, seed=1)
    with pytest.raises(pkg.B200Error):
        pkg.Engine(cfg, tp_rank=0, tp_size=2, nccl_id=pkg.tp_unique_id())   # k-quant weights are not sharded
    cfg = tiny_cfg(pkg, pkg.WEIGHTS_BF16, n_layers=2, seed=1, n_heads=16, n_kv_heads=4)
    with pytest.raises(pkg.B200Error):
        pkg.Engine(cfg, tp_rank=0, tp_size=3, nccl_id=pkg.tp_unique_id())   # 3 does not divide the heads
