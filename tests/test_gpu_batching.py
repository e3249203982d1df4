"""Continuous batching: several peers' requests decode together on the tensor-core path (numerics mode 1);
token ids must equal the CPU oracle's for every sequence, whatever the others in the batch are doing."""
import threading

import numpy as np
import pytest

import __graft_entry__ as ge
from oracle import orc
from tests.util import engine_and_oracle, tiny_cfg

pytestmark = pytest.mark.gpu
TOL = 3e-3  # same logit tolerance as the prefill tests (bf16 GEMM inputs)


@pytest.fixture(scope="module")
def pkg():
    p = ge.load_package()
    p.lib()
    assert p.device_count() >= 1
    return p


def oracle_batched_greedy(ora, prompt, n_new):
    """Prefill in GEMM numerics (first token from the mode-0 lm_head), then every decode step in GEMM numerics."""
    ora.reset()
    lg = ora.forward(prompt, mode=1)
    out, margins = [], []
    for _ in range(n_new):
        tok, m = orc.argmax(lg)
        out.append(tok)
        margins.append(m)
        lg = ora.forward([tok], mode=1, all_logits=True)[0]
    return out, min(margins)


def decisive_prompts(ora, cfg, lengths, n_new):
    prompts, wants = [], []
    seed = 100
    for n in lengths:
        while True:
            seed += 1
            p = np.random.default_rng(seed).integers(0, cfg.vocab, n).astype(np.int32)
            w, m = oracle_batched_greedy(ora, p, n_new)
            if m > 4 * TOL:
                break
        prompts.append(p)
        wants.append(w)
    return prompts, wants


@pytest.mark.parametrize("fmt", [0, 1])
def test_batched_decode_matches_oracle(pkg, fmt):
    eng, ora, cfg = engine_and_oracle(pkg, fmt, n_layers=2, seed=71, source="synth", max_ctx=128)
    eng.close()
    cfg.max_seqs, cfg.reserved, cfg.seed = 8, 1, 71          # reserved bit 0: always decode on the batched path
    n_new = 12
    prompts, wants = decisive_prompts(ora, cfg, [3, 17, 40, 9, 64, 5, 23, 31, 12, 2], n_new)
    with pkg.Engine(cfg) as eng:
        sp = pkg.default_sampling()
        sp.temperature, sp.num_predict, sp.prefill_mode = 0.0, n_new, pkg.MODE_GEMM
        reqs = [eng.submit(p, sp) for p in prompts]       # 10 requests on 8 slots: two wait for a free slot
        got = [list(r.tokens()) for r in reqs]
        stats = [r.stats() for r in reqs]
        [r.free() for r in reqs]
        for i, (g, w) in enumerate(zip(got, wants)):
            assert g == w, (i, g, w)
        assert all(s.eval_count == n_new and s.done_reason == 2 and s.prompt_eval_count == len(p) for s, p in zip(stats, prompts))


def test_batched_decode_long_contexts(pkg):
    """Batched decode with contexts of several hundred to ~1000 tokens and very different lengths in one step: the
    tensor-core decode attention walks 2..16 KV tiles per sequence (ragged last tiles, per-sequence block tables)."""
    eng, ora, cfg = engine_and_oracle(pkg, 0, n_layers=2, seed=73, source="synth", max_ctx=1100)
    eng.close()
    cfg.max_seqs, cfg.reserved, cfg.seed = 4, 1, 73
    n_new = 10
    prompts, wants = decisive_prompts(ora, cfg, [1000, 130, 577, 65], n_new)
    with pkg.Engine(cfg) as eng:
        sp = pkg.default_sampling()
        sp.temperature, sp.num_predict, sp.prefill_mode = 0.0, n_new, pkg.MODE_GEMM
        reqs = [eng.submit(p, sp) for p in prompts]
        got = [list(r.tokens()) for r in reqs]
        [r.free() for r in reqs]
        for i, (g, w) in enumerate(zip(got, wants)):
            assert g == w, (i, g, w)


def test_mixed_traffic_completes_and_single_request_uses_fast_path(pkg):
    eng, ora, cfg = engine_and_oracle(pkg, 0, n_layers=2, seed=72, source="synth", max_ctx=128)
    eng.close()
    cfg.max_seqs, cfg.seed = 4, 72
    with pkg.Engine(cfg) as eng:
        sp = pkg.default_sampling()
        sp.temperature, sp.num_predict, sp.prefill_mode = 0.0, 20, pkg.MODE_GEMV
        # a lone request decodes on the batch-1 path: exact mode-0 parity
        p0 = np.random.default_rng(1).integers(0, cfg.vocab, 11).astype(np.int32)
        r = eng.submit(p0, sp)
        got = list(r.tokens())
        r.free()
        want, _ = ora.generate_greedy(p0, 20, prefill_mode=0)
        assert got == want.tolist()
        # bursts from several threads, some cancelled, some sampled: everything terminates with the right counts
        results = {}

        def peer(i):
            s = pkg.default_sampling()
            s.num_predict, s.seed, s.prefill_mode = 5 + i, i, pkg.MODE_GEMM
            s.temperature = 0.0 if i % 2 else 0.8
            rq = eng.submit(np.random.default_rng(i).integers(0, cfg.vocab, 4 + 3 * i).astype(np.int32), s)
            if i == 3:
                rq.cancel()
            toks = []
            try:
                for t in rq.tokens():
                    toks.append(t)
            except pkg.B200Error:
                pass
            results[i] = (toks, rq.stats().done_reason)
            rq.free()
        ts = [threading.Thread(target=peer, args=(i,)) for i in range(9)]
        [t.start() for t in ts]
        [t.join(timeout=120) for t in ts]
        assert len(results) == 9
        for i, (toks, reason) in results.items():
            if i == 3:
                assert reason in (2, 3)
            else:
                assert len(toks) == 5 + i and reason == 2, (i, toks, reason)
        # and afterwards the engine still serves a lone request exactly
        r = eng.submit(p0, sp)
        assert list(r.tokens()) == want.tolist()
        r.free()


def test_bench_batched_hook_is_deterministic(pkg):
    """b200_bench_batched: the batched step driven synchronously (what bench.py times for BASELINE config 5's
    bs32 line).  Fresh engine: every sequence sees the same (zero) history, so all rows pick the same token, and a
    second engine repeats the first one's ids."""
    def run():
        cfg = tiny_cfg(pkg, 0, n_layers=2, seed=71, max_ctx=128)
        cfg.max_seqs = 8
        with pkg.Engine(cfg) as eng:
            ms, ids = eng.bench_batched(8, 40, 5)
            with pytest.raises(pkg.B200Error):
                eng.bench_batched(9, 40, 5)   # more sequences than slots
        return ms, ids
    ms, ids = run()
    assert ms > 0 and ((0 <= ids) & (ids < 1024)).all()
    assert len(set(ids.tolist())) == 1
    assert run()[1].tolist() == ids.tolist()


def test_one_slot_engine_serves_queued_sampled_requests_like_solo_runs(pkg):
    """An engine with a single sequence slot (LLMD_MAX_SEQS=1) never has the batched path's per-slot state: a
    request that finds another one waiting behind it must keep decoding on the batch-1 graph, with its own sampling
    options, exactly as if it were alone.  Two sampled requests queued back to back == the same two run one by one."""
    cfg = tiny_cfg(pkg, 0, n_layers=2, seed=77, max_ctx=128)
    cfg.max_seqs, cfg.reserved = 1, 2   # reserved bit 1: no prefix reuse, so the repeat runs recompute exactly what the solo runs did
    prompts = [np.random.default_rng(s).integers(0, cfg.vocab, n).astype(np.int32) for s, n in ((1, 9), (2, 14))]

    def sampling(i):
        s = pkg.default_sampling()                     # temperature 0.8, top_k 40, top_p 0.9, repeat penalty 1.1
        s.num_predict, s.seed, s.prefill_mode = 24, 1000 + i, pkg.MODE_GEMM
        return s
    with pkg.Engine(cfg) as eng:
        solo = []
        for i, p in enumerate(prompts):
            r = eng.submit(p, sampling(i))
            solo.append(list(r.tokens()))
            r.free()
        reqs = [eng.submit(p, sampling(i)) for i, p in enumerate(prompts)]   # the second waits while the first decodes
        both = [list(r.tokens()) for r in reqs]
        [r.free() for r in reqs]
    assert both == solo
    assert len(set(solo[0])) > 4                       # really sampled, not a constant


def test_set_tensor_after_a_gemm_forward_refreshes_the_tensor_core_copy(pkg):
    """k-quant engines keep bf16 copies of the matrices for the tensor-core path.  Replacing a tensor after those
    copies exist must reach both paths: set -> GEMM forward -> set again -> GEMM forward == oracle with the new tensor."""
    from tests.util import random_tensor
    eng, ora, cfg = engine_and_oracle(pkg, 1, n_layers=2, seed=78, source="synth", max_ctx=128)
    with eng:
        toks = np.random.default_rng(5).integers(0, cfg.vocab, 40).astype(np.int32)
        before = eng.forward_logits(toks, pkg.MODE_GEMM)          # builds the bf16 shadows
        rng = np.random.default_rng(6)
        for name in ("blk.0.ffn_up.weight", "blk.1.attn_v.weight", "blk.1.attn_output.weight"):
            t, r, c, _ = eng.tensor_info(name)
            arr = random_tensor(rng, t, r, c)
            eng.set_tensor(name, t, arr, r, c)
            ora.set_tensor(name, t, arr, r, c)
        after = eng.forward_logits(toks, pkg.MODE_GEMM)
        ora.reset()
        want = ora.forward(toks, mode=1)
        assert np.abs(after - want).max() <= TOL * max(1.0, np.abs(want).max())
        assert np.abs(after - before).max() > 10 * TOL             # the replacement really changed the model
        gemv = eng.forward_logits(toks, pkg.MODE_GEMV)             # and the decode path sees the same new weights
        ora.reset()
        for tk in toks:
            want0 = ora.forward([tk], mode=0)
        assert np.abs(gemv - want0).max() <= 1e-3 * max(1.0, np.abs(want0).max())


def test_prefix_kv_reuse_across_a_peers_successive_suggestions(pkg):
    """the UI re-sends the same chat history with one more message per click
    (web/streamlit_app.py:117-133, :161-165).  A request whose prompt starts with what a free slot's cache still holds
    is prefilled from the first differing token only: prompt_eval_count drops to the new tokens, the answer is the one
    a cold engine gives (checked against the CPU oracle), and an unrelated prompt is not affected."""
    eng, ora, cfg = engine_and_oracle(pkg, 0, n_layers=2, seed=79, source="synth", max_ctx=512)
    eng.close()
    cfg.max_seqs, cfg.seed = 2, 79
    n_new = 10
    rng = np.random.default_rng(9)
    history = rng.integers(0, cfg.vocab, 300).astype(np.int32)
    sp = pkg.default_sampling()
    sp.temperature, sp.num_predict, sp.prefill_mode = 0.0, n_new, pkg.MODE_GEMM

    def run(eng, prompt):
        r = eng.submit(prompt, sp)
        toks = list(r.tokens())
        st = r.stats()
        r.free()
        return toks, st

    def oracle(prompt):  # lone request: tensor-core prefill (mode 1), then the batch-1 decode path (mode 0)
        w, m = ora.generate_greedy(prompt, n_new, prefill_mode=1)
        return w.tolist(), m.min()
    with pkg.Engine(cfg) as eng:
        first, st1 = run(eng, history)
        assert st1.prompt_eval_count == 300
        # same history + the first reply + a new message of 20 tokens
        longer = np.concatenate([history, np.array(first[:-1], np.int32), rng.integers(0, cfg.vocab, 20).astype(np.int32)])
        second, st2 = run(eng, longer)
        assert st2.prompt_eval_count == len(longer) - (300 + n_new - 1), st2.prompt_eval_count   # only the new tokens
        assert st2.load_ns + st2.prompt_eval_ns < st1.load_ns + st1.prompt_eval_ns                 # and a shorter TTFT
        want, margin = oracle(longer)
        if margin > 4 * TOL:
            assert second == want
        # an identical prompt again: everything but its last token is cached
        third, st3 = run(eng, longer)
        assert st3.prompt_eval_count == 1 and (third == second or margin <= 4 * TOL)
        # a different peer's prompt shares nothing: full prefill, right answer
        other = rng.integers(0, cfg.vocab, 120).astype(np.int32)
        got, st4 = run(eng, other)
        assert st4.prompt_eval_count == 120
        want_o, m_o = oracle(other)
        if m_o > 4 * TOL:
            assert got == want_o
        # ... and it went to the other slot, so the first peer's history is still warm
        again, st5 = run(eng, np.concatenate([longer, np.array(second[:3], np.int32)]))
        assert st5.prompt_eval_count <= 4
    cfg.reserved = 2                                   # bit 1: prefix reuse off
    with pkg.Engine(cfg) as eng:
        a, s1 = run(eng, history)
        b, s2 = run(eng, history)
        assert s2.prompt_eval_count == 300 and a == b == first
