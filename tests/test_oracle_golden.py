"""Pins the CPU oracle against the independent known-answer sources (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from oracle import orc
from tests.util import random_tensor, tiny_from_hf_fixture


@pytest.mark.parametrize("name,wtype", [("q4k", orc.Q4_K), ("q6k", orc.Q6_K)])
def test_dequant_bit_exact_vs_gguf_py_fixture(golden_dir, name, wtype):
    z = np.load(os.path.join(golden_dir, f"dequant_{name}.npz"))
    got = orc.dequant(z["blocks"], wtype).reshape(z["out"].shape)
    assert got.view(np.uint32).tolist() == z["out"].view(np.uint32).tolist()  # bit-exact, incl. -0.0


@pytest.mark.parametrize("name,wtype,bs", [("q4k", orc.Q4_K, 144), ("q6k", orc.Q6_K, 210)])
def test_dequant_bit_exact_vs_gguf_py_live(name, wtype, bs):
    gguf = pytest.importorskip("gguf")
    from gguf import quants
    rng = np.random.default_rng(123)
    blk = rng.integers(0, 256, (257, bs), dtype=np.uint8)
    d = np.exp(rng.uniform(np.log(1e-5), np.log(1e-1), (257, 2))).astype(np.float16)
    if wtype == orc.Q4_K:
        blk[:, 0:4] = d.view(np.uint8).reshape(257, 4)
    else:
        blk[:, 208:210] = d[:, :1].copy().view(np.uint8).reshape(257, 2)
    qt = gguf.GGMLQuantizationType.Q4_K if wtype == orc.Q4_K else gguf.GGMLQuantizationType.Q6_K
    want = quants.dequantize(blk, qt).astype(np.float32).reshape(-1)
    got = orc.dequant(blk, wtype)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_whole_model_logits_vs_hf(golden_dir):
    """fp32 oracle (bf16-representable weights, fp32 KV) vs HuggingFace LlamaForCausalLM logits."""
    z = np.load(os.path.join(golden_dir, "hf_tiny_llama.npz"))
    model, cfg = tiny_from_hf_fixture(z, kv_bf16=0)
    toks = z["tokens"]
    want = z["logits"]
    # prefill all at once (mode 0 == exact fp32 activations for bf16 weights)
    got = model.forward(toks, mode=0, all_logits=True)
    err = np.abs(got - want).max()
    assert err < 2e-4 * max(1.0, np.abs(want).max()), err
    # token-by-token must give the same thing as batched prefill
    model.reset()
    for i, t in enumerate(toks):
        lg = model.forward([t], mode=0)
        assert np.abs(lg - want[i]).max() < 2e-4 * max(1.0, np.abs(want).max())


def test_rope_table_matches_hf_inv_freq(golden_dir):
    z = np.load(os.path.join(golden_dir, "hf_tiny_llama.npz"))
    _, cfg = tiny_from_hf_fixture(z, kv_bf16=0)
    tab = orc.rope_table(cfg)
    inv = z["inv_freq"].astype(np.float64)
    pos = np.arange(cfg.max_ctx)[:, None].astype(np.float64)
    assert np.abs(tab[..., 0] - np.cos(pos * inv)).max() < 1e-5
    assert np.abs(tab[..., 1] - np.sin(pos * inv)).max() < 1e-5


@pytest.mark.parametrize("wtype", [orc.Q4_K, orc.Q6_K])
@pytest.mark.parametrize("K", [4096, 14336])          # the two contraction lengths of Llama-3.1-8B
def test_kquant_int16_activation_path_is_tied_to_plain_dequant(wtype, K):
    """The decode contract for k-quant weights (oracle mode 0, mirrored by the CUDA GEMV) multiplies int16-quantised
    activations with the integer weights and applies the scales in fp32.  This test ties that contract to the plain
    definition y = dequant(W) . x evaluated in float64, so oracle and kernel cannot drift together:
        |y0 - y64| <= sum_k |w_k| * amax(block of k) / 65534      (rounding x to int16: half a step of amax/32767)
                      + 1e-6 * sum_k |w_k x_k|                    (fp32 products / scale roundings, ~8 ulp)
    and the error actually seen stays near the statistical expectation (a loose bound alone could hide a bug)."""
    rng = np.random.default_rng(K + wtype)
    rows = 192
    W = random_tensor(rng, wtype, rows, K)
    Wd = orc.dequant(W, wtype).reshape(rows, K).astype(np.float64)
    for scale in (1.0, 37.0, 1e-3):
        x = (rng.standard_normal(K) * scale * np.exp(rng.standard_normal(K))).astype(np.float32)   # heavy-tailed
        y0 = orc.matmul(wtype, W, rows, K, x, 0)[0].astype(np.float64)
        y64 = Wd @ x.astype(np.float64)
        amax = np.abs(x.reshape(-1, 32)).max(axis=1).astype(np.float64).repeat(32)
        bound = np.abs(Wd) @ (amax / 65534.0) + 1e-6 * (np.abs(Wd) @ np.abs(x.astype(np.float64)))
        err = np.abs(y0 - y64)
        assert (err <= bound).all(), (wtype, K, scale, (err / bound).max())
        # random rounding errors add in quadrature: expect ~ sqrt(sum w_k^2 step_k^2 / 12), far below the worst case
        rms_expected = np.sqrt((Wd ** 2) @ ((amax / 32767.0) ** 2 / 12.0))
        assert np.median(err / rms_expected) < 2.0, (wtype, K, scale, np.median(err / rms_expected))
        # and the quantised path is NOT the plain path in disguise: it differs by more than fp32 noise somewhere
        assert err.max() > 0


def test_matmul_gemm_mode_blocked_loop_matches_plain_float64():
    """Mode 1 (prefill numerics): bf16-rounded inputs, exact products, wide accumulation — checked for token counts
    around the 4-token blocking of the oracle's loop (n = 1..9) against numpy float64."""
    rng = np.random.default_rng(7)
    rows, K = 96, 1024
    W = orc.f32_to_bf16((rng.standard_normal((rows, K)) / 32).astype(np.float32))
    Wf = orc.bf16_to_f32(W).astype(np.float64)
    for n in range(1, 10):
        X = rng.standard_normal((n, K)).astype(np.float32)
        Xb = orc.bf16_to_f32(orc.f32_to_bf16(X)).astype(np.float64)
        got = orc.matmul(orc.BF16, W, rows, K, X, 1)
        want = Xb @ Wf.T
        assert np.abs(got - want).max() <= 1e-6 * np.abs(want).max() + 1e-7, n
