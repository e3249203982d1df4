"""Pins the CPU oracle against the independent known-answer sources (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from oracle import orc
from tests.util import tiny_from_hf_fixture


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
