"""GPU parity, kernel level: each sm_100a kernel through the C ABI vs the CPU oracle on the same inputs.
Bit-exact for integer/byte work (dequantisation, activation quantisation); stated tolerance for fp32 sums."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
from oracle import orc
from tests.util import random_tensor

pytestmark = pytest.mark.gpu

# fp32 accumulation in another order than the oracle's double sums: |err| <= TOL * sum_k |w_k x_k|
GEMV_RTOL = 1e-5


@pytest.fixture(scope="module")
def pkg():
    p = ge.load_package()
    p.lib()
    assert p.device_count() >= 1, "gpu tests need a CUDA device"
    return p


@pytest.mark.parametrize("name,wtype", [("q4k", orc.Q4_K), ("q6k", orc.Q6_K)])
def test_dequant_bit_exact_golden(pkg, golden_dir, name, wtype):
    z = np.load(os.path.join(golden_dir, f"dequant_{name}.npz"))
    got = pkg.dequant(z["blocks"], wtype).reshape(z["out"].shape)
    assert np.array_equal(got.view(np.uint32), z["out"].view(np.uint32))


@pytest.mark.parametrize("wtype,bs", [(orc.Q4_K, 144), (orc.Q6_K, 210)])
@pytest.mark.parametrize("nb", [1, 7, 4096])
def test_dequant_bit_exact_vs_oracle(pkg, wtype, bs, nb):
    rng = np.random.default_rng(nb + bs)
    blk = random_tensor(rng, wtype, nb, 256, sigma=0.05)
    assert blk.size == nb * bs
    got, want = pkg.dequant(blk, wtype), orc.dequant(blk, wtype)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_dequant_empty(pkg):
    assert pkg.dequant(np.zeros(0, np.uint8), pkg.Q4_K).size == 0


@pytest.mark.parametrize("n", [32, 4096, 14336])
def test_quant_q16_bit_exact(pkg, n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * np.exp(rng.uniform(-8, 4, n))).astype(np.float32)
    x[:32] = 0.0                      # an all-zero block (d = 0)
    if n > 64:
        x[40] = 3.0e38                # huge magnitude
    q, d = pkg.quant_q16(x)
    q0, d0 = orc.quant_q16(x)
    assert np.array_equal(q, q0) and np.array_equal(d.view(np.uint32), d0.view(np.uint32))


def _gemv_case(pkg, wtype, rows, cols, seed, norm=False):
    rng = np.random.default_rng(seed)
    W = random_tensor(rng, wtype, rows, cols)
    x = rng.standard_normal(cols).astype(np.float32) * 1.5
    if norm:
        nw = (1 + 0.1 * rng.standard_normal(cols)).astype(np.float32)
        got = pkg.gemv(wtype, W, rows, cols, x, norm_w=nw, eps=1e-5)
        xin = orc.rmsnorm(x, nw, 1e-5)
    else:
        got = pkg.gemv(wtype, W, rows, cols, x)
        xin = x
    want = orc.matmul(wtype, W, rows, cols, xin, mode=0)[0]
    # error budget from the magnitude of the terms, not of the (cancelling) result
    if wtype in (orc.Q4_K, orc.Q6_K):
        wabs = np.abs(orc.dequant(W, wtype).reshape(rows, cols))
    elif wtype == orc.F16:
        wabs = np.abs(W.view(np.float16).astype(np.float32).reshape(rows, cols))
    else:
        wabs = np.abs(orc.bf16_to_f32(W).reshape(rows, cols))
    budget = wabs @ np.abs(xin).astype(np.float64)
    err = np.abs(got.astype(np.float64) - want)
    tol = (2 if norm else 1) * GEMV_RTOL * budget  # norm: the fp32 inverse rms differs in the last ulp
    assert (err <= tol + 1e-30).all(), (float((err / (budget + 1e-30)).max()), wtype, rows, cols)
    return got, want


@pytest.mark.parametrize("rows,cols", [(4, 256), (64, 512), (1028, 4096), (4096, 14336), (6144, 4096)])
def test_gemv_bf16(pkg, rows, cols):
    _gemv_case(pkg, orc.BF16, rows, cols, seed=rows + cols)


@pytest.mark.parametrize("rows,cols", [(4, 256), (1028, 4096), (4096, 14336)])
def test_gemv_f16(pkg, rows, cols):
    """F16 tensors (the matrices of an F16 GGUF file): same kernel, the widening is a convert; used exactly in mode 0."""
    _gemv_case(pkg, orc.F16, rows, cols, seed=rows + cols + 1)


@pytest.mark.parametrize("rows,cols", [(4, 2048), (520, 4096), (4096, 14336)])
def test_gemv_q4k(pkg, rows, cols):
    _gemv_case(pkg, orc.Q4_K, rows, cols, seed=rows + cols)


@pytest.mark.parametrize("rows,cols", [(2, 2048), (514, 4096), (4096, 14336)])
def test_gemv_q6k(pkg, rows, cols):
    _gemv_case(pkg, orc.Q6_K, rows, cols, seed=rows + cols)


@pytest.mark.parametrize("wtype,rows,cols", [(orc.BF16, 512, 4096), (orc.Q4_K, 512, 4096), (orc.Q6_K, 512, 4096)])
def test_gemv_fused_rmsnorm(pkg, wtype, rows, cols):
    _gemv_case(pkg, wtype, rows, cols, seed=99, norm=True)


def test_gemv_linearity_full_size(pkg):
    """Size-independent property at the 8B lm_head shape: W.(a+b) == W.a + W.b (bf16 weights, fp32 x)."""
    rows, cols = 128256, 4096
    rng = np.random.default_rng(5)
    W = rng.integers(0, 65536, (rows, cols), dtype=np.uint16)
    W &= np.uint16(0xBFFF)  # clear the top exponent bit: no inf/nan, |w| < 2
    a = rng.standard_normal(cols).astype(np.float32)
    b = rng.standard_normal(cols).astype(np.float32)
    ya, yb, yab = (pkg.gemv(orc.BF16, W, rows, cols, v) for v in (a, b, a + b))
    scale = np.abs(ya) + np.abs(yb) + 1.0
    assert (np.abs(yab - (ya + yb)) <= 3e-4 * scale).all()


def test_gemv_rejects_bad_shapes(pkg):
    with pytest.raises(pkg.B200Error):
        pkg.gemv(orc.BF16, np.zeros((4, 100), np.uint16), 4, 100, np.zeros(100, np.float32))
    with pytest.raises(pkg.B200Error):
        pkg.gemv(orc.Q4_K, np.zeros(144 * 4 * 4, np.uint8), 4, 1024, np.zeros(1024, np.float32))


@pytest.mark.parametrize("n_ctx,n_heads,n_kv", [(1, 4, 1), (17, 8, 2), (640, 32, 8), (1000, 8, 1), (33, 2, 2)])
def test_attn_decode(pkg, n_ctx, n_heads, n_kv):
    rng = np.random.default_rng(n_ctx)
    q = rng.standard_normal(n_heads * 128).astype(np.float32)
    K = orc.bf16_to_f32(orc.f32_to_bf16(rng.standard_normal((n_ctx, n_kv * 128)).astype(np.float32)))
    V = orc.bf16_to_f32(orc.f32_to_bf16(rng.standard_normal((n_ctx, n_kv * 128)).astype(np.float32)))
    got = pkg.attn_decode(q, K, V, n_heads, n_kv)
    want = orc.attention(q, K, V, n_heads, n_kv, 128)
    assert np.abs(got - want).max() <= 2e-5 * max(1.0, np.abs(want).max())
    # run it again: the ticket counters must have reset themselves
    got2 = pkg.attn_decode(q, K, V, n_heads, n_kv)
    assert np.array_equal(got, got2)


# tcgen05 GEMM (GEMM numerics, oracle mode 1): bf16 inputs, exact products, fp32 accumulation in TMEM
@pytest.mark.parametrize("wtype,rows,cols,n", [
    (orc.BF16, 128, 64, 1), (orc.BF16, 256, 512, 7), (orc.BF16, 1024, 4096, 128), (orc.BF16, 512, 2048, 129),
    (orc.BF16, 6144, 4096, 512), (orc.BF16, 4096, 14336, 300), (orc.Q4_K, 256, 2048, 33), (orc.Q6_K, 256, 2048, 64),
    (orc.BF16, 200, 256, 40), (orc.BF16, 1032, 512, 300), (orc.BF16, 6144, 4096, 24), (orc.F16, 512, 2048, 129)])  # ragged last row tile (vocab-parallel lm_head shards: 16032 rows)
def test_gemm_tcgen05(pkg, wtype, rows, cols, n):
    rng = np.random.default_rng(rows + cols + n)
    W = random_tensor(rng, wtype, rows, cols)
    X = rng.standard_normal((n, cols)).astype(np.float32)
    got = pkg.gemm(wtype, W, rows, cols, X)
    want = orc.matmul(wtype, W, rows, cols, X, mode=1)
    if wtype in (orc.Q4_K, orc.Q6_K):
        Wf = orc.dequant(W, wtype).reshape(rows, cols)
    elif wtype == orc.F16:
        Wf = W.view(np.float16).astype(np.float32).reshape(rows, cols)
    else:
        Wf = orc.bf16_to_f32(W).reshape(rows, cols)
    budget = np.abs(X).astype(np.float64) @ np.abs(Wf).astype(np.float64).T
    err = np.abs(got.astype(np.float64) - want)
    assert (err <= 1e-5 * budget + 1e-30).all(), float((err / (budget + 1e-30)).max())


# the 256 x 256-tile variant (two accumulators filling TMEM), forced for shapes the heuristic leaves to the
# 128-row kernel; ragged token counts exercise the zero-filled part of the token tile
@pytest.mark.parametrize("rows,cols,n", [(256, 64, 129), (512, 2048, 256), (1024, 4096, 300), (3584, 1024, 513)])
def test_gemm_tcgen05_m256(pkg, rows, cols, n, monkeypatch):
    monkeypatch.setenv("B200_GEMM_M256", "1")
    rng = np.random.default_rng(7 * rows + cols + n)
    W = random_tensor(rng, orc.BF16, rows, cols)
    X = rng.standard_normal((n, cols)).astype(np.float32)
    got = pkg.gemm(orc.BF16, W, rows, cols, X)
    want = orc.matmul(orc.BF16, W, rows, cols, X, mode=1)
    Wf = orc.bf16_to_f32(W).reshape(rows, cols)
    budget = np.abs(X).astype(np.float64) @ np.abs(Wf).astype(np.float64).T
    err = np.abs(got.astype(np.float64) - want)
    assert (err <= 1e-5 * budget + 1e-30).all(), float((err / (budget + 1e-30)).max())
    monkeypatch.setenv("B200_GEMM_M256", "0")
    assert np.array_equal(got, pkg.gemm(orc.BF16, W, rows, cols, X))  # same K order: same bits as the 128-row tiles


# stream-K schedule of the 256-row tiles for up to 128 tokens (persistent CTAs, tiles cut across CTAs, partials
# through a workspace, summed by the tile's owner in CTA order): forced on small shapes, where a tile is shared by up
# to 32 CTAs; repeated calls must give the same bits (deterministic reduction, flags re-armed by the kernel)
@pytest.mark.parametrize("rows,cols,n", [  # one token tile: the weight-streaming regime of batched decode (32- and 128-token variants)
    (256, 64, 7), (512, 2048, 33), (1024, 4096, 128), (6144, 4096, 1), (28672, 4096, 8), (4096, 1024, 100), (28672, 4096, 32),
    # 256-token tiles (prefill): gate/up at 512 tokens = 224 tiles on 148 CTAs; small shapes forced, ragged token counts
    (28672, 4096, 512), (1024, 2048, 300), (512, 4096, 700), (28672, 256, 1100)])
def test_gemm_tcgen05_streamk(pkg, rows, cols, n, monkeypatch):
    rng = np.random.default_rng(11 * rows + cols + n)
    W = random_tensor(rng, orc.BF16, rows, cols)
    X = rng.standard_normal((n, cols)).astype(np.float32)
    monkeypatch.setenv("B200_GEMM_STREAMK", "1")
    monkeypatch.setenv("B200_GEMM_SK256", "1")
    got = pkg.gemm(orc.BF16, W, rows, cols, X)
    again = pkg.gemm(orc.BF16, W, rows, cols, X)
    assert np.array_equal(got, again)
    monkeypatch.setenv("B200_GEMM_STREAMK", "0")
    monkeypatch.setenv("B200_GEMM_SK256", "0")
    monkeypatch.setenv("B200_NO_STREAMK", "1")
    plain = pkg.gemm(orc.BF16, W, rows, cols, X)
    Wf = orc.bf16_to_f32(W).reshape(rows, cols)
    budget = np.abs(X).astype(np.float64) @ np.abs(Wf).astype(np.float64).T
    if rows * cols * n <= 1 << 31:  # the oracle in seconds
        want = orc.matmul(orc.BF16, W, rows, cols, X, mode=1)
        err = np.abs(got.astype(np.float64) - want)
        assert (err <= 1e-5 * budget + 1e-30).all(), float((err / (budget + 1e-30)).max())
    # against the one-CTA-per-tile schedule: same products, fp32 sums grouped differently
    assert (np.abs(got.astype(np.float64) - plain) <= 1e-5 * budget + 1e-30).all()


def _swiglu_bf16_ref(gu):
    """h = bf16(silu(g) * u) from fp32 gate/up columns interleaved g0, u0, g1, u1, ... (fp32 arithmetic, as the kernels)"""
    g, u = gu[:, 0::2].astype(np.float32), gu[:, 1::2].astype(np.float32)
    h = (g / (np.float32(1.0) + np.exp(-g, dtype=np.float32))) * u
    return orc.f32_to_bf16(h.astype(np.float32))


# SwiGLU in the GEMM epilogue (gate/up rows interleaved): every schedule that can carry it — 128-row tiles, 256 x 256
# tiles, stream-K at <= 32 tokens and at >= 256 tokens (tiles cut across CTAs, partials pulled in by the owner) —
# and the split-K fallback at 33..128 tokens; against the oracle's GEMM + fp32 SwiGLU, and against the unfused GPU path
@pytest.mark.parametrize("rows,cols,n,env", [
    (256, 512, 7, {}), (1024, 2048, 129, {}), (3584, 1024, 300, {"B200_GEMM_M256": "1"}), (28672, 4096, 8, {}),
    (2048, 1024, 100, {}), (28672, 1024, 512, {}), (28672, 512, 700, {}), (4096, 512, 300, {"B200_GEMM_SK256": "1"}), (19456, 256, 2100, {})])
def test_gemm_swiglu_epilogue(pkg, rows, cols, n, env, monkeypatch):
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    rng = np.random.default_rng(13 * rows + cols + n)
    W = random_tensor(rng, orc.BF16, rows, cols)
    X = rng.standard_normal((n, cols)).astype(np.float32)
    got = pkg.gemm_swiglu(orc.BF16, W, rows, cols, X)
    assert np.array_equal(got, pkg.gemm_swiglu(orc.BF16, W, rows, cols, X))  # deterministic (flags re-armed, fixed partial order)
    monkeypatch.setenv("B200_NO_GEMM_SWIGLU", "1")
    unfused = pkg.gemm_swiglu(orc.BF16, W, rows, cols, X)  # fp32 GEMM (split-K where the schedule wants it) + swiglu_bf16_kernel
    gu = pkg.gemm(orc.BF16, W, rows, cols, X)
    a, b, r = orc.bf16_to_f32(got), orc.bf16_to_f32(unfused), orc.bf16_to_f32(_swiglu_bf16_ref(gu))
    Wf = orc.bf16_to_f32(W).reshape(rows, cols)
    # the same products everywhere; fp32 sums are grouped differently where K is cut (split-K, stream-K), and
    # |d h| <= |silu'(g)| |u| dg + |silu(g)| du with dg, du <= 1e-5 x the GEMM's error budget, plus the bf16 rounding of h
    tol = 2.0 ** -7 * np.abs(r) + 1e-5 * ((np.abs(X) @ np.abs(Wf[0::2]).T) * 1.1 * (np.abs(gu[:, 1::2]) + 1) +
                                          (np.abs(X) @ np.abs(Wf[1::2]).T) * (np.abs(gu[:, 0::2]) + 1))
    assert (np.abs(b - r) <= tol).all(), float(np.abs(b - r).max())
    assert (np.abs(a - r) <= tol).all(), float(np.abs(a - r).max())
    assert (got == unfused).mean() > 0.98
    if rows * cols * n <= 1 << 31:
        want = _swiglu_bf16_ref(orc.matmul(orc.BF16, W, rows, cols, X, mode=1).astype(np.float32))
        assert (np.abs(a - orc.bf16_to_f32(want)) <= tol).all()
        assert (got == want).mean() > 0.98


def test_gemm_rejects_bad_shapes(pkg):
    with pytest.raises(pkg.B200Error):
        pkg.gemm(orc.BF16, np.zeros((100, 64), np.uint16), 100, 64, np.zeros((1, 64), np.float32))
