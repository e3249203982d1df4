"""b200_attn_decode_noalloc: the decode-attention kernel for callers that may not allocate device memory.

Who needs it: a host process that has already committed all of HBM to its own allocator (an llmd replica whose
KV pool fills the GPU, a PyTorch caching allocator).  There an allocation made inside this library can fail
outright, and cudaMalloc stalls every stream on the device even when it succeeds.  The entry point takes device
pointers for everything and keeps its scratch in a fixed array inside libb200llm.so.

The file is named so that pytest collects it last: a run that stops at the first failure has then already
executed every other test.  It passes alone in a fresh process and after the other GPU tests.
Checked here, through the C ABI: same output as the existing entry point (b200_attn_decode) AND as the CPU oracle's
attention (so the check is not CUDA against CUDA only), within the attention tolerance the other tests use, on two
calls in a row (the second call finds the previous call's
partials in the shared scratch and must not be disturbed by them), and sizes beyond the built-in maximum are
refused with a clear error.
"""
import numpy as np
import pytest

import __graft_entry__ as ge
from oracle import orc

pytestmark = pytest.mark.gpu

ATTN_TOL = 2e-5  # same bar as tests/test_gpu_kernels.py::test_attn_decode


@pytest.fixture(scope="module")
def pkg():
    p = ge.load_package()
    p.lib()
    assert p.device_count() >= 1
    return p


def _paged_cache(torch, K, V, n_kv, perm):
    """fp32 [n_ctx][n_kv*128] (bf16-representable) -> bf16 pages [page][n_kv][16][128] + int32 block table."""
    n_ctx = K.shape[0]
    n_pages = (n_ctx + 15) // 16
    kp = np.zeros((n_pages, n_kv, 16, 128), np.uint16)
    vp = np.zeros((n_pages, n_kv, 16, 128), np.uint16)
    Kb = orc.f32_to_bf16(K).reshape(n_ctx, n_kv, 128)
    Vb = orc.f32_to_bf16(V).reshape(n_ctx, n_kv, 128)
    for t in range(n_ctx):
        kp[perm[t // 16], :, t % 16, :] = Kb[t]
        vp[perm[t // 16], :, t % 16, :] = Vb[t]
    dev = torch.device("cuda", 0)
    to = lambda a, dt: torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).to(dev).view(dt).contiguous()
    return to(kp, torch.bfloat16), to(vp, torch.bfloat16), torch.from_numpy(perm.astype(np.int32)).to(dev)


@pytest.mark.parametrize("n_ctx,n_heads,n_kv", [(640, 32, 8), (37, 8, 2), (1000, 64, 8)])
def test_noalloc_matches_existing_entry_point_twice(pkg, n_ctx, n_heads, n_kv):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(n_ctx + n_heads)
    n_pages = (n_ctx + 15) // 16
    perm = rng.permutation(n_pages)
    results = []
    for call in range(2):  # two calls in a row, different data, one shared scratch
        q = rng.standard_normal(n_heads * 128).astype(np.float32)
        K = orc.bf16_to_f32(orc.f32_to_bf16(rng.standard_normal((n_ctx, n_kv * 128)).astype(np.float32)))
        V = orc.bf16_to_f32(orc.f32_to_bf16(rng.standard_normal((n_ctx, n_kv * 128)).astype(np.float32)))
        want = pkg.attn_decode(q, K, V, n_heads, n_kv)  # existing entry point (allocates its own staging)
        kc, vc, bt = _paged_cache(torch, K, V, n_kv, perm)
        qd = torch.from_numpy(q).to(kc.device)
        out = torch.full((n_heads * 128,), float("nan"), device=kc.device, dtype=torch.float32)
        stream = torch.cuda.current_stream(kc.device)
        torch.cuda.synchronize()
        pkg.attn_decode_noalloc(qd.data_ptr(), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), n_ctx, n_heads, n_kv,
                                out.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        got = out.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= ATTN_TOL * max(1.0, np.abs(want).max()), (call, n_ctx, n_heads, n_kv)
        ref = orc.attention(q, K, V, n_heads, n_kv, 128)  # the CPU restatement, same bf16-representable K / V
        assert np.abs(got - ref).max() <= ATTN_TOL * max(1.0, np.abs(ref).max()), (call, n_ctx, n_heads, n_kv)
        results.append(got)
    assert not np.array_equal(results[0], results[1])  # the two calls really computed different problems


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch(pkg):
    torch = pytest.importorskip("torch")
    t = torch.zeros(16, device="cuda")
    with pytest.raises(pkg.B200Error) as ei:
        pkg.attn_decode_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 16, 128, 16, t.data_ptr(), 0)
    assert ei.value.code == -7 and "B200_NOALLOC_MAX_HEADS" in str(ei.value)
    with pytest.raises(pkg.B200Error):
        pkg.attn_decode_noalloc(0, t.data_ptr(), t.data_ptr(), t.data_ptr(), 16, 8, 2, t.data_ptr(), 0)
