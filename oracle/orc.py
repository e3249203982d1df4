"""ctypes wrapper over oracle/_build/liboracle.so — TEST INFRASTRUCTURE ONLY (see oracle.cpp header).

Importers allowed: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")

F32, F16, Q4_K, Q6_K, BF16 = 0, 1, 12, 14, 30
QK_K, Q4K_BYTES, Q6K_BYTES = 256, 144, 210


def build(force=False):
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(
            os.path.join(HERE, "oracle.cpp")):
        subprocess.check_call(["make", "-C", HERE, "-s"] + (["-B"] if force else []))
    return LIB_PATH


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("n_layers", "d_model", "n_heads", "n_kv_heads", "head_dim", "d_ffn", "vocab", "max_ctx")] + \
               [(n, C.c_float) for n in
                ("rms_eps", "rope_theta", "rope_factor", "rope_low_freq_factor", "rope_high_freq_factor")] + \
               [("rope_orig_ctx", C.c_int32), ("kv_bf16", C.c_int32)]


class Sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("top_p", C.c_float), ("repeat_penalty", C.c_float),
                ("top_k", C.c_int32), ("repeat_last_n", C.c_int32), ("seed", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, i64, i32, f32 = C.c_void_p, C.c_int64, C.c_int, C.c_float
        L.orc_dequant_q4k.argtypes = [vp, i64, vp]
        L.orc_dequant_q6k.argtypes = [vp, i64, vp]
        L.orc_quant_q16.argtypes = [vp, i64, vp, vp]
        L.orc_f32_to_bf16.argtypes = [vp, i64, vp]
        L.orc_matmul.argtypes = [i32, vp, i64, i64, vp, i32, i32, vp]
        L.orc_rmsnorm.argtypes = [vp, vp, i32, f32, vp]
        L.orc_rope_table.argtypes = [C.POINTER(Config), vp]
        L.orc_rope_apply.argtypes = [vp, i32, i32, vp]
        L.orc_swiglu.argtypes = [vp, vp, i32, vp]
        L.orc_attention.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp]
        L.orc_model_new.argtypes = [C.POINTER(Config)]
        L.orc_model_new.restype = vp
        L.orc_model_free.argtypes = [vp]
        L.orc_model_reset.argtypes = [vp]
        L.orc_model_n_past.argtypes = [vp]
        L.orc_model_truncate.argtypes = [vp, i32]
        L.orc_model_set_tensor.argtypes = [vp, C.c_char_p, i32, vp, i64, i64]
        L.orc_model_forward.argtypes = [vp, vp, i32, i32, vp, i32]
        L.orc_argmax.argtypes = [vp, i32, vp]
        L.orc_sample.argtypes = [vp, i32, C.POINTER(Sampling), vp, i32, C.c_uint64]
        L.orc_generate_greedy.argtypes = [vp, vp, i32, i32, i32, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def dequant(blocks, wtype):
    blocks = _c(blocks, np.uint8).reshape(-1)
    bs = Q4K_BYTES if wtype == Q4_K else Q6K_BYTES
    nb = blocks.size // bs
    out = np.empty(nb * QK_K, np.float32)
    (lib().orc_dequant_q4k if wtype == Q4_K else lib().orc_dequant_q6k)(_p(blocks), nb, _p(out))
    return out


def quant_q16(x):
    x = _c(x, np.float32).reshape(-1)
    q = np.empty(x.size, np.int16)
    d = np.empty(x.size // 32, np.float32)
    lib().orc_quant_q16(_p(x), x.size, _p(q), _p(d))
    return q, d


def f32_to_bf16(x):
    x = _c(x, np.float32)
    y = np.empty(x.shape, np.uint16)
    lib().orc_f32_to_bf16(_p(x), x.size, _p(y))
    return y


def bf16_to_f32(h):
    return (np.asarray(h, np.uint16).astype(np.uint32) << 16).view(np.float32)


def matmul(wtype, W, rows, cols, X, mode):
    """Y[n, rows] = X[n, cols] @ W^T under numerics `mode` (0 gemv, 1 gemm)."""
    X = _c(X, np.float32).reshape(-1, cols)
    W = np.ascontiguousarray(W)
    Y = np.empty((X.shape[0], rows), np.float32)
    lib().orc_matmul(wtype, _p(W), rows, cols, _p(X), X.shape[0], mode, _p(Y))
    return Y


def rmsnorm(x, w, eps):
    x, w = _c(x, np.float32), _c(w, np.float32)
    y = np.empty_like(x)
    lib().orc_rmsnorm(_p(x), _p(w), x.size, eps, _p(y))
    return y


def rope_table(cfg):
    t = np.empty((cfg.max_ctx, cfg.head_dim // 2, 2), np.float32)
    lib().orc_rope_table(C.byref(cfg), _p(t))
    return t


def rope_apply(v, n_heads, head_dim, tab_pos):
    v = _c(v, np.float32).copy()
    tab_pos = _c(tab_pos, np.float32)
    lib().orc_rope_apply(_p(v), n_heads, head_dim, _p(tab_pos))
    return v


def swiglu(g, u):
    g, u = _c(g, np.float32), _c(u, np.float32)
    h = np.empty_like(g)
    lib().orc_swiglu(_p(g), _p(u), g.size, _p(h))
    return h


def attention(q, K, V, n_heads, n_kv_heads, head_dim):
    q, K, V = _c(q, np.float32), _c(K, np.float32), _c(V, np.float32)
    n_ctx = K.reshape(-1, n_kv_heads * head_dim).shape[0]
    out = np.empty(n_heads * head_dim, np.float32)
    lib().orc_attention(_p(q), _p(K), _p(V), n_ctx, n_heads, n_kv_heads, head_dim, _p(out))
    return out


def argmax(logits):
    logits = _c(logits, np.float32)
    m = C.c_float()
    i = lib().orc_argmax(_p(logits), logits.size, C.byref(m))
    return i, m.value


def sample(logits, sp, history, step):
    logits = _c(logits, np.float32)
    history = _c(history, np.int32)
    return lib().orc_sample(_p(logits), logits.size, C.byref(sp), _p(history), history.size, step)


class Model:
    """Holds numpy tensors alive and mirrors the oracle model API."""

    def __init__(self, cfg):
        self.cfg = cfg
        self.h = lib().orc_model_new(C.byref(cfg))
        self._keep = {}

    def set_tensor(self, name, wtype, arr, rows, cols):
        arr = np.ascontiguousarray(arr)
        self._keep[name] = arr
        rc = lib().orc_model_set_tensor(self.h, name.encode(), wtype, _p(arr), rows, cols)
        if rc:
            raise ValueError(f"orc_model_set_tensor({name}) -> {rc}")

    def reset(self):
        lib().orc_model_reset(self.h)

    def truncate(self, n):
        if lib().orc_model_truncate(self.h, n):
            raise ValueError(f"cannot rewind to {n} positions")

    def forward(self, toks, mode, all_logits=False):
        toks = _c(toks, np.int32).reshape(-1)
        out = np.empty((toks.size if all_logits else 1, self.cfg.vocab), np.float32)
        rc = lib().orc_model_forward(self.h, _p(toks), toks.size, mode, _p(out), int(all_logits))
        if rc:
            raise RuntimeError(f"orc_model_forward -> {rc}")
        return out if all_logits else out[0]

    def generate_greedy(self, prompt, n_new, prefill_mode):
        prompt = _c(prompt, np.int32)
        out = np.empty(n_new, np.int32)
        mg = np.empty(n_new, np.float32)
        rc = lib().orc_generate_greedy(self.h, _p(prompt), prompt.size, n_new, prefill_mode, _p(out), _p(mg))
        if rc:
            raise RuntimeError(f"orc_generate_greedy -> {rc}")
        return out, mg

    def __del__(self):
        try:
            lib().orc_model_free(self.h)
        except Exception:
            pass
