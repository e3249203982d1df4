"""p2p-llm-chat-go_b200 — Python host side of libb200llm.so (ctypes over the C ABI in include/b200llm.h).

This package holds no arithmetic: every number comes out of the sm_100a kernels in csrc/.  It exists so
that tests and bench.py (and tests/ref_client.py, the Python mirror of the reference's caller at
web/streamlit_app.py:89-101) can drive the same entry points `go/cmd/llmd` binds through cgo.

The directory name follows the repo convention and is not a valid Python identifier; load it with
`__graft_entry__.load_package()` (registers it as module `p2p_llm_chat_go_b200`).

There is NO fallback: if libb200llm.so is missing, importing `lib()` raises; if no CUDA device is present,
every compute entry point returns B200_ENODEV and the wrappers raise `B200Error`.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200llm.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "b200llm.h")

F32, F16, Q4_K, Q6_K, BF16 = 0, 1, 12, 14, 30
WEIGHTS_BF16, WEIGHTS_Q4_K_M = 0, 1
MODE_GEMV, MODE_GEMM = 0, 1
ENODEV = -2
ERANGE = -7
QK_K, Q4K_BYTES, Q6K_BYTES = 256, 144, 210


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200llm error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("n_layers", "d_model", "n_heads", "n_kv_heads", "head_dim", "d_ffn", "vocab", "max_ctx")] + \
               [(n, C.c_float) for n in
                ("rms_eps", "rope_theta", "rope_factor", "rope_low_freq_factor", "rope_high_freq_factor")] + \
               [("rope_orig_ctx", C.c_int32), ("weight_format", C.c_int32), ("device", C.c_int32),
                ("max_seqs", C.c_int32), ("seed", C.c_int64), ("eos_token", C.c_int32), ("reserved", C.c_int32)]


class Sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("top_p", C.c_float), ("repeat_penalty", C.c_float),
                ("top_k", C.c_int32), ("repeat_last_n", C.c_int32), ("seed", C.c_uint64),
                ("num_predict", C.c_int32), ("prefill_mode", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("total_ns", C.c_int64), ("load_ns", C.c_int64), ("prompt_eval_ns", C.c_int64),
                ("eval_ns", C.c_int64), ("prompt_eval_count", C.c_int32), ("eval_count", C.c_int32),
                ("done_reason", C.c_int32), ("reserved", C.c_int32)]


_lib = None


def lib():
    """dlopen libb200llm.so (built in-tree by __graft_entry__.build()); raises if it is not there."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built — run `python -c 'import __graft_entry__ as g; g.build()'`; "
                          "there is no Python/CPU fallback for the inference path")
    L = C.CDLL(LIB_PATH)
    vp, i64, i32, f32, u64, sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_uint64, C.c_size_t
    cfgp, smpp = C.POINTER(Config), C.POINTER(Sampling)
    sig = {
        "b200_last_error": (C.c_char_p, []),
        "b200_version": (C.c_char_p, []),
        "b200_device_count": (i32, []),
        "b200_default_sampling": (None, [smpp]),
        "b200_llama31_8b": (None, [cfgp]),
        "b200_llama31_70b": (None, [cfgp]),
        "b200_init": (i32, [cfgp, C.POINTER(vp)]),
        "b200_tp_unique_id": (i32, [vp]),
        "b200_init_tp": (i32, [cfgp, i32, i32, vp, C.POINTER(vp)]),
        "b200_shutdown": (None, [vp]),
        "b200_get_config": (i32, [vp, cfgp]),
        "b200_set_tensor": (i32, [vp, C.c_char_p, i32, vp, i64, i64]),
        "b200_tensor_info": (i32, [vp, C.c_char_p, C.POINTER(i32), C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
        "b200_get_tensor": (i32, [vp, C.c_char_p, vp, i64]),
        "b200_weight_bytes": (i64, [vp]),
        "b200_tokenize": (i32, [vp, C.c_char_p, sz, i32, vp, sz, C.POINTER(sz)]),
        "b200_detokenize": (i32, [vp, vp, sz, vp, sz, C.POINTER(sz)]),
        "b200_gguf_info": (i32, [C.c_char_p, cfgp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
        "b200_load_gguf": (i32, [C.c_char_p, i32, i32, i32, C.POINTER(vp)]),
        "b200_tok_load_gguf": (i32, [C.c_char_p, C.POINTER(vp)]),
        "b200_tok_free": (None, [vp]),
        "b200_tok_encode": (i32, [vp, C.c_char_p, sz, i32, vp, sz, C.POINTER(sz)]),
        "b200_pretokenize": (i32, [C.c_char_p, sz, vp, sz, C.POINTER(sz)]),
        "b200_tok_decode": (i32, [vp, vp, sz, vp, sz, C.POINTER(sz)]),
        "b200_submit": (i32, [vp, vp, sz, smpp, C.POINTER(vp)]),
        "b200_next": (i32, [vp, C.POINTER(C.c_int32), C.POINTER(i32), i32]),
        "b200_stats_get": (i32, [vp, C.POINTER(Stats)]),
        "b200_cancel": (None, [vp]),
        "b200_free": (None, [vp]),
        "b200_launch_count": (C.c_ulonglong, []),
        "b200_chat_template": (i32, [vp, vp, sz, C.POINTER(sz)]),
        "b200_stop_ids": (i32, [vp, vp, sz, C.POINTER(sz)]),
        "b200_special_token": (i32, [vp, C.c_char_p]),
        "b200_prefix_match": (i32, [vp, vp, sz]),
        "b200_forward_logits": (i32, [vp, vp, sz, i32, vp]),
        "b200_forward_logits_forced": (i32, [vp, vp, sz, i32, vp, sz, vp]),
        "b200_generate_greedy": (i32, [vp, vp, sz, vp, sz, i32, C.POINTER(f32), C.POINTER(f32)]),
        "b200_dequant": (i32, [i32, vp, i64, vp]),
        "b200_quant_q16": (i32, [vp, i64, vp, vp]),
        "b200_gemv": (i32, [i32, vp, i64, i64, vp, vp]),
        "b200_gemv_rmsnorm": (i32, [i32, vp, i64, i64, vp, vp, f32, vp]),
        "b200_gemm": (i32, [i32, vp, i64, i64, vp, i64, vp]),
        "b200_gemm_swiglu": (i32, [i32, vp, i64, i64, vp, i64, vp]),
        "b200_attn_decode": (i32, [vp, vp, vp, i32, i32, i32, vp]),
        "b200_attn_decode_noalloc": (i32, [vp, vp, vp, vp, i32, i32, i32, vp, vp]),
        "b200_sample": (i32, [vp, i32, smpp, vp, i32, u64, C.POINTER(C.c_int32)]),
        "b200_bench_decode": (i32, [vp, i32, i32, C.POINTER(f32), C.POINTER(i32)]),
        "b200_bench_batched": (i32, [vp, i32, i32, i32, C.POINTER(f32), vp]),
        "b200_bench_kernel": (i32, [vp, C.c_char_p, i32, C.POINTER(f32), C.POINTER(i64)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)           # AttributeError here == header/library mismatch
        fn.restype, fn.argtypes = res, args
    L._signatures = sig
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise B200Error(rc, lib().b200_last_error().decode(errors="replace"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def device_count():
    return lib().b200_device_count()


def llama31_8b():
    c = Config()
    lib().b200_llama31_8b(C.byref(c))
    return c


def llama31_70b():
    c = Config()
    lib().b200_llama31_70b(C.byref(c))
    return c


def default_sampling():
    s = Sampling()
    lib().b200_default_sampling(C.byref(s))
    return s


def row_bytes(wtype, cols):
    return {F32: cols * 4, BF16: cols * 2, Q4_K: cols // QK_K * Q4K_BYTES, Q6_K: cols // QK_K * Q6K_BYTES}[wtype]


# ---- stand-alone kernel hooks (host arrays in, host arrays out) ---------------------------------------
def dequant(blocks, wtype):
    blocks = _c(blocks, np.uint8).reshape(-1)
    nb = blocks.size // (Q4K_BYTES if wtype == Q4_K else Q6K_BYTES)
    out = np.empty(nb * QK_K, np.float32)
    _check(lib().b200_dequant(wtype, _p(blocks), nb, _p(out)))
    return out


def quant_q16(x):
    x = _c(x, np.float32).reshape(-1)
    q, d = np.empty(x.size, np.int16), np.empty(x.size // 32, np.float32)
    _check(lib().b200_quant_q16(_p(x), x.size, _p(q), _p(d)))
    return q, d


def gemv(wtype, W, rows, cols, x, norm_w=None, eps=1e-5):
    W, x = np.ascontiguousarray(W), _c(x, np.float32)
    y = np.empty(rows, np.float32)
    if norm_w is None:
        _check(lib().b200_gemv(wtype, _p(W), rows, cols, _p(x), _p(y)))
    else:
        norm_w = _c(norm_w, np.float32)
        _check(lib().b200_gemv_rmsnorm(wtype, _p(W), rows, cols, _p(x), _p(norm_w), eps, _p(y)))
    return y


def gemm(wtype, W, rows, cols, X):
    W, X = np.ascontiguousarray(W), _c(X, np.float32).reshape(-1, cols)
    Y = np.empty((X.shape[0], rows), np.float32)
    _check(lib().b200_gemm(wtype, _p(W), rows, cols, _p(X), X.shape[0], _p(Y)))
    return Y


def gemm_swiglu(wtype, W, rows, cols, X):
    """ffn gate/up GEMM with the SwiGLU epilogue; returns bf16 bit patterns [n][rows/2] (uint16)."""
    W, X = np.ascontiguousarray(W), _c(X, np.float32).reshape(-1, cols)
    H = np.empty((X.shape[0], rows // 2), np.uint16)
    _check(lib().b200_gemm_swiglu(wtype, _p(W), rows, cols, _p(X), X.shape[0], _p(H)))
    return H


def attn_decode(q, K, V, n_heads, n_kv_heads):
    q, K, V = _c(q, np.float32), _c(K, np.float32), _c(V, np.float32)
    n_ctx = K.size // (n_kv_heads * 128)
    out = np.empty(n_heads * 128, np.float32)
    _check(lib().b200_attn_decode(_p(q), _p(K), _p(V), n_ctx, n_heads, n_kv_heads, _p(out)))
    return out


def launch_count():
    return int(lib().b200_launch_count())


def attn_decode_noalloc(q_ptr, k_cache_ptr, v_cache_ptr, block_table_ptr, n_ctx, n_heads, n_kv_heads, out_ptr, stream=0):
    """Device pointers (ints) in, nothing allocated; asynchronous on `stream` (a cudaStream_t value)."""
    _check(lib().b200_attn_decode_noalloc(q_ptr, k_cache_ptr, v_cache_ptr, block_table_ptr, n_ctx, n_heads, n_kv_heads,
                                          out_ptr, stream))


def sample(logits, sp, history, step):
    logits, history = _c(logits, np.float32), _c(history, np.int32)
    tok = C.c_int32()
    _check(lib().b200_sample(_p(logits), logits.size, C.byref(sp), _p(history), history.size, step, C.byref(tok)))
    return tok.value


# ---- engine -------------------------------------------------------------------------------------------
class Request:
    def __init__(self, eng, handle):
        self.eng, self.h = eng, handle

    def next(self, timeout_ms=-1):
        tok, done = C.c_int32(), C.c_int()
        _check(lib().b200_next(self.h, C.byref(tok), C.byref(done), timeout_ms))
        return tok.value, bool(done.value)

    def tokens(self):
        while True:
            tok, done = self.next()
            yield tok
            if done:
                return

    def stats(self):
        s = Stats()
        _check(lib().b200_stats_get(self.h, C.byref(s)))
        return s

    def cancel(self):
        lib().b200_cancel(self.h)

    def free(self):
        if self.h:
            lib().b200_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def gguf_info(path):
    """(config, n_tensors, n_vocab_tokens, n_merges) of a GGUF file; needs no GPU."""
    cfg, a, b, c = Config(), C.c_int64(), C.c_int64(), C.c_int64()
    _check(lib().b200_gguf_info(path.encode(), C.byref(cfg), C.byref(a), C.byref(b), C.byref(c)))
    return cfg, a.value, b.value, c.value


def pretokenize(text):
    """Pieces of the llama-bpe pre-tokeniser (a str in, list of str out); raw bytes are accepted too (list of bytes)."""
    b = text.encode() if isinstance(text, str) else bytes(text)
    ends = np.empty(len(b) + 1, np.int32)
    n = C.c_size_t()
    _check(lib().b200_pretokenize(b, len(b), _p(ends), ends.size, C.byref(n)))
    cuts = [0] + ends[:n.value].tolist()
    pieces = [b[cuts[i]:cuts[i + 1]] for i in range(n.value)]
    return [p.decode() for p in pieces] if isinstance(text, str) else pieces


class Tok:
    """Stand-alone tokenizer handle over a GGUF vocabulary."""

    def __init__(self, path):
        self.h = C.c_void_p()
        _check(lib().b200_tok_load_gguf(path.encode(), C.byref(self.h)))

    def encode(self, text, add_bos=False):
        b = text.encode()
        ids = np.empty(len(b) + 8, np.int32)
        n = C.c_size_t()
        _check(lib().b200_tok_encode(self.h, b, len(b), int(add_bos), _p(ids), ids.size, C.byref(n)))
        return ids[:n.value].tolist()

    def decode(self, ids):
        ids = _c(ids, np.int32)
        buf = C.create_string_buffer(max(16, ids.size * 64))
        n = C.c_size_t()
        _check(lib().b200_tok_decode(self.h, _p(ids), ids.size, buf, len(buf), C.byref(n)))
        return buf.raw[:n.value].decode(errors="replace")

    def __del__(self):
        try:
            lib().b200_tok_free(self.h)
        except Exception:
            pass


def tp_unique_id():
    """128-byte NCCL id (rank 0 creates it; the host hands it to every rank)."""
    buf = C.create_string_buffer(128)
    _check(lib().b200_tp_unique_id(buf))
    return buf.raw


class Engine:
    def __init__(self, cfg, tp_rank=0, tp_size=1, nccl_id=None):
        h = C.c_void_p()
        if tp_size > 1:
            idbuf = C.create_string_buffer(nccl_id, 128)
            _check(lib().b200_init_tp(C.byref(cfg), tp_rank, tp_size, idbuf, C.byref(h)))
        else:
            _check(lib().b200_init(C.byref(cfg), C.byref(h)))
        self.h, self.cfg = h, cfg

    @classmethod
    def from_gguf(cls, path, device=0, max_ctx=2048, max_seqs=1):
        self = cls.__new__(cls)
        h = C.c_void_p()
        _check(lib().b200_load_gguf(path.encode(), device, max_ctx, max_seqs, C.byref(h)))
        self.h, self.cfg = h, Config()
        _check(lib().b200_get_config(h, C.byref(self.cfg)))
        return self

    def close(self):
        if self.h:
            lib().b200_shutdown(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_tensor(self, name, wtype, arr, rows, cols):
        arr = np.ascontiguousarray(arr)
        _check(lib().b200_set_tensor(self.h, name.encode(), wtype, _p(arr), rows, cols))

    def tensor_info(self, name):
        t, r, c, n = C.c_int(), C.c_int64(), C.c_int64(), C.c_int64()
        _check(lib().b200_tensor_info(self.h, name.encode(), C.byref(t), C.byref(r), C.byref(c), C.byref(n)))
        return t.value, r.value, c.value, n.value

    def get_tensor(self, name):
        t, r, c, n = self.tensor_info(name)
        buf = np.empty(n, np.uint8)
        _check(lib().b200_get_tensor(self.h, name.encode(), _p(buf), n))
        return t, r, c, buf

    def tensor_names(self):
        c = self.cfg
        names = ["token_embd.weight", "output_norm.weight", "output.weight"]
        for l in range(c.n_layers):
            names += [f"blk.{l}.{s}.weight" for s in ("attn_norm", "attn_q", "attn_k", "attn_v", "attn_output",
                                                      "ffn_norm", "ffn_gate", "ffn_up", "ffn_down")]
        return names

    def weight_bytes(self):
        return lib().b200_weight_bytes(self.h)

    def tokenize(self, text, add_bos=True):
        b = text.encode()
        ids = np.empty(len(b) + 8, np.int32)
        n = C.c_size_t()
        _check(lib().b200_tokenize(self.h, b, len(b), int(add_bos), _p(ids), ids.size, C.byref(n)))
        return ids[:n.value].copy()

    def detokenize(self, ids):
        ids = _c(ids, np.int32)
        buf = C.create_string_buffer(max(16, ids.size * 64))
        n = C.c_size_t()
        rc = lib().b200_detokenize(self.h, _p(ids), ids.size, buf, len(buf), C.byref(n))
        if rc == ERANGE:                      # a token longer than the guess: n holds the size needed
            buf = C.create_string_buffer(n.value)
            rc = lib().b200_detokenize(self.h, _p(ids), ids.size, buf, len(buf), C.byref(n))
        _check(rc)
        return buf.raw[:n.value].decode(errors="replace")

    def chat_template(self):
        n = C.c_size_t()
        lib().b200_chat_template(self.h, None, 0, C.byref(n))
        buf = C.create_string_buffer(max(1, n.value))
        _check(lib().b200_chat_template(self.h, buf, n.value, C.byref(n)))
        return buf.raw[:n.value].decode(errors="replace")

    def stop_ids(self):
        ids = np.empty(64, np.int32)
        n = C.c_size_t()
        _check(lib().b200_stop_ids(self.h, _p(ids), ids.size, C.byref(n)))
        return ids[:n.value].tolist()

    def special_token(self, text):
        return lib().b200_special_token(self.h, text.encode())

    def prefix_match(self, ids):
        ids = _c(ids, np.int32)
        return lib().b200_prefix_match(self.h, _p(ids), ids.size)

    def submit(self, prompt_ids, sampling=None):
        ids = _c(prompt_ids, np.int32)
        h = C.c_void_p()
        sp = C.byref(sampling) if sampling is not None else None
        _check(lib().b200_submit(self.h, _p(ids), ids.size, sp, C.byref(h)))
        return Request(self, h)

    def forward_logits(self, ids, mode=MODE_GEMV):
        ids = _c(ids, np.int32)
        out = np.empty(self.cfg.vocab, np.float32)
        _check(lib().b200_forward_logits(self.h, _p(ids), ids.size, mode, _p(out)))
        return out

    def forward_logits_forced(self, prompt, forced, prefill_mode=MODE_GEMM):
        """Logits [len(forced) + 1, vocab]: row 0 after the prompt, row k + 1 after forced[k] was fed (teacher forcing)."""
        prompt, forced = _c(prompt, np.int32), _c(forced, np.int32)
        out = np.empty((forced.size + 1, self.cfg.vocab), np.float32)
        _check(lib().b200_forward_logits_forced(self.h, _p(prompt), prompt.size, prefill_mode, _p(forced), forced.size, _p(out)))
        return out

    def generate_greedy(self, prompt, n_new, prefill_mode=MODE_GEMM):
        prompt = _c(prompt, np.int32)
        out = np.empty(n_new, np.int32)
        ttft, dec = C.c_float(), C.c_float()
        _check(lib().b200_generate_greedy(self.h, _p(prompt), prompt.size, _p(out), n_new, prefill_mode,
                                          C.byref(ttft), C.byref(dec)))
        return out, ttft.value, dec.value

    def bench_decode(self, ctx, n_steps):
        ms, nl = C.c_float(), C.c_int()
        _check(lib().b200_bench_decode(self.h, ctx, n_steps, C.byref(ms), C.byref(nl)))
        return ms.value, nl.value

    def bench_batched(self, n_seqs, ctx, n_steps):
        """(ms per step, ids of the last step) of n_seqs sequences decoding together from a ctx-token context."""
        ms = C.c_float()
        ids = np.zeros(n_seqs, np.int32)
        _check(lib().b200_bench_batched(self.h, n_seqs, ctx, n_steps, C.byref(ms), _p(ids)))
        return ms.value, ids

    def bench_kernel(self, which, iters):
        ms, nb = C.c_float(), C.c_int64()
        _check(lib().b200_bench_kernel(self.h, which.encode(), iters, C.byref(ms), C.byref(nb)))
        return ms.value, nb.value
