"""Shared helpers for the tests: tiny-model construction, synthetic weights in every storage type."""
import ast

import numpy as np

from oracle import orc

GGUF_LAYER = {"self_attn.q_proj": "attn_q", "self_attn.k_proj": "attn_k", "self_attn.v_proj": "attn_v",
              "self_attn.o_proj": "attn_output", "mlp.gate_proj": "ffn_gate", "mlp.up_proj": "ffn_up",
              "mlp.down_proj": "ffn_down", "input_layernorm": "attn_norm", "post_attention_layernorm": "ffn_norm"}


def make_cfg(n_layers, d_model, n_heads, n_kv_heads, head_dim, d_ffn, vocab, max_ctx, rms_eps=1e-5,
             rope_theta=500000.0, rope_factor=8.0, rope_low=1.0, rope_high=4.0, rope_orig_ctx=8192, kv_bf16=1):
    c = orc.Config()
    c.n_layers, c.d_model, c.n_heads, c.n_kv_heads, c.head_dim = n_layers, d_model, n_heads, n_kv_heads, head_dim
    c.d_ffn, c.vocab, c.max_ctx, c.rms_eps, c.rope_theta = d_ffn, vocab, max_ctx, rms_eps, rope_theta
    c.rope_factor, c.rope_low_freq_factor, c.rope_high_freq_factor = rope_factor, rope_low, rope_high
    c.rope_orig_ctx, c.kv_bf16 = rope_orig_ctx, kv_bf16
    return c


def hf_to_pairs(w, n_heads, head_dim):
    """HF rotate-half row order -> GGUF adjacent-pair row order (row 2i <- i, row 2i+1 <- i + hd/2)."""
    w = w.reshape(n_heads, 2, head_dim // 2, -1)
    return np.ascontiguousarray(w.transpose(0, 2, 1, 3)).reshape(n_heads * head_dim, -1)


def tiny_from_hf_fixture(z, kv_bf16):
    t = ast.literal_eval(str(z["config"][0]))
    cfg = make_cfg(t["n_layers"], t["d_model"], t["n_heads"], t["n_kv_heads"], t["head_dim"], t["d_ffn"],
                   t["vocab"], t["max_ctx"], t["rms_eps"], t["rope_theta"], t["rope_factor"], t["rope_low"],
                   t["rope_high"], t["rope_orig_ctx"], kv_bf16)
    m = orc.Model(cfg)
    for key in z.files:
        if not key.startswith("w:"):
            continue
        name, arr = key[2:], z[key]
        if name == "model.embed_tokens.weight":
            g = "token_embd.weight"
        elif name == "model.norm.weight":
            g = "output_norm.weight"
        elif name == "lm_head.weight":
            g = "output.weight"
        else:
            parts = name.split(".")          # model.layers.N.<...>.weight
            g = f"blk.{parts[2]}.{GGUF_LAYER['.'.join(parts[3:-1])]}.weight"
        if arr.dtype == np.uint16:
            if g.endswith("attn_q.weight"):
                arr = hf_to_pairs(arr, cfg.n_heads, cfg.head_dim)
            elif g.endswith("attn_k.weight"):
                arr = hf_to_pairs(arr, cfg.n_kv_heads, cfg.head_dim)
            m.set_tensor(g, orc.BF16, arr, arr.shape[0], arr.shape[1])
        else:
            m.set_tensor(g, orc.F32, arr.astype(np.float32), 1, arr.size)
    return m, cfg


# ---------------------------------------------------------------------------------------------------------
# Engine <-> oracle twins (GPU tests, smoke, bench): the same tensors on both sides.
def random_tensor(rng, wtype, rows, cols, sigma=None):
    """A random tensor in GGUF layout of `wtype` (uint8 view for block types / bf16 as uint16 / f32)."""
    sigma = sigma if sigma is not None else 1.0 / np.sqrt(cols)
    if wtype == orc.F32:
        return (1.0 + 0.1 * rng.standard_normal(rows * cols)).astype(np.float32)
    if wtype == orc.BF16:
        return orc.f32_to_bf16((rng.standard_normal((rows, cols)) * sigma).astype(np.float32))
    if wtype == orc.F16:   # IEEE half bits, as F16 GGUF files store them
        return (rng.standard_normal((rows, cols)) * sigma).astype(np.float16).view(np.uint16)
    nb = rows * cols // 256
    if wtype == orc.Q4_K:
        blk = rng.integers(0, 256, (nb, 144), dtype=np.uint8)
        d = (sigma / 258.3 * rng.uniform(0.75, 1.25, nb)).astype(np.float16)
        dmin = (7.5 * d.astype(np.float32)).astype(np.float16)
        blk[:, 0:2] = d.view(np.uint8).reshape(nb, 2)
        blk[:, 2:4] = dmin.view(np.uint8).reshape(nb, 2)
        return blk.reshape(-1)
    if wtype == orc.Q6_K:
        blk = rng.integers(0, 256, (nb, 210), dtype=np.uint8)   # scales: full int8 range
        d = (sigma / 1600.0 * rng.uniform(0.75, 1.25, nb)).astype(np.float16)
        blk[:, 208:210] = d.view(np.uint8).reshape(nb, 2)
        return blk.reshape(-1)
    raise ValueError(wtype)


def tiny_cfg(pkg, fmt, n_layers=2, d_model=2048, n_heads=16, n_kv_heads=4, d_ffn=2048, vocab=1024, max_ctx=256,
             seed=-1, eos=-1):
    c = pkg.Config()
    c.n_layers, c.d_model, c.n_heads, c.n_kv_heads, c.head_dim = n_layers, d_model, n_heads, n_kv_heads, 128
    c.d_ffn, c.vocab, c.max_ctx, c.rms_eps, c.rope_theta = d_ffn, vocab, max_ctx, 1e-5, 500000.0
    c.rope_factor, c.rope_low_freq_factor, c.rope_high_freq_factor, c.rope_orig_ctx = 8.0, 1.0, 4.0, 64
    c.weight_format, c.device, c.max_seqs, c.seed, c.eos_token = fmt, 0, 1, seed, eos
    return c


def oracle_cfg_from(c, kv_bf16=1):
    return make_cfg(c.n_layers, c.d_model, c.n_heads, c.n_kv_heads, c.head_dim, c.d_ffn, c.vocab, c.max_ctx,
                    c.rms_eps, c.rope_theta, c.rope_factor, c.rope_low_freq_factor, c.rope_high_freq_factor,
                    c.rope_orig_ctx, kv_bf16)


def oracle_from_engine(eng, names=None):
    """Build the oracle twin by reading every tensor back through b200_get_tensor (GGUF layout)."""
    ora = orc.Model(oracle_cfg_from(eng.cfg))
    for name in names or eng.tensor_names():
        t, r, c, buf = eng.get_tensor(name)
        ora.set_tensor(name, t, buf, r, c)
    return ora


def engine_and_oracle(pkg, fmt, n_layers=2, seed=0, source="synth", **kw):
    """source='synth': device-side random init, read back for the oracle.
       source='numpy': host-side random tensors pushed through b200_set_tensor (independent of synth)."""
    cfg = tiny_cfg(pkg, fmt, n_layers=n_layers, seed=seed if source == "synth" else -1, **kw)
    eng = pkg.Engine(cfg)
    if source == "synth":
        return eng, oracle_from_engine(eng), cfg
    rng = np.random.default_rng(seed)
    ora = orc.Model(oracle_cfg_from(cfg))
    for name in eng.tensor_names():
        t, r, c, _ = eng.tensor_info(name)
        arr = random_tensor(rng, t, r, c, sigma=1.0 if name == "token_embd.weight" else None)
        eng.set_tensor(name, t, arr, r, c)
        ora.set_tensor(name, t, arr, r, c)
    return eng, ora, cfg
