"""Generates the golden fixtures that pin oracle/oracle.cpp.  Run in the build container:

    python tests/golden/make_golden.py

Sources of truth (the reference repo itself holds no test, fixture or golden vector for the inference path,
SURVEY.md §4 / §8c, so these are the independent known-answer sources available offline):
  * gguf-py 0.19.0  gguf.quants.dequantize      -> dequant_q4k.npz, dequant_q6k.npz (bit-exact targets)
  * transformers 5.5.0 LlamaForCausalLM (fp32, rope_type=llama3) -> hf_tiny_llama.npz (logit targets)
Nothing here reads /root/reference; the fixtures travel to the GPU box, the generators need not.
"""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def rand_fp16_scales(rng, n, lo=1e-4, hi=2e-2, signed=True):
    v = np.exp(rng.uniform(np.log(lo), np.log(hi), n)).astype(np.float16)
    if signed:
        v = np.where(rng.random(n) < 0.25, -v, v).astype(np.float16)
    return v


def make_q4k(rng, nb):
    blk = rng.integers(0, 256, (nb, 144), dtype=np.uint8)
    d = rand_fp16_scales(rng, nb)
    dmin = rand_fp16_scales(rng, nb)
    # edge cases: zeros, max nibbles/scales, fp16 subnormal d, largest normal-ish d
    blk[0, :] = 0
    blk[1, 4:] = 0xFF
    d[2] = np.float16(6e-8)
    dmin[2] = np.float16(3e-7)
    d[3] = np.float16(1000.0)
    dmin[3] = np.float16(-512.0)
    blk[:, 0:2] = d.view(np.uint8).reshape(nb, 2)
    blk[:, 2:4] = dmin.view(np.uint8).reshape(nb, 2)
    blk[0, 0:4] = 0
    return blk


def make_q6k(rng, nb):
    blk = rng.integers(0, 256, (nb, 210), dtype=np.uint8)
    d = rand_fp16_scales(rng, nb)
    blk[0, :] = 0
    blk[1, :208] = 0xFF
    blk[2, 192:208] = 0x80  # int8 scale -128
    d[3] = np.float16(6e-8)
    blk[:, 208:210] = d.view(np.uint8).reshape(nb, 2)
    blk[0, 208:210] = 0
    return blk


def golden_dequant():
    import gguf
    from gguf import quants
    rng = np.random.default_rng(20261106)
    for name, mk, qt in (("q4k", make_q4k, gguf.GGMLQuantizationType.Q4_K),
                         ("q6k", make_q6k, gguf.GGMLQuantizationType.Q6_K)):
        blk = mk(rng, 96)
        out = quants.dequantize(blk, qt).astype(np.float32).reshape(blk.shape[0], 256)
        assert np.isfinite(out).all()
        np.savez_compressed(os.path.join(HERE, f"dequant_{name}.npz"), blocks=blk, out=out,
                            source=f"gguf-py {getattr(gguf, '__version__', '0.19.0')} quants.dequantize")
        print(name, blk.shape, out.shape, float(np.abs(out).max()))


TINY = dict(n_layers=2, d_model=64, n_heads=4, n_kv_heads=2, head_dim=16, d_ffn=128, vocab=128, max_ctx=64,
            rms_eps=1e-5, rope_theta=500000.0, rope_factor=8.0, rope_low=1.0, rope_high=4.0, rope_orig_ctx=16)


def bf16_round(a):
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) >> 16
    return u.astype(np.uint16)


def golden_hf():
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    t = TINY
    cfg = LlamaConfig(
        vocab_size=t["vocab"], hidden_size=t["d_model"], intermediate_size=t["d_ffn"],
        num_hidden_layers=t["n_layers"], num_attention_heads=t["n_heads"], num_key_value_heads=t["n_kv_heads"],
        head_dim=t["head_dim"], max_position_embeddings=t["max_ctx"], rms_norm_eps=t["rms_eps"],
        rope_theta=t["rope_theta"], tie_word_embeddings=False, attention_bias=False, mlp_bias=False,
        rope_scaling={"rope_type": "llama3", "factor": t["rope_factor"], "low_freq_factor": t["rope_low"],
                      "high_freq_factor": t["rope_high"], "original_max_position_embeddings": t["rope_orig_ctx"]},
    )
    cfg._attn_implementation = "eager"
    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg).float().eval()
    rng = np.random.default_rng(7)
    saved = {}
    with torch.no_grad():
        for name, p in model.named_parameters():
            if p.ndim == 2:
                w = rng.standard_normal(p.shape).astype(np.float32) * (2.0 / np.sqrt(p.shape[1]))
                h = bf16_round(w)                      # bf16-representable => BF16 storage is exact
                w = (h.astype(np.uint32) << 16).view(np.float32)
                saved[name] = h
            else:
                w = (1.0 + 0.2 * rng.standard_normal(p.shape)).astype(np.float32)
                saved[name] = w
            p.copy_(torch.from_numpy(w.copy()))
        toks = rng.integers(0, t["vocab"], 40).astype(np.int32)
        logits = model(torch.from_numpy(toks.astype(np.int64))[None]).logits[0].numpy().astype(np.float32)
    inv_freq = model.model.rotary_emb.inv_freq.numpy().astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "hf_tiny_llama.npz"), tokens=toks, logits=logits, inv_freq=inv_freq,
                        config=np.array([repr(t)]), **{"w:" + k: v for k, v in saved.items()})
    print("hf logits", logits.shape, float(logits.std()), "inv_freq", inv_freq)


if __name__ == "__main__":
    golden_dequant()
    golden_hf()
