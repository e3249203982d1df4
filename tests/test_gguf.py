"""GGUF model files (the on-disk format either side of the path): the reader, the
byte-level BPE tokenizer it feeds (CPU tests), and loading a whole model onto the GPU (gpu test).
Files are written with gguf-py's own GGUFWriter, i.e. by the format's reference implementation."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
from oracle import orc
from tests.util import random_tensor

gguf = pytest.importorskip("gguf")


def bytes_to_unicode():
    bs = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
    cs, n = bs[:], 0
    for b in range(256):
        if b not in bs:
            bs.append(b); cs.append(256 + n); n += 1
    return {b: chr(c) for b, c in zip(bs, cs)}


B2U = bytes_to_unicode()
U = lambda raw: "".join(B2U[b] for b in raw)
MERGES = [(b"h", b"e"), (b"l", b"l"), (b"he", b"ll"), (b"hell", b"o"), (b" ", b"w"), (b" w", b"o"), (b"r", b"l"), (b" wo", b"rl"),
          (b" worl", b"d"), (b"1", b"2"), (b"12", b"3"), (b"\xc3", b"\xa9"), (b" ", b"t"), (b" t", b"h"), (b" th", b"e")]


WHOLE_WORD = b" xyzzy"


def vocab():
    toks = [U(bytes([b])) for b in range(256)]
    for a, b in MERGES:
        toks.append(U(a + b))
    toks.append(U(WHOLE_WORD))       # in the vocabulary, but no sequence of merges builds it
    toks += ["<|begin_of_text|>", "<|eot_id|>"]
    types = [1] * (len(toks) - 2) + [3, 3]
    return toks, types, [f"{U(a)} {U(b)}" for a, b in MERGES]


def py_bpe(word, ranks, tok_id, ignore_merges=True):
    """Textbook BPE; ignore_merges (llama.cpp's setting for llama-bpe): a word that is a token is that token."""
    if ignore_merges and word in tok_id:
        return [tok_id[word]]
    sym = [bytes([b]) for b in word]
    while len(sym) > 1:
        best = min(((ranks.get((sym[i], sym[i + 1]), 1 << 30), i) for i in range(len(sym) - 1)))
        if best[0] == 1 << 30:
            break
        i = best[1]
        sym[i:i + 2] = [sym[i] + sym[i + 1]]
    return [tok_id[s] for s in sym]


def write_model(path, fmt, seed=5):
    """Tiny llama-architecture GGUF: 2 layers, d=2048, 16/4 heads, ffn 2048, vocab = len(tokens) padded to 512."""
    toks, types, merges = vocab()
    V, d, ff, nh, nkv, L = 512, 2048, 2048, 16, 4, 2
    toks += [f"<|reserved_{i}|>" for i in range(V - len(toks))]
    types += [3] * (V - len(types))
    w = gguf.GGUFWriter(path, "llama")
    w.add_block_count(L); w.add_embedding_length(d); w.add_feed_forward_length(ff); w.add_head_count(nh); w.add_head_count_kv(nkv)
    w.add_layer_norm_rms_eps(1e-5); w.add_rope_freq_base(500000.0); w.add_context_length(8192); w.add_vocab_size(V)
    w.add_tokenizer_model("gpt2"); w.add_tokenizer_pre("llama-bpe"); w.add_token_list(toks); w.add_token_types(types); w.add_token_merges(merges)
    w.add_bos_token_id(V - 2 if False else toks.index("<|begin_of_text|>")); w.add_eos_token_id(toks.index("<|eot_id|>"))
    rng = np.random.default_rng(seed)
    Q = gguf.GGMLQuantizationType
    tmat = {0: orc.BF16, 1: orc.Q4_K, 2: orc.F16}[fmt]      # fmt 2: an F16 file (every matrix IEEE half)
    tensors = {}

    def add(name, t, rows, cols):
        arr = random_tensor(rng, t, rows, cols, sigma=1.0 if name.startswith("token_embd") else None)
        tensors[name] = (t, arr, rows, cols)
        if t == orc.F32:
            w.add_tensor(name, arr.reshape(cols))
        elif t == orc.BF16:
            w.add_tensor(name, arr.view(np.uint8).reshape(rows, cols * 2), raw_dtype=Q.BF16)
        elif t == orc.F16:
            w.add_tensor(name, arr.view(np.float16).reshape(rows, cols))
        else:
            w.add_tensor(name, arr.reshape(rows, -1), raw_dtype=Q.Q4_K if t == orc.Q4_K else Q.Q6_K)
    add("token_embd.weight", tmat, V, d)
    add("output_norm.weight", orc.F32, 1, d)
    add("output.weight", {0: orc.BF16, 1: orc.Q6_K, 2: orc.F16}[fmt], V, d)
    w.add_tensor("rope_freqs.weight", np.ones(64, np.float32))
    for l in range(L):
        tv = tmat if fmt != 1 else (orc.Q6_K if l == 1 else orc.Q4_K)   # a per-tensor mix, dictated by the file
        for n, t, r, c in (("attn_q", tmat, nh * 128, d), ("attn_k", tmat, nkv * 128, d), ("attn_v", tv, nkv * 128, d),
                           ("attn_output", tmat, d, nh * 128), ("ffn_gate", tmat, ff, d), ("ffn_up", tmat, ff, d), ("ffn_down", tv, d, ff)):
            add(f"blk.{l}.{n}.weight", t, r, c)
        add(f"blk.{l}.attn_norm.weight", orc.F32, 1, d)
        add(f"blk.{l}.ffn_norm.weight", orc.F32, 1, d)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    return tensors, (V, d, ff, nh, nkv, L)


@pytest.fixture(scope="module")
def pkg():
    return ge.build()


@pytest.fixture(scope="module")
def model_file(tmp_path_factory):
    p = str(tmp_path_factory.mktemp("gguf") / "tiny-q4.gguf")
    tensors, dims = write_model(p, fmt=1)
    return p, tensors, dims


def test_gguf_info(pkg, model_file):
    path, tensors, (V, d, ff, nh, nkv, L) = model_file
    cfg, n_tensors, n_tok, n_merges = pkg.gguf_info(path)
    assert (cfg.n_layers, cfg.d_model, cfg.d_ffn, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.vocab) == (L, d, ff, nh, nkv, 128, V)
    assert abs(cfg.rope_theta - 500000.0) < 1 and abs(cfg.rms_eps - 1e-5) < 1e-9 and cfg.rope_factor == 8.0
    assert cfg.weight_format == pkg.WEIGHTS_Q4_K_M and n_tensors == len(tensors) + 1 and n_tok == V and n_merges == len(MERGES)
    with pytest.raises(pkg.B200Error):
        pkg.gguf_info(__file__)            # not a GGUF file


def test_bpe_tokenizer_matches_reference_bpe(pkg, model_file):
    tok = pkg.Tok(model_file[0])
    toks, _, _ = vocab()
    tok_id = {bytes([b]): b for b in range(256)}
    for i, (a, b) in enumerate(MERGES):
        tok_id[a + b] = 256 + i
    ranks = {m: i for i, m in enumerate(MERGES)}
    assert tok.encode("hello world") == [tok_id[b"hello"], tok_id[b" world"]]
    assert tok.encode("hello") == py_bpe(b"hello", ranks, tok_id)
    assert tok.encode("123456") == [tok_id[b"123"]] + py_bpe(b"456", ranks, tok_id)      # digits split in threes
    # " the" is a vocabulary entry: emitted whole (plain BPE would give " t" + "he" because (h,e) has the lowest rank)
    assert tok.encode("the the") == py_bpe(b"the", ranks, tok_id) + [tok_id[b" the"]]
    assert py_bpe(b" the", ranks, tok_id, ignore_merges=False) == [tok_id[b" t"], tok_id[b"he"]]
    assert tok.encode(" thel") == [tok_id[b" t"], tok_id[b"he"], ord("l")]                 # not a token: lowest rank first
    for text in ["hello world", "café 中文!", "it's  1234\n\nnew  line ", "", "a\tb"]:
        assert tok.decode(tok.encode(text)) == text                                        # lossless round trip
    assert tok.encode("x", add_bos=True)[0] == toks.index("<|begin_of_text|>")
    assert tok.decode([toks.index("<|eot_id|>"), ord("A")]) == "A"                          # specials render as nothing


def test_bpe_whole_word_lookup_and_long_words(pkg, model_file):
    """(1) llama-bpe's ignore_merges: a pre-token that is a vocabulary entry is that id, whatever the merge table says
    (llama.cpp sets ignore_merges for this pre-tokeniser; tiktoken behaves the same).  (2) The merge loop is a
    priority-queue walk: it must give what the textbook loop gives on words with many competing merges, and a
    pathological 200k-letter "word" must tokenise in well under a second per 100k letters."""
    import random
    import time
    tok = pkg.Tok(model_file[0])
    toks, _, _ = vocab()
    assert tok.encode("a xyzzy") == [ord("a"), toks.index(U(WHOLE_WORD))]
    assert tok.encode(" xyzz") == [ord(c) for c in " xyzz"]            # its prefixes are not reachable by merges
    tok_id = {bytes([b]): b for b in range(256)}
    for i, (a, b) in enumerate(MERGES):
        tok_id[a + b] = 256 + i
    ranks = {m: i for i, m in enumerate(MERGES)}
    rng = random.Random(1)
    for _ in range(300):
        word = "".join(rng.choice("helowrdt") for _ in range(rng.randint(1, 40)))
        assert tok.encode(word) == py_bpe(word.encode(), ranks, tok_id), word
        assert tok.encode(" " + word) == py_bpe(b" " + word.encode(), ranks, tok_id), word
    long_word = "hello" * 40000
    t0 = time.perf_counter()
    ids = tok.encode(long_word)
    assert time.perf_counter() - t0 < 5.0
    assert ids == [tok_id[b"hello"]] * 40000


def _mini_gguf(alignment=32, dims=(256,), ttype=0, offset=0, payload=2048):
    """A hand-packed GGUF v3 file: two metadata keys and one tensor directory entry."""
    import struct
    def s(x):
        return struct.pack("<Q", len(x)) + x
    b = b"GGUF" + struct.pack("<IQQ", 3, 1, 2)
    b += s(b"general.architecture") + struct.pack("<I", 8) + s(b"llama")
    b += s(b"general.alignment") + struct.pack("<II", 4, alignment)
    b += s(b"token_embd.weight") + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", d) for d in dims)
    b += struct.pack("<IQ", ttype, offset)
    return b + b"\0" * payload


@pytest.mark.parametrize("kw", [dict(alignment=0), dict(alignment=24), dict(offset=2**64 - 64), dict(offset=2**63),
                                dict(dims=(2**62, 2**62)), dict(dims=(2**33, 2**33, 2**33)), dict(dims=(2**64 - 1,))])
def test_crafted_headers_are_refused(pkg, tmp_path, kw):
    """Header values a fuzzer finds first: alignment 0 (division by zero), an offset that wraps the bounds check
    around 2^64, dimension products that overflow int64.  Each must be an error return, not a signal."""
    p = str(tmp_path / "m.gguf")
    open(p, "wb").write(_mini_gguf(**kw))
    with pytest.raises(pkg.B200Error):
        pkg.gguf_info(p)
    # and the same file with sane values is read as far as the reader goes (it then lacks llama.* keys / tensors)
    open(p, "wb").write(_mini_gguf())
    with pytest.raises(pkg.B200Error, match="attn_q"):
        pkg.gguf_info(p)


def test_corrupt_files_are_rejected_not_crashed(pkg, model_file, tmp_path):
    """Model files are untrusted input: truncations and random byte / length-field damage in the metadata must come
    back as an error code (or parse, if the damage was harmless), never as a crash or a hang."""
    import random
    data = open(model_file[0], "rb").read()
    hdr_end = data.find(b"tokenizer.ggml.merges") + 4096
    head = data[:hdr_end + (1 << 16)]           # metadata + the start of the tensor data: cheap to rewrite 150 times
    rng = random.Random(3)
    p2 = str(tmp_path / "c.gguf")
    clean = 0
    for it in range(150):
        b = bytearray(head)
        mode = it % 4
        if mode == 0:
            b = b[:rng.randint(0, hdr_end)]
        elif mode == 1:
            for _ in range(rng.randint(1, 8)):
                b[rng.randint(0, hdr_end - 1)] = rng.randint(0, 255)
        elif mode == 2:
            pos = rng.randint(0, hdr_end - 9)
            b[pos:pos + 8] = rng.choice([0, 1, 2**31, 2**63 - 1, 2**64 - 1, 2**40]).to_bytes(8, "little")
        else:
            pos = rng.randint(0, hdr_end - 5)
            b[pos:pos + 4] = rng.choice([0, 9, 13, 255, 2**32 - 1]).to_bytes(4, "little")
        open(p2, "wb").write(bytes(b))
        for fn in (lambda: pkg.gguf_info(p2), lambda: pkg.Tok(p2).encode("hello wörld 123")):
            try:
                fn()
            except pkg.B200Error:
                clean += 1
    assert clean >= 150        # every file here is at least truncated: gguf_info must have refused each one


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", [0, 1, 2])
def test_load_gguf_and_match_oracle(pkg, tmp_path, fmt):
    assert pkg.device_count() >= 1
    path = str(tmp_path / f"tiny-{fmt}.gguf")
    tensors, (V, d, ff, nh, nkv, L) = write_model(path, fmt, seed=9)
    eng = pkg.Engine.from_gguf(path, max_ctx=128)
    with eng:
        from tests.util import oracle_cfg_from
        ora = orc.Model(oracle_cfg_from(eng.cfg))
        for name, (t, arr, r, c) in tensors.items():
            ora.set_tensor(name, t, arr, r, c)
        ids = np.array(eng.tokenize("hello world 123", add_bos=False), np.int32)
        assert eng.detokenize(ids) == "hello world 123" and len(ids) < 8                   # merges were applied
        got = eng.forward_logits(ids, pkg.MODE_GEMV)
        ora.reset()
        for t in ids:
            want = ora.forward([t], mode=0)
        assert np.abs(got - want).max() <= 1e-3 * max(1.0, np.abs(want).max())
        out, _, _ = eng.generate_greedy(ids, 8, pkg.MODE_GEMV)
        want_ids, _ = ora.generate_greedy(ids, 8, prefill_mode=0)
        assert out.tolist() == want_ids.tolist()
        # the tensor-core path on the same file (k-quant and F16 tensors enter it rounded to bf16)
        long = np.array(eng.tokenize("hello world 123 the the hello " * 4, add_bos=False), np.int32)
        got1 = eng.forward_logits(long, pkg.MODE_GEMM)
        ora.reset()
        want1 = ora.forward(long, mode=1)
        assert np.abs(got1 - want1).max() <= 3e-3 * max(1.0, np.abs(want1).max())
