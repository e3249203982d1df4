"""The llama-bpe pre-tokeniser (csrc/tokenizer.cpp, evaluated on Unicode code points) against the Llama-3 regex run by
the `regex` module - the pattern llama.cpp applies before BPE for tokenizer.ggml.pre = "llama-bpe".  CPU only."""
import random

import pytest

import __graft_entry__ as ge

regex = pytest.importorskip("regex")

LLAMA3 = regex.compile(r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")


@pytest.fixture(scope="module")
def pkg():
    p = ge.load_package()
    p.lib()
    return p


FIXED = [
    "", "a", " ", "  ", "   x", "hello world", "Hello, World!  It's 12345 o'clock...\n\n  next\tline  ",
    "I'M HERE, we'Ve THEY'LL she'D 'sup it'S", "café 中文 ١٢٣٤ ½ ² Ⅻ — naïve façade",
    "x\r\n\r\ny \n z\n", "tabs\t\there\t", "emoji \U0001F600\U0001F600 mixed\U0001F642text", "price: $1,234.56!!\n",
    "日本語のテキストと123４５６", "a b c　d", "   x", "été'S ſ 'ſ",
    "...\n\n\n", "  \n  \n  a", "don't stop 'til 99999", "Ünïcödé ΑΒΓ абв أهلا שלום हिन्दी",
    "１２３４５６７", "a1b22c333d4444", "line sep paranel",
]


def test_fixed_strings(pkg):
    for t in FIXED:
        assert pkg.pretokenize(t) == LLAMA3.findall(t), repr(t)


def test_random_strings(pkg):
    rng = random.Random(7)
    alphabet = (list("abcXYZ 019'\n\r\t.,!?-_$") +
                ["é", "ß", "中", "文", "١", "½", " ", " ", "　", "\U0001F600", "ſ", "И",
                 "ش", "", " ", "ǅ", "ª", "ⅷ", "〇", "́", "_", "​", "'s", "'RE", "'ll ", "  ", "\n\n"])
    for _ in range(3000):
        t = "".join(rng.choice(alphabet) for _ in range(rng.randint(0, 24)))
        assert pkg.pretokenize(t) == LLAMA3.findall(t), repr(t)


def test_pieces_cover_every_byte_even_invalid_utf8(pkg):
    raw = b"ok \xff\xfe caf\xc3\xa9 \xc3( \xe2\x82 \xf0\x9f\x98\x80 end\xc3"
    pieces = pkg.pretokenize(raw)
    assert b"".join(pieces) == raw and all(pieces)
