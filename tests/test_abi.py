"""CPU-side checks of the drop-in boundary: libb200llm.so loads, exports every symbol include/b200llm.h
declares, and fails loudly (no CPU fallback) when no CUDA device is present."""
import ctypes
import os
import re

import numpy as np
import pytest

import __graft_entry__ as ge


@pytest.fixture(scope="module")
def pkg():
    return ge.build()


def header_symbols():
    src = open(os.path.join(ge.ROOT, "include", "b200llm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", src)


def test_every_declared_symbol_is_exported(pkg):
    names = header_symbols()
    assert len(names) >= 30 and len(set(names)) == len(names)
    L = ctypes.CDLL(pkg.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # and the ctypes binding covers the same set
    assert set(pkg.lib()._signatures) == set(names)


def test_struct_layouts_match_header(pkg):
    assert ctypes.sizeof(pkg.Config) == 88  # 17 x 4 B, 4 B pad, int64 seed, 2 x 4 B
    assert ctypes.sizeof(pkg.Sampling) == 40
    assert ctypes.sizeof(pkg.Stats) == 48
    c = pkg.llama31_8b()
    assert (c.n_layers, c.d_model, c.n_heads, c.n_kv_heads, c.head_dim, c.d_ffn, c.vocab) == (32, 4096, 32, 8, 128, 14336, 128256)
    c = pkg.llama31_70b()
    assert (c.n_layers, c.d_model, c.n_heads, c.d_ffn) == (80, 8192, 64, 28672)
    s = pkg.default_sampling()   # Ollama defaults (the reference sends no options)
    assert (round(s.temperature, 3), s.top_k, round(s.top_p, 3), round(s.repeat_penalty, 3), s.repeat_last_n) == (0.8, 40, 0.9, 1.1, 64)


def test_no_cpu_fallback_without_a_device(pkg):
    if pkg.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.B200Error) as ei:
        pkg.Engine(pkg.llama31_8b())
    assert ei.value.code == pkg.ENODEV
    with pytest.raises(pkg.B200Error) as ei:
        pkg.dequant(np.zeros(144, np.uint8), pkg.Q4_K)
    assert ei.value.code == pkg.ENODEV
    with pytest.raises(pkg.B200Error):
        pkg.gemv(pkg.BF16, np.zeros((4, 256), np.uint16), 4, 256, np.zeros(256, np.float32))


def test_product_does_not_touch_the_oracle():
    """The shipped path may not include, link or import anything under oracle/."""
    bad = []
    for root, _, files in os.walk(ge.PKG_DIR):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".cpp", ".py", ".go")) or f == "Makefile":
                txt = open(os.path.join(root, f), errors="replace").read()
                if re.search(r"#include[^\n]*oracle|liboracle|-loracle|from oracle|import oracle|\borc_\w+\s*\(|dlopen[^\n]*oracle", txt):
                    bad.append(os.path.join(root, f))
    assert not bad, bad


def test_go_twin_binds_declared_symbols_and_serves_the_same_routes():
    """go/cmd/llmd/main.go cannot be compiled on this image (no Go toolchain), so what can be checked is checked
    statically: every C.b200_* function it calls is declared in include/b200llm.h (a renamed or removed entry point
    would otherwise only show up on a maintainer's machine), and it registers the routes csrc/llmd.cpp serves."""
    root = os.path.dirname(ge.PKG_DIR)
    go = open(os.path.join(root, "go", "cmd", "llmd", "main.go")).read()
    hdr = open(os.path.join(root, "include", "b200llm.h")).read()
    cpp = open(os.path.join(ge.PKG_DIR, "csrc", "llmd.cpp")).read()
    declared = set(re.findall(r"\b(b200_\w+)\s*\(", hdr)) | set(re.findall(r"\b(b200_\w+)\b(?=\s*;|\s*\{)", hdr))
    called = set(re.findall(r"\bC\.(b200_\w+)\(", go))
    assert called, "the twin calls nothing?"
    missing = sorted(f for f in called if f not in declared)
    assert not missing, missing
    for route in ("/api/generate", "/api/chat", "/api/tags", "/api/ps", "/api/version", "/metrics", "/suggest"):
        assert f'"{route}"' in go, route
        assert f'"{route}"' in cpp, route
    # the same environment knobs on both servers
    for knob in re.findall(r'"(LLMD_[A-Z_]+)"', cpp):
        assert knob in go, knob
