"""The C-ABI library loads on a machine without a GPU and exports every symbol include/skrnn.h
declares; compute calls fail loudly instead of falling back to the CPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "skrnn.h")).read()
    return sorted(set(re.findall(r"SKRNN_API[^;]*?\b(skrnn_[a-z_0-9]+)\s*\(", src)))


def test_header_declares_what_python_binds():
    from sketch_rnn_b200 import _lib
    assert _declared_symbols() == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol():
    from sketch_rnn_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in _lib.load().skrnn_version()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from sketch_rnn_b200 import _lib
    with pytest.raises(_lib.SkrnnError, match="no CPU fallback"):
        _lib.Handle(model="lstm", rnn_size=16, num_layers=2, num_mixture=3)


def test_product_does_not_import_oracle():
    """oracle/ is test infrastructure: nothing under sketch_rnn_b200/, train.py or sample.py may use it."""
    bad = []
    files = [os.path.join(ROOT, f) for f in ("train.py", "sample.py") if os.path.exists(os.path.join(ROOT, f))]
    for d, _, fs in os.walk(os.path.join(ROOT, "sketch_rnn_b200")):
        files += [os.path.join(d, f) for f in fs if f.endswith((".py", ".cu", ".cuh", ".h"))]
    for f in files:
        txt = open(f).read()
        if re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M) or "oracle/" in txt and f.endswith(".py"):
            bad.append(f)
    assert not bad, bad
