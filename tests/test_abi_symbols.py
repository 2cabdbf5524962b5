"""CPU-side checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads,
and exports every entry point include/ssb200.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "ssb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ssb_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def library():
    from silent_speech_b200 import build
    path = build.build(verbose=False)
    return ctypes.CDLL(path)


def test_header_symbols_exported(library):
    names = _declared()
    assert len(names) >= 10
    missing = [n for n in names if not hasattr(library, n)]
    assert not missing, missing


def test_version_and_error_channel(library):
    assert library.ssb_version() >= 100
    library.ssb_last_error.restype = ctypes.c_char_p
    # argument validation happens before any CUDA call: a NULL pointer is reported, not dereferenced
    rc = library.ssb_colsum(None, ctypes.c_int64(4), ctypes.c_int64(4), ctypes.c_int64(4), None, None)
    assert rc != 0 and b"null pointer" in library.ssb_last_error()


def test_no_cpu_fallback():
    """The product path refuses CPU tensors instead of silently computing elsewhere."""
    from silent_speech_b200 import BiGRUClassifier
    from silent_speech_b200._lib import SsbError
    m = BiGRUClassifier(180, 10, use_roi=False).eval()
    with pytest.raises(SsbError):
        m(torch.zeros(1, 5, 180), torch.tensor([5]))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "silent_speech_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
