"""CPU checks of the boundary: the C-ABI library loads, exports every symbol include/tts_b200.h declares, the ctypes
signature table mirrors the header, and argument validation fails loudly without a GPU (no compute calls here)."""
import ctypes as C
import os
import re

import pytest
import torch

from transformertts_b200 import _ext

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "tts_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tts_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _ext.load()
    names = _declared()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tts_b200.h but not exported by libtts_b200.so"


def test_ctypes_table_matches_header():
    assert sorted(_ext.SIGNATURES) == _declared()


def test_abi_version_and_workspace_size():
    lib = _ext.load()
    assert lib.tts_abi_version() == 3
    assert lib.tts_workspace_bytes(0, 10, 10) == 0
    a, b = lib.tts_workspace_bytes(4, 150, 800), lib.tts_workspace_bytes(8, 150, 800)
    assert 0 < a < b and a % 1024 == 0


def test_forward_rejects_bad_arguments_without_touching_the_gpu():
    lib = _ext.load()
    w = _ext.Weights()
    assert lib.tts_forward(C.byref(w), None, None, None, None, 0, 10, 10, 0, None, None, None, None, None, 0, None) != 0
    assert b"positive" in lib.tts_last_error()
    assert lib.tts_forward(C.byref(w), None, None, None, None, 1, 6000, 10, 0, None, None, None, None, None, 0, None) != 0
    assert b"positional-encoding" in lib.tts_last_error()


def test_product_package_does_not_import_the_oracle():
    import transformertts_b200
    pkg = os.path.dirname(transformertts_b200.__file__)
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f)).read()
                assert "tts_oracle" not in text and "import oracle" not in text, f


def test_model_without_cuda_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from transformertts_b200 import TransformerTTS
    with pytest.raises(Exception):
        m = TransformerTTS(device="cuda").eval()
        m(torch.zeros(1, 4, dtype=torch.long), torch.zeros(1, 8, 80), torch.tensor([4]), torch.tensor([8]))
