"""The C-ABI library loads without a GPU and exports every symbol include/ptb.h declares; argument
validation that needs no device behaves as documented (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ptb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ptb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(ptb):
    L = ptb.load_library()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"libptb.so does not export {n}"
    assert sorted(ptb.SYMBOLS) == names, "api.SYMBOLS and include/ptb.h disagree"


def test_struct_sizes_match_header(ptb):
    # ptb_camera: 9 floats; ptb_params: 15 int32 (include/ptb.h)
    assert ctypes.sizeof(ptb.Camera) == 36
    assert ctypes.sizeof(ptb.Params) == 15 * 4


def test_create_fails_loudly_without_gpu(ptb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ptb.PtbError) as e:
        ptb.PathTracer(0)
    assert "no CPU fallback" in str(e.value)


def test_null_context_is_rejected(ptb):
    L = ptb.load_library()
    assert L.ptb_render(None, 1) == -1
    assert L.ptb_build_bvh(None) == -1
    assert L.ptb_structure_bytes(None, 1) == 0
    assert L.ptb_last_error(None) is not None


def test_product_does_not_import_the_oracle():
    """Only tests/, smoke() and bench.py's CPU legs may touch oracle/."""
    pkg = os.path.join(ROOT, "bachelor-thesis-pathtracer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "ptoracle" not in txt and "oracle_lib" not in txt and "oracle/" not in txt, f
