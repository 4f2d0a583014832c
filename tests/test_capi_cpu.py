"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/splat_cuda.h declares, and fails loudly (never silently falls back) without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from gaussian_splat_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "splat_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(splat_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = capi.lib()
    names = _declared_functions()
    assert len(names) >= 13
    for n in names:
        assert hasattr(L, n), n
    assert sorted(capi.EXPORTS) == names
    assert L.splat_version() >= 100


def test_struct_layouts_match_header():
    assert C.sizeof(capi.SplatOpts) == 32
    assert C.sizeof(capi.SplatStatus) == 56
    assert capi.RECORD_DTYPE.itemsize == 48
    # SplatInstance: src/instanced_splat_renderer.h:6-11 = 14 floats, 56 bytes, no padding
    from gaussian_splat_b200 import scene
    inst, _ = scene.make_instances(10, 1)
    assert inst.dtype == np.float32 and inst.shape == (10, 14) and inst.strides == (56, 4)


def test_argument_errors_do_not_need_a_gpu():
    L = capi.lib()
    assert L.splat_create(0, None) == -1
    assert b"NULL" in L.splat_last_error_string()
    assert L.splat_upload(None, None, 0) == -1
    assert L.splat_render(None, None, None, None, None, 16, 16, None) == -1
    assert L.splat_destroy(None) == 0


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.SplatError):
        capi.Context(0)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under gaussian_splat_b200/ may reference it."""
    pkg = os.path.join(ROOT, "gaussian_splat_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.lower() or f in ("scene.py",) and "from oracle" not in txt, os.path.join(dp, f)
