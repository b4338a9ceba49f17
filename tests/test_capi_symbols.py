"""CPU test: the C-ABI library loads and exports every symbol include/splatloc_b200.h declares
(and the ctypes table in splatloc_b200/_capi.py binds exactly those).  No compute calls."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "splatloc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sl_[a-z0-9_]+)\s*\(", text)) - {"sl_alloc_fn"})


def test_library_exports_every_declared_symbol():
    from splatloc_b200 import _capi, build
    build.build()
    assert os.path.exists(_capi.LIB_PATH)
    lib = ctypes.CDLL(_capi.LIB_PATH)
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(_capi.SIGNATURES) == names


def test_size_queries_and_error_text():
    from splatloc_b200 import _capi
    lib = _capi.load()
    assert lib.sl_version().startswith(b"splatloc_b200")
    assert lib.sl_geometry_bytes(1000) > 79 * 1000
    assert lib.sl_binning_bytes(1000) > 24 * 1000 + 12 * 1000
    assert lib.sl_image_bytes(640, 480) > 4 * 640 * 480 + 8 * 1200
    assert lib.sl_knn_scratch_bytes(1000) > 16 * 1000
    # argument errors are reported without touching a GPU
    rc = lib.sl_knn(-1, None, None, None, 0, None)
    assert rc == -2 and b"P < 0" in lib.sl_last_error()


def test_product_path_has_no_cpu_fallback():
    import pytest
    import torch
    from splatloc_b200 import _C
    with pytest.raises(RuntimeError, match="CUDA"):
        _C.rasterize_gaussians(torch.zeros(4), torch.zeros(8, 3), torch.zeros(8, 4), torch.zeros(8, 1), torch.zeros(8, 3),
                               torch.zeros(8, 4), 1.0, torch.Tensor([]), torch.Tensor([]).bool(), torch.eye(4),
                               torch.eye(4), 1.0, 1.0, 32, 32, torch.Tensor([]), 0, torch.zeros(3), False, False)
    with pytest.raises(RuntimeError, match="CUDA"):
        _C.distCUDA2(torch.zeros(8, 3))
    # and nothing under splatloc_b200/ imports the oracle
    pkg = os.path.join(ROOT, "splatloc_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src.replace("no oracle", ""), f"{f} mentions the oracle"
