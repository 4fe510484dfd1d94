"""The C-ABI boundary without a GPU: the library loads, exports every symbol include/gsr_b200.h declares,
the ctypes table mirrors the header one to one, size queries are sane, and argument errors are reported
before any device work (no compute calls here)."""
import ctypes as C
import re

import pytest

from conftest import ROOT
from monogs_b200 import _abi

HEADER = (ROOT / "include" / "gsr_b200.h").read_text()


def _declared():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    return sorted(set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", body)))


def test_header_symbols_exported(built):
    lib = _abi.load()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gsr_b200.h but not exported"


def test_ctypes_table_matches_header(built):
    assert sorted(_abi.SIGNATURES) == _declared()
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    for name, (_, args) in _abi.SIGNATURES.items():
        m = re.search(name + r"\s*\((.*?)\)\s*;", body, flags=re.S)
        params = m.group(1).strip()
        n = 0 if params in ("", "void") else len(params.split(","))
        assert n == len(args), (name, n, len(args))


def test_no_torch_types_in_header():
    assert "at::" not in HEADER and "Tensor" not in HEADER and 'extern "C"' in HEADER


def test_version_and_arch(built):
    lib = _abi.load()
    assert b"sm_100a" in lib.gsr_version()


def test_workspace_sizes(built):
    lib = _abi.load()
    P, W, H = 100_000, 640, 480
    g = lib.gsr_geom_bytes(P, W, H)
    assert g % 256 == 0 and 48 * P <= g <= 120 * P
    im = lib.gsr_image_bytes(W, H)
    assert im % 256 == 0 and im >= 8 * W * H + 8 * 1200
    assert lib.gsr_point_list_bytes(0) >= 4 and lib.gsr_point_list_bytes(1000) >= 4000
    assert lib.gsr_binning_bytes(1_000_000) >= 16 * 1_000_000
    assert lib.gsr_backward_scratch_bytes(P) >= 48 * P
    assert lib.gsr_knn_bytes(P) >= 32 * P
    assert lib.gsr_geom_bytes(2 * P, W, H) > g  # monotone


def test_argument_errors_without_device_work(built):
    lib = _abi.load()
    R = C.c_int64(-7)
    # P == 0: nothing to do, R = 0 (reference skips the kernels: rasterize_points.cu:85)
    assert lib.gsr_forward_plan(None, 0, 0, 0, 64, 64, *[None] * 5, 1.0, *[None] * 5, 1.0, 1.0, 0, None, None,
                                C.byref(R), 0) == 0
    assert R.value == 0
    # bad sizes
    assert lib.gsr_forward_plan(None, -1, 0, 0, 64, 64, *[None] * 5, 1.0, *[None] * 5, 1.0, 1.0, 0, None, None,
                                C.byref(R), 0) == -1
    assert lib.gsr_forward_plan(None, 4, 0, 0, 0, 64, *[None] * 5, 1.0, *[None] * 5, 1.0, 1.0, 0, None, None,
                                C.byref(R), 0) == -1
    # missing required pointers
    assert lib.gsr_forward_plan(None, 4, 0, 0, 64, 64, *[None] * 5, 1.0, *[None] * 5, 1.0, 1.0, 0, None, None,
                                C.byref(R), 0) == -1
    assert lib.gsr_forward_render(None, 0, 64, 64, None, None, None, None, None, None, 0, None, None, None, None,
                                  0) == 0
    assert lib.gsr_forward_render(None, 4, 64, 64, None, None, None, None, None, None, 0, None, None, None, None,
                                  0) == -1
    assert lib.gsr_mark_visible(None, 0, None, None, None, None) == 0
    assert lib.gsr_mark_visible(None, 3, None, None, None, None) == -1
    assert lib.gsr_knn_dist2(None, 0, None, None, None) == 0
    assert lib.gsr_knn_dist2(None, 5, None, None, None) == -1
    assert lib.gsr_sort_pairs_u64(None, 0, None, None, None, None, 40, None) == 0
    assert lib.gsr_sort_pairs_u64(None, 5, None, None, None, None, 0, None) == -1


def test_error_strings():
    with pytest.raises(RuntimeError, match="GSR_ENOCOLOR"):
        _abi.check(-2, "x")
    with pytest.raises(RuntimeError, match="CUDA error 700"):
        _abi.check(700, "x")
    _abi.check(0, "x")
