"""Host-side mirror of the reference interface, no GPU: names, fields, argument validation, aliases, and the
loud failure when the CUDA library is missing (there is no CPU fallback on the product path)."""
import sys

import pytest
import torch

import monogs_b200
from monogs_b200 import _abi
from monogs_b200.diff_gaussian_rasterization import (GaussianRasterizationSettings, GaussianRasterizer,
                                                     rasterize_gaussians)


def test_settings_fields_match_reference():
    # submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:173-186
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "projmatrix_raw", "sh_degree", "campos", "prefiltered", "debug")


def test_forward_signature_matches_reference():
    import inspect
    sig = inspect.signature(GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                                    "rotations", "cov3D_precomp", "theta", "rho"]
    assert callable(rasterize_gaussians) and hasattr(GaussianRasterizer, "markVisible")


def test_validation_messages_cpu():
    s = GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), torch.eye(4), 0,
                                      torch.zeros(3), False, False)
    r = GaussianRasterizer(s)
    m = torch.zeros(2, 3)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(m, m, torch.zeros(2, 1), shs=torch.zeros(2, 1, 3), colors_precomp=m, scales=m, rotations=torch.zeros(2, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, torch.zeros(2, 1), colors_precomp=m, scales=m, rotations=torch.zeros(2, 4), cov3D_precomp=torch.zeros(2, 6))


def test_cpu_tensors_are_rejected_not_silently_computed(built):
    """The product path never computes on the CPU: CPU inputs raise."""
    s = GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), torch.eye(4), 0,
                                      torch.zeros(3), False, False)
    m = torch.zeros(2, 3)
    with pytest.raises(RuntimeError):
        GaussianRasterizer(s)(m, m, torch.ones(2, 1), colors_precomp=m, scales=m + 1, rotations=torch.ones(2, 4))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_abi, "_lib", None)
    monkeypatch.setattr(_abi, "_LIB_PATH", tmp_path / "libgsr_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _abi.load()


def test_install_aliases():
    for k in ("diff_gaussian_rasterization", "diff_gaussian_rasterization._C", "simple_knn", "simple_knn._C"):
        sys.modules.pop(k, None)
    monogs_b200.install_aliases()
    from diff_gaussian_rasterization import GaussianRasterizationSettings as S2, GaussianRasterizer as R2  # noqa
    from simple_knn._C import distCUDA2  # noqa: F401
    assert S2 is GaussianRasterizationSettings and R2 is GaussianRasterizer


def test_product_package_never_imports_oracle():
    import pathlib, re
    root = pathlib.Path(monogs_b200.__file__).parent
    for p in root.rglob("*.py"):
        assert not re.search(r"^\s*(from|import)\s+oracle", p.read_text(), flags=re.M), p


def test_tracking_graph_rejects_cpu_tensors():
    import types
    import torch
    from monogs_b200 import tracking
    ci = types.SimpleNamespace(fx=torch.tensor([100.0]), fy=torch.tensor([100.0]), cx=8.0, cy=8.0, width=16, height=16)
    z = torch.zeros
    with pytest.raises(RuntimeError):
        tracking.TrackingGraph(ci, z(4, 3), z(4, 4), z(4, 3), z(4, 1), z(4, 3), z(3), z(3, 16, 16))


def test_window_chunks_cover_the_rows_in_order():
    from monogs_b200 import window
    for P in (0, 1, 4095, 4096, 100_000, 1_000_003):
        for n in (1, 2, 4, 7):
            ch = window.window_chunks(P, n)
            if P == 0:
                assert ch == []
                continue
            assert ch[0][0] == 0 and ch[-1][1] == P
            assert all(a[1] == b[0] for a, b in zip(ch[:-1], ch[1:]))          # contiguous, ascending
            assert all(c[0] % 256 == 0 and c[1] > c[0] for c in ch)            # gsr_backward_rows needs 256-aligned starts
    sl = window.column_slices()
    assert sum(s.stop - s.start for s in sl.values()) == window.N_COLS == 16
