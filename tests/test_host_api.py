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
