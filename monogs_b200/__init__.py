"""monogs_b200 -- B200-native (sm_100a) differentiable Gaussian-splatting rasterizer with SE(3) pose
gradients + distCUDA2, a drop-in for the rasterizer hot path of s-esposito/MonoGS.

  monogs_b200.diff_gaussian_rasterization   API-identical to the reference package
  monogs_b200.simple_knn._C.distCUDA2       API-identical to simple_knn._C
  monogs_b200.install_aliases()             registers both under their reference import names
"""
from __future__ import annotations

import importlib
import sys

__version__ = "0.1.0"


def install_aliases() -> None:
    """Make `import diff_gaussian_rasterization` / `from simple_knn._C import distCUDA2` resolve to this
    package, so gaussian_renderer/__init__.py:13-16 and gaussian_model.py:18 of MonoGS run unchanged."""
    dgr = importlib.import_module("monogs_b200.diff_gaussian_rasterization")
    knn = importlib.import_module("monogs_b200.simple_knn")
    knn_c = importlib.import_module("monogs_b200.simple_knn._C")
    sys.modules.setdefault("diff_gaussian_rasterization", dgr)
    sys.modules.setdefault("diff_gaussian_rasterization._C", importlib.import_module(
        "monogs_b200.diff_gaussian_rasterization._C"))
    sys.modules.setdefault("simple_knn", knn)
    sys.modules.setdefault("simple_knn._C", knn_c)
