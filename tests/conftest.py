"""pytest configuration: the `gpu` marker + shared helpers (tests/ is the only place besides smoke() and
bench.py's CPU-baseline legs that may import oracle/)."""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """Build the product library and the CPU oracle once per session (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    from monogs_b200 import build as gsr_build
    from oracle import cpu_oracle
    gsr_build.build()
    cpu_oracle.load()
    return ge


@pytest.fixture(scope="session")
def golden_mod():
    spec = importlib.util.spec_from_file_location("make_golden", ROOT / "tests" / "golden" / "make_golden.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def rel_inf(a, b, floor=1e-30):
    """||a-b||_inf / max(||b||_inf, floor) in float64 (the per-tensor metric of SURVEY 8c)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), floor))
