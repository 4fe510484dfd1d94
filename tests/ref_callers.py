"""TEST INFRASTRUCTURE: loads the reference's UNMODIFIED caller files (oracle/_ref/callers/, copied byte for byte
from /root/reference by `make -C oracle callers`, sha256 recorded) on top of a chosen rasterizer backend.

  backend "ours": `diff_gaussian_rasterization` resolves to monogs_b200.diff_gaussian_rasterization
                  (what monogs_b200.install_aliases() registers) -> the sm_100a library of this repo.
  backend "ref" : `diff_gaussian_rasterization` resolves to the reference's own Python wrapper
                  (submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py, unmodified)
                  whose `_C` is oracle/ref_cuda.py = the unmodified reference CUDA kernels (oracle/_ref/libref_dgr.so).

The files loaded (reference paths): gaussian_splatting/gaussian_renderer/__init__.py (render), utils/camera_utils.py
(CameraExtrinsics, CameraIntrinsics), utils/pose_utils.py (update_pose), utils/slam_utils.py (get_loss_tracking,
get_loss_mapping), gaussian_splatting/utils/graphics_utils.py.
"""
from __future__ import annotations

import hashlib
import importlib
import importlib.util
import sys
import types
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
CALLERS = ROOT / "oracle" / "_ref" / "callers"


def available() -> bool:
    return (CALLERS / "SHA256SUMS").exists()


def verify_unmodified() -> int:
    """Every shipped file still has the sha256 `make -C oracle callers` recorded from /root/reference."""
    n = 0
    for line in (CALLERS / "SHA256SUMS").read_text().splitlines():
        digest, name = line.split(None, 1)
        got = hashlib.sha256((CALLERS / name.strip()).read_bytes()).hexdigest()
        assert got == digest, f"{name}: shipped copy differs from the reference file"
        n += 1
    return n


def _load_file(name: str, path: Path, package_dir: Path | None = None, preset: dict | None = None):
    spec = importlib.util.spec_from_file_location(
        name, path, submodule_search_locations=[str(package_dir)] if package_dir else None)
    mod = importlib.util.module_from_spec(spec)
    for k, v in (preset or {}).items():
        setattr(mod, k, v)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_stacks: dict = {}


def load_stack(backend: str) -> types.SimpleNamespace:
    """-> namespace(render, CameraExtrinsics, CameraIntrinsics, update_pose, get_loss_tracking, get_loss_mapping,
    dgr) with the reference's render() bound to `backend`'s rasterizer."""
    if backend in _stacks:
        return _stacks[backend]
    if str(CALLERS) not in sys.path:
        sys.path.insert(0, str(CALLERS))
    if backend == "ours":
        import monogs_b200
        monogs_b200.install_aliases()
        dgr = importlib.import_module("monogs_b200.diff_gaussian_rasterization")
    elif backend == "ref":
        from oracle import ref_cuda
        ref_cuda.load()
        ref_cuda.set_ring(8)
        wdir = CALLERS / "ref_wrapper" / "diff_gaussian_rasterization"
        # the reference wrapper does `from . import _C`: hand it the ctypes binding of the reference CUDA library
        sys.modules["ref_wrapper_dgr._C"] = ref_cuda
        dgr = _load_file("ref_wrapper_dgr", wdir / "__init__.py", wdir, preset={"_C": ref_cuda})
    else:
        raise ValueError(backend)
    saved = sys.modules.get("diff_gaussian_rasterization")
    sys.modules["diff_gaussian_rasterization"] = dgr
    try:
        rdir = CALLERS / "gaussian_splatting" / "gaussian_renderer"
        renderer = _load_file(f"ref_gaussian_renderer_{backend}", rdir / "__init__.py", rdir)
    finally:
        if saved is not None:
            sys.modules["diff_gaussian_rasterization"] = saved
        else:
            sys.modules.pop("diff_gaussian_rasterization", None)
    assert renderer.GaussianRasterizer is dgr.GaussianRasterizer
    camera_utils = importlib.import_module("utils.camera_utils")
    pose_utils = importlib.import_module("utils.pose_utils")
    slam_utils = importlib.import_module("utils.slam_utils")
    for m in (camera_utils, pose_utils, slam_utils):
        assert Path(m.__file__).resolve().is_relative_to(CALLERS.resolve()), m.__file__
    ns = types.SimpleNamespace(
        render=renderer.render, CameraExtrinsics=camera_utils.CameraExtrinsics,
        CameraIntrinsics=camera_utils.CameraIntrinsics, update_pose=pose_utils.update_pose,
        get_loss_tracking=slam_utils.get_loss_tracking, get_loss_mapping=slam_utils.get_loss_mapping, dgr=dgr)
    _stacks[backend] = ns
    return ns
