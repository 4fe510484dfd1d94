"""Golden vectors for the Gaussian-insertion path: outputs of the REFERENCE's own
GaussianModel.create_viewpoint_pcd (/root/reference/gaussian_splatting/scene/gaussian_model.py:121-319), imported in
this container and run on CPU tensors.  The three packages its module imports that are absent here are stubbed:
`open3d` and `plyfile` (unused by this method) and `simple_knn._C.distCUDA2`, which is served by the CPU oracle's
restatement of simple_knn.cu (oracle/cpu_ref.cpp, pinned bit-exact against the reference CUDA kNN in
tests/golden/knn.npz).  Run from the repo root (only where /root/reference exists):

    python tests/golden/make_insert_golden.py        -> tests/golden/insert.npz
"""
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
REF = Path("/root/reference")
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(REF))


def _stub_modules():
    from oracle import cpu_oracle
    o3d = types.ModuleType("open3d")
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    knn_pkg = types.ModuleType("simple_knn")
    knn_c = types.ModuleType("simple_knn._C")
    knn_c.distCUDA2 = lambda pts: torch.from_numpy(cpu_oracle.knn(pts.detach().cpu().numpy()))
    knn_pkg._C = knn_c
    sys.modules.update({"open3d": o3d, "plyfile": ply, "simple_knn": knn_pkg, "simple_knn._C": knn_c})


def make_inputs(seed, W=160, H=120):
    g = torch.Generator().manual_seed(seed)
    rgb = torch.rand(3, H, W, generator=g)
    low = 0.6 + 4.0 * torch.rand(1, 1, H // 8 + 2, W // 8 + 2, generator=g)
    depth = torch.nn.functional.interpolate(low, size=(H, W), mode="bilinear", align_corners=False)[0, 0]
    depth = depth * (torch.rand(H, W, generator=g) > 0.12).float()           # invalid (zero) depth pixels
    render_depth = (depth + 0.02 * torch.randn(H, W, generator=g)
                    + 0.8 * (torch.rand(H, W, generator=g) > 0.93).float())   # a few pixels far behind the truth
    render_opacity = torch.rand(H, W, generator=g).clamp(0, 1) ** 0.7         # mostly opaque, some holes
    seg = torch.randint(0, 7, (H, W), generator=g, dtype=torch.int32)
    ang = torch.tensor([0.05, -0.08, 0.03])
    from monogs_b200 import synth
    R = synth.so3_exp(ang)
    T = torch.tensor([0.1, -0.05, 0.2])
    intr = dict(fx=133.85, fy=134.8, cx=80.03, cy=61.9)
    return dict(rgb=rgb, depth=depth, render_depth=render_depth[None], render_opacity=render_opacity[None],
                segmentation=seg, R=R, T=T, exposure_a=torch.tensor([0.07]), exposure_b=torch.tensor([-0.02]), **intr)


def main():
    _stub_modules()
    from gaussian_splatting.scene.gaussian_model import GaussianModel  # the reference's own code
    out = {}
    for case, (seed, init, isotropic) in {"init": (1, True, True), "extend": (2, False, True),
                                          "extend_aniso": (3, False, False)}.items():
        inp = make_inputs(seed)
        vp = types.SimpleNamespace(rgb=inp["rgb"], depth=inp["depth"], segmentation=inp["segmentation"], R=inp["R"],
                                   T=inp["T"], exposure_a=inp["exposure_a"], exposure_b=inp["exposure_b"])
        intr = types.SimpleNamespace(fx=inp["fx"], fy=inp["fy"], cx=inp["cx"], cy=inp["cy"])
        me = types.SimpleNamespace(device="cpu", isotropic=isotropic)
        torch.manual_seed(100 + seed)   # the randperm draw of :256
        res = GaussianModel.create_viewpoint_pcd(me, vp, intr, render_depth=None if init else inp["render_depth"].clone(),
                                                 render_opacity=None if init else inp["render_opacity"].clone(),
                                                 init=init)
        # replay the draw to store the indices themselves
        names = ("points_3d", "features", "scales", "rots", "opacities", "points_ids")
        for k, v in inp.items():
            out[f"{case}/in/{k}"] = v.numpy() if torch.is_tensor(v) else np.float32(v)
        for k, v in zip(names, res):
            out[f"{case}/out/{k}"] = v.numpy()
        out[f"{case}/seed"] = np.int64(100 + seed)
        out[f"{case}/init"] = np.bool_(init)
        out[f"{case}/isotropic"] = np.bool_(isotropic)
        print(case, {k: tuple(v.shape) for k, v in zip(names, res)})
    np.savez_compressed(ROOT / "tests" / "golden" / "insert.npz", **out)


if __name__ == "__main__":
    main()
