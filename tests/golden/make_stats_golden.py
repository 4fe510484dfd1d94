"""Golden vectors for the mapper statistics (row f4): produced by EXECUTING the reference's own code --
Mapper.optimize_map (/root/reference/utils/slam_mapper.py:244-500: the occlusion-aware visibility dict :399-404, the
observation count + prune mask :407-440, the max_radii_2d update :452-456) and
GaussianModel.add_densification_stats (/root/reference/gaussian_splatting/scene/gaussian_model.py:888-892) -- on CPU
tensors in this container.  The renderer and the loss are replaced by stand-ins that hand optimize_map seeded render
packages (radii, n_touched, a viewspace tensor whose .grad is known after backward); everything between those and
the statistics is the reference's unmodified code.  Packages absent here that its modules import but these code
paths never use are stubbed (open3d, plyfile, cv2, fused_ssim, simple_knn, diff_gaussian_rasterization).

    python tests/golden/make_stats_golden.py        -> tests/golden/stats.npz
"""
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
REF = Path("/root/reference")
P, K = 20_011, 8
WINDOW = [19, 17, 14, 11, 9, 6, 3, 0]          # keyframe ids of a full window (descending, as the mapper keeps them)


def case_inputs(seed):
    """Seeded per-keyframe rasterizer outputs + the model's running statistics (CPU tensors)."""
    g = torch.Generator().manual_seed(seed)
    radii = [torch.randint(-3, 40, (P,), generator=g, dtype=torch.int32).clamp_min(0) for _ in range(K)]
    n_touched = [(torch.randint(0, 3, (P,), generator=g, dtype=torch.int32) *
                  torch.randint(0, 2, (P,), generator=g, dtype=torch.int32)) for _ in range(K)]
    grads = [torch.randn(P, 3, generator=g) * 1e-3 for _ in range(K)]
    max_radii = torch.rand(P, generator=g) * 30
    accum = torch.rand(P, 1, generator=g)
    denom = torch.randint(0, 9, (P, 1), generator=g).float()
    kf_idx = torch.randint(0, 20, (P,), generator=g, dtype=torch.int32)
    return dict(radii=radii, n_touched=n_touched, grads=grads, max_radii=max_radii, accum=accum, denom=denom,
                kf_idx=kf_idx)


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def main():
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(REF))
    _stub("open3d")
    _stub("plyfile", PlyData=object, PlyElement=object)
    _stub("fused_ssim", fused_ssim=None)
    try:
        import cv2  # noqa: F401
    except Exception:
        _stub("cv2")
    knn = _stub("simple_knn")
    knn._C = _stub("simple_knn._C", distCUDA2=None)
    _stub("diff_gaussian_rasterization", GaussianRasterizationSettings=None, GaussianRasterizer=None)
    from gaussian_splatting.scene.gaussian_model import GaussianModel   # the reference's own code
    from utils import slam_mapper                                       # the reference's own code

    out = {}
    for case, (seed, prune, first_time) in {"stats": (1, False, True), "prune_first": (2, True, False),
                                            "prune_later": (3, True, True)}.items():
        c = case_inputs(seed)

        class Model:  # the attributes optimize_map and add_densification_stats touch, nothing else
            add_densification_stats = GaussianModel.add_densification_stats
            get_xyz = get_rotation = get_scaling = get_opacity = get_features = None

            def __init__(self):
                self.max_radii_2d = c["max_radii"].clone()
                self.xyz_gradient_accum = c["accum"].clone()
                self.denom = c["denom"].clone()
                self.kf_idx = c["kf_idx"].clone()
                self.nr_obs = torch.zeros(P).int()
                self.pruned = None
                self.optimizer = types.SimpleNamespace(step=lambda: None, zero_grad=lambda set_to_none=True: None)

            def prune_points(self, mask):
                self.pruned = mask.clone()

            def update_learning_rate(self, it):
                pass

            def reset_opacity_nonvisible(self, v):
                raise AssertionError("not on this path")

            def densify_and_prune(self, *a):
                raise AssertionError("not on this path")

        model = Model()
        calls = {"i": 0}
        leaf = torch.zeros(1, requires_grad=True)

        def fake_render(*args, **kw):
            i = calls["i"]
            calls["i"] += 1
            vsp = torch.zeros(P, 3, requires_grad=True) + 0
            vsp.retain_grad()
            return {"render": vsp, "viewspace_points": vsp, "visibility_filter": c["radii"][i] > 0,
                    "radii": c["radii"][i], "depth": i, "opacity": None, "n_touched": c["n_touched"][i]}

        def fake_loss(render_image, render_depth, viewpoint, init=False, invert_depth=False):
            # a loss whose gradient w.r.t. the viewspace tensor is exactly the seeded gradient of this keyframe
            return (render_image * c["grads"][render_depth]).sum() + 0 * leaf.sum()

        slam_mapper.render = fake_render
        slam_mapper.get_loss_mapping = fake_loss
        slam_mapper.update_pose = lambda vp: None
        slam_mapper.Log = lambda *a, **k: None
        tensor_cuda = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self           # no GPU in this container: `.cuda()` is the identity
        try:
            m = object.__new__(slam_mapper.Mapper)               # no process / queues: only the loop's own state
            m.viewpoints_dict = {k: types.SimpleNamespace(frame_idx=k) for k in WINDOW}
            m.gaussians = model
            m.cam_intrinsics = m.background = None
            m.nr_iters = 0
            m.window_size = K
            m.gaussian_update_every, m.gaussian_update_offset, m.gaussian_reset = 150, 50, 2001
            m.keyframe_optimizers = types.SimpleNamespace(step=lambda: None, zero_grad=lambda set_to_none=True: None)
            m.first_time_pruned = first_time
            m.occ_aware_visibility_dict = {}
            m.optimize_map(list(WINDOW), prune=prune, iters=1)
        finally:
            torch.Tensor.cuda = tensor_cuda
        if prune:
            out[f"{case}/nr_obs"] = model.nr_obs.numpy()
            out[f"{case}/to_prune"] = model.pruned.numpy()
            out[f"{case}/min_kf"] = np.int32(0 if not first_time else sorted(WINDOW, reverse=True)[2])
            assert m.first_time_pruned
        else:
            out[f"{case}/max_radii_2d"] = model.max_radii_2d.numpy()
            out[f"{case}/xyz_gradient_accum"] = model.xyz_gradient_accum.numpy()
            out[f"{case}/denom"] = model.denom.numpy()
            vis = np.stack([m.occ_aware_visibility_dict[k].numpy() for k in WINDOW])
            out[f"{case}/occ_visibility"] = vis.astype(np.uint8)
        print(case, {k: (v.shape, float(np.asarray(v, dtype=np.float64).sum())) for k, v in out.items()
                     if k.startswith(case)})
    np.savez_compressed(ROOT / "tests" / "golden" / "stats.npz", **out)


if __name__ == "__main__":
    main()
