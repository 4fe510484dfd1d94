"""Row b2 of the scope table: the reference's own callers run UNCHANGED on this package.

The files under oracle/_ref/callers/ are byte-for-byte copies of /root/reference's
  gaussian_splatting/gaussian_renderer/__init__.py:26-168   render()
  utils/camera_utils.py                                     CameraExtrinsics / CameraIntrinsics / get_full_proj_transform
  utils/pose_utils.py:76-93                                 update_pose
  utils/slam_utils.py:58-146                                get_loss_tracking / get_loss_mapping
(sha256 re-checked here).  Each scenario below is executed twice with the same inputs -- once with
`diff_gaussian_rasterization` resolving to this package (monogs_b200.install_aliases()), once with it resolving to the
reference's own Python wrapper over the unmodified reference CUDA kernels (oracle/_ref) -- and everything a caller
can observe is compared: the returned dict, the .grad of every leaf incl. cam_rot_delta / cam_trans_delta /
exposure, and the pose after an optimisation loop.
  (i)   one render() + get_loss_mapping + backward
  (ii)  the tracking loop of utils/slam_tracker.py:132-188 (Adam on the pose deltas + exposure, update_pose)
  (iii) the window loop of utils/slam_mapper.py:273-322,394 (summed loss over 3 keyframes, ONE backward)
Bars: integers exact; images 1e-5 relative; gradients 1e-4 relative.
"""
import numpy as np
import pytest
import torch

import ref_callers
from conftest import rel_inf
from monogs_b200 import synth
from oracle import ref_cuda

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not (ref_callers.available() and ref_cuda.available()),
                                 reason="oracle/_ref (reference callers + reference CUDA library) not shipped")]
DEV = "cuda:0"
KIND = "tum_half"


def n(t):
    return t.detach().cpu().numpy()


class Map:
    """The Gaussian tensors render() receives (`gaussians.get_xyz` ... in the callers): activations already applied.
    isotropic=True: scale [P,1] like MonoGS's default config (render() repeats it to 3 columns)."""

    def __init__(self, P, cam, seed, isotropic=True):
        sc = synth.make_scene(P, cam, seed=seed, anisotropic=not isotropic, device=DEV)
        scale = sc["scales"][:, :1].contiguous() if isotropic else sc["scales"]
        # opacities lifted to [0.5, 1): dense enough for get_loss_tracking's `opacity > 0.99` mask to select pixels
        self.leaves = dict(xyz=sc["means3D"], rot=sc["rotations"], scale=scale,
                           opacity=(sc["opacities"] * 0.5 + 0.5).contiguous(), rgb=sc["colors"])
        for k in self.leaves:
            self.leaves[k] = self.leaves[k].clone().requires_grad_(True)

    def args(self):
        L = self.leaves
        return L["xyz"], L["rot"], L["scale"], L["opacity"], L["rgb"]


def make_viewpoint(S, idx, cam, rgb, depth, seed):
    """A CameraExtrinsics exactly as the dataset path builds it (camera_utils.py:128-156), pose set by update_RT."""
    g = torch.Generator().manual_seed(seed)
    mask = (torch.rand(cam.H, cam.W, generator=g) > 0.1).to(DEV)
    vp = S.CameraExtrinsics(idx, device=DEV, rgb=rgb, depth=depth, mask=mask)
    vp.update_RT(cam.w2c[:3, :3].clone(), cam.w2c[:3, 3].clone())
    vp.grad_mask = (torch.rand(1, cam.H, cam.W, generator=g) > 0.4).to(DEV)
    with torch.no_grad():
        vp.exposure_a.fill_(0.03)
        vp.exposure_b.fill_(-0.01)
    return vp


def make_intrinsics(S):
    W, H, fx, fy, cx, cy = synth.INTRINSICS[KIND]
    return S.CameraIntrinsics(fx, fy, cx, cy, H, W, device=DEV)


def ground_truth(cam_true, P, seed):
    """Target image / depth: this repo's renderer at the true pose (the same tensors are fed to both backends)."""
    from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    sc = synth.make_scene(P, cam_true, seed=seed, device=DEV)
    view, proj, proj_raw, campos = cam_true.matrices(DEV)
    s = GaussianRasterizationSettings(cam_true.H, cam_true.W, cam_true.tanfovx, cam_true.tanfovy,
                                      torch.zeros(3, device=DEV), 1.0, view, proj, proj_raw, 0, campos, False, False)
    with torch.no_grad():
        col, _, dep, opa, _ = GaussianRasterizer(s)(
            means3D=sc["means3D"], means2D=torch.zeros_like(sc["means3D"]), colors_precomp=sc["colors"],
            opacities=(sc["opacities"] * 0.5 + 0.5).contiguous(), scales=sc["scales"][:, :1].repeat(1, 3),
            rotations=sc["rotations"])
    g = torch.Generator().manual_seed(seed + 1)
    # the rasterizer's depth is the un-normalised sum of depth * alpha * T and the callers compare it to gt as is
    depth = dep[0] * (torch.rand(cam_true.H, cam_true.W, generator=g) > 0.1).to(DEV)
    return col.clamp(0, 1).contiguous(), depth.contiguous()


def compare_pkg(a, b):
    assert set(a) == set(b) == {"render", "viewspace_points", "visibility_filter", "radii", "depth", "opacity",
                                "n_touched"}
    assert torch.equal(a["radii"], b["radii"]) and torch.equal(a["n_touched"], b["n_touched"])
    assert torch.equal(a["visibility_filter"], b["visibility_filter"])
    for k in ("render", "depth", "opacity"):
        assert a[k].shape == b[k].shape
        assert rel_inf(n(a[k]), n(b[k])) <= 1e-5, k          # tolerance: 1e-5 relative


def compare_grads(A, B, names, tol=1e-4):
    for k in names:
        ga, gb = A[k].grad, B[k].grad
        assert ga is not None and gb is not None, k
        assert ga.shape == gb.shape, k
        r = rel_inf(n(ga), n(gb))
        assert r <= tol, (k, r)                                # tolerance: 1e-4 relative


def test_shipped_caller_files_are_the_reference_files():
    assert ref_callers.verify_unmodified() == 7


def test_render_and_mapping_loss_unchanged(built):
    P = 30_000
    cam_true = synth.make_camera(KIND)
    cam = synth.make_camera(KIND, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    rgb, depth = ground_truth(cam_true, P, seed=11)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    out = {}
    for backend in ("ours", "ref"):
        S = ref_callers.load_stack(backend)
        m = Map(P, cam_true, seed=11, isotropic=False)   # anisotropic: the rotation gradient is non-trivial
        vp = make_viewpoint(S, 3, cam, rgb, depth, seed=5)
        pkg = S.render(vp, make_intrinsics(S), *m.args(), bg)
        loss = S.get_loss_mapping(pkg["render"], pkg["depth"], vp, init=False)
        loss.backward()
        leaves = dict(m.leaves, theta=vp.cam_rot_delta, rho=vp.cam_trans_delta, a=vp.exposure_a, b=vp.exposure_b,
                      viewspace=pkg["viewspace_points"])
        out[backend] = (pkg, float(loss), leaves)
    compare_pkg(out["ours"][0], out["ref"][0])
    assert abs(out["ours"][1] - out["ref"][1]) <= 1e-5 * abs(out["ref"][1])
    compare_grads(out["ours"][2], out["ref"][2], ["xyz", "rot", "scale", "opacity", "rgb", "theta", "rho", "a", "b",
                                                  "viewspace"])
    # the densification statistic the mapper reads from viewspace_points.grad (gaussian_model.py:888-892)
    vis = out["ref"][0]["visibility_filter"]
    assert int(vis.sum()) > P // 2 and float(out["ours"][2]["viewspace"].grad[vis, :2].norm(dim=-1).max()) > 0


def test_tracking_loop_unchanged(built):
    """utils/slam_tracker.py:102-188: Adam over cam_rot_delta / cam_trans_delta / exposure_a / exposure_b, render,
    get_loss_tracking, backward, step, update_pose -- 10 iterations from a perturbed pose."""
    P = 30_000
    cam_true = synth.make_camera(KIND)
    cam0 = synth.make_camera(KIND, rot=(0.006, -0.008, 0.004), trans=(0.03, -0.015, 0.02))
    rgb, depth = ground_truth(cam_true, P, seed=21)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    out = {}
    for backend in ("ours", "ref"):
        S = ref_callers.load_stack(backend)
        m = Map(P, cam_true, seed=21)
        for t in m.leaves.values():
            t.requires_grad_(False)                       # tracking runs against a fixed map
        vp = make_viewpoint(S, 7, cam0, rgb, depth, seed=6)
        with torch.no_grad():
            vp.exposure_a.zero_()
            vp.exposure_b.zero_()
        intr = make_intrinsics(S)
        opt = torch.optim.Adam([                          # learning rates of configs/mono/tum/base_config.yaml
            {"params": [vp.cam_rot_delta], "lr": 0.003}, {"params": [vp.cam_trans_delta], "lr": 0.001},
            {"params": [vp.exposure_a], "lr": 0.01}, {"params": [vp.exposure_b], "lr": 0.01}])
        losses, first_grads = [], None
        for it in range(10):
            opt.zero_grad()
            pkg = S.render(vp, intr, *m.args(), bg)
            loss = S.get_loss_tracking(pkg["render"], pkg["depth"], pkg["opacity"], vp, invert_depth=False)
            loss.backward()
            if first_grads is None:
                first_grads = {k: getattr(vp, k).grad.clone() for k in
                               ("cam_rot_delta", "cam_trans_delta", "exposure_a", "exposure_b")}
            with torch.no_grad():
                opt.step()
                S.update_pose(vp, converged_threshold=1e-4)
            losses.append(float(loss))
        out[backend] = (losses, vp.R.clone(), vp.T.clone(), first_grads, float(vp.exposure_a), float(vp.exposure_b))
    lo, Ro, To, go, ao, bo = out["ours"]
    lr, Rr, Tr, gr, ar, br = out["ref"]
    for k in gr:
        assert rel_inf(n(go[k]), n(gr[k])) <= 1e-4, (k, n(go[k]), n(gr[k]))   # pose / exposure gradient, first iteration
    assert lr[-1] < lr[0] and lo[-1] < lo[0]                                  # the loop does optimise the pose
    assert max(abs(a - b) / abs(b) for a, b in zip(lo, lr)) <= 1e-3, (lo, lr)
    # Adam turns gradients into lr-sized steps (3e-3 rad / 1e-3 m per iteration): after 10 iterations the two
    # trajectories must agree to a small fraction of ONE step
    assert float((Ro - Rr).abs().max()) <= 2e-4, float((Ro - Rr).abs().max())
    assert float((To - Tr).abs().max()) <= 2e-4, float((To - Tr).abs().max())
    assert abs(ao - ar) <= 1e-3 and abs(bo - br) <= 1e-3
    moved = float((To - cam0.w2c[:3, 3].to(DEV)).abs().max())
    assert moved > 2e-3                                                        # and the pose did move


def test_mapping_window_loop_unchanged(built):
    """utils/slam_mapper.py:273-322,394: loss_mapping = sum over the window's keyframes of get_loss_mapping(render(kf)),
    ONE backward; Gaussian gradients are the sums over keyframes, pose / exposure gradients stay per keyframe."""
    P = 40_000
    cams = synth.arc_cameras(KIND, 3)
    targets = [ground_truth(c, P, seed=31) for c in cams]
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    out = {}
    for backend in ("ours", "ref"):
        S = ref_callers.load_stack(backend)
        m = Map(P, cams[1], seed=31)
        intr = make_intrinsics(S)
        # keyframe poses a little off the poses the targets were rendered from -> non-trivial pose gradients
        vps = []
        for i, (c, (rgb, depth)) in enumerate(zip(cams, targets)):
            off = synth.make_camera(KIND, rot=(0.002 * (i + 1), -0.003, 0.001), trans=(0.01, -0.005 * i, 0.008))
            c2 = synth.Camera(c.W, c.H, c.fx, c.fy, c.cx, c.cy, off.w2c @ c.w2c)
            vps.append(make_viewpoint(S, i, c2, rgb, depth, seed=40 + i))
        loss_mapping = 0
        pkgs = []
        for vp in vps:
            pkg = S.render(vp, intr, *m.args(), bg)
            loss_mapping += S.get_loss_mapping(pkg["render"], pkg["depth"], vp, init=False, invert_depth=False)
            pkgs.append(pkg)
        loss_mapping.backward()
        out[backend] = (pkgs, float(loss_mapping), m.leaves, vps)
    for a, b in zip(out["ours"][0], out["ref"][0]):
        compare_pkg(a, b)
    assert abs(out["ours"][1] - out["ref"][1]) <= 1e-5 * abs(out["ref"][1])
    compare_grads(out["ours"][2], out["ref"][2], ["xyz", "rot", "scale", "opacity", "rgb"])
    for va, vb, pa, pb in zip(out["ours"][3], out["ref"][3], out["ours"][0], out["ref"][0]):
        for k in ("cam_rot_delta", "cam_trans_delta", "exposure_a", "exposure_b"):
            assert rel_inf(n(getattr(va, k).grad), n(getattr(vb, k).grad)) <= 1e-4, k
        assert rel_inf(n(pa["viewspace_points"].grad), n(pb["viewspace_points"].grad)) <= 1e-4
    # occlusion-aware visibility the mapper derives (slam_mapper.py:399-404)
    for pa, pb in zip(out["ours"][0], out["ref"][0]):
        assert torch.equal((pa["n_touched"] > 0).long(), (pb["n_touched"] > 0).long())
