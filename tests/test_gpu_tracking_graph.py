"""monogs_b200.tracking.TrackingGraph: one tracking iteration captured into a CUDA graph must give the results of the
eager fused path (render() + fused_loss.render_l1_loss(kind="tracking") + backward) for every replayed pose:
loss to 1e-6 relative, pose gradient to 1e-5 relative (fp32 atomics order), images bit-identical."""
import numpy as np
import pytest
import torch

from conftest import rel_inf
from monogs_b200 import fused_loss, render as R, synth, tracking
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings

pytestmark = pytest.mark.gpu
DEV = "cuda"


class _Obj:
    def __init__(self, **kw):
        self.__dict__.update(kw)


def _eager(ci, sc, bg, Rm, Tv, gt_rgb, gt_depth, ea, eb):
    tanfovx, tanfovy, proj_T = R.camera_constants(ci)
    view, proj, campos = R.camera_matrices(Rm, Tv, proj_T)
    s = GaussianRasterizationSettings(image_height=ci.height, image_width=ci.width, tanfovx=tanfovx, tanfovy=tanfovy,
                                      bg=bg, scale_modifier=1.0, viewmatrix=view, projmatrix=proj,
                                      projmatrix_raw=proj_T, sh_degree=0, campos=campos, prefiltered=False, debug=False)
    th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
    a, b = ea.clone().requires_grad_(True), eb.clone().requires_grad_(True)
    m2d = torch.zeros_like(sc["means3D"])
    loss, color, radii, depth, opacity, n_touched = fused_loss.render_l1_loss(
        s, sc["means3D"], m2d, sc["opacities"], colors_precomp=sc["colors"], scales=sc["scales"],
        rotations=sc["rotations"], theta=th, rho=rh, gt_rgb=gt_rgb, gt_depth=gt_depth, exposure_a=a, exposure_b=b,
        kind="tracking")
    loss.backward()
    return loss.detach(), rh.grad.reshape(3), th.grad.reshape(3), a.grad, b.grad, color, depth, n_touched


def test_graph_replays_match_eager(built):
    cam = synth.make_camera("tum")
    P = 60_000
    sc = synth.make_scene(P, cam, seed=4, device=DEV)
    fx, fy = 535.4, 539.2
    ci = _Obj(fx=torch.tensor([fx], device=DEV), fy=torch.tensor([fy], device=DEV), cx=320.1, cy=247.6,
              height=cam.H, width=cam.W)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    gt_rgb, gt_depth = synth.make_targets(cam, seed=2, device=DEV)
    ea, eb = torch.tensor([0.03], device=DEV), torch.tensor([-0.01], device=DEV)
    poses = []
    for k in range(4):
        c = synth.make_camera("tum", rot=(0.004 * k, -0.006 * k, 0.003), trans=(0.02 * k, -0.01, 0.015 * k))
        w2c = c.w2c.to(DEV)
        poses.append((w2c[:3, :3].contiguous(), w2c[:3, 3].contiguous()))
    tg = tracking.TrackingGraph(ci, sc["means3D"], sc["rotations"], sc["scales"], sc["opacities"], sc["colors"], bg,
                                gt_rgb, gt_depth, R0=poses[0][0], T0=poses[0][1])
    for Rm, Tv in poses + poses[:1]:
        loss, g_rho, g_theta, terms = tg.step(Rm, Tv, ea, eb)
        got = [t.clone() for t in (loss, g_rho, g_theta, terms, tg.color, tg.depth, tg.n_touched)]
        assert not tg.overflowed()
        ref = _eager(ci, sc, bg, Rm, Tv, gt_rgb, gt_depth, ea, eb)
        assert abs(float(got[0]) - float(ref[0])) <= 1e-6 * abs(float(ref[0]))
        assert rel_inf(got[1].cpu().numpy(), ref[1].cpu().numpy()) <= 1e-5
        assert rel_inf(got[2].cpu().numpy(), ref[2].cpu().numpy()) <= 1e-5
        assert abs(float(got[3][5]) - float(ref[3])) <= 1e-5 * max(abs(float(ref[3])), 1e-8)
        assert abs(float(got[3][6]) - float(ref[4])) <= 1e-5 * max(abs(float(ref[4])), 1e-8)
        assert torch.equal(got[4], ref[5]) and torch.equal(got[5], ref[6]) and torch.equal(got[6], ref[7])


def test_graph_reports_capacity_overflow(built):
    cam = synth.make_camera("tum")
    sc = synth.make_scene(20_000, cam, seed=1, device=DEV)
    ci = _Obj(fx=torch.tensor([535.4], device=DEV), fy=torch.tensor([539.2], device=DEV), cx=320.1, cy=247.6,
              height=cam.H, width=cam.W)
    bg = torch.zeros(3, device=DEV)
    gt_rgb, gt_depth = synth.make_targets(cam, seed=2, device=DEV)
    far = synth.make_camera("tum", trans=(0.0, 0.0, 40.0))      # from 40 m behind: almost nothing on screen
    w2c = far.w2c.to(DEV)
    args = (ci, sc["means3D"], sc["rotations"], sc["scales"], sc["opacities"], sc["colors"], bg, gt_rgb, gt_depth)
    tg = tracking.TrackingGraph(*args, R0=w2c[:3, :3].contiguous(), T0=w2c[:3, 3].contiguous(),
                                capacity_factor=1.0, capacity_slack=0)
    loss, g_rho, g_theta, _ = tg.step(w2c[:3, :3].contiguous(), w2c[:3, 3].contiguous())
    assert not tg.overflowed() and bool(torch.isfinite(loss)) and bool(torch.isfinite(g_rho).all())
    near = cam.w2c.to(DEV)
    Rn, Tn = near[:3, :3].contiguous(), near[:3, 3].contiguous()
    loss, g_rho, g_theta, _ = tg.step(Rn, Tn)
    # the near pose has far more instances than the capacity captured for the far pose holds: the replay itself
    # must have poisoned its loss and pose gradient (device side), and overflowed() must say so
    assert tg.overflowed() and int(tg.total[0]) > tg.capacity
    assert bool(torch.isnan(loss)) and bool(torch.isnan(g_rho).all()) and bool(torch.isnan(g_theta).all())
    tg.recapture()                                               # sized for the near pose now
    loss, g_rho, g_theta, _ = tg.step(Rn, Tn)
    assert not tg.overflowed() and bool(torch.isfinite(loss)) and bool(torch.isfinite(g_theta).all())
    ea, eb = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    ref = _eager(ci, sc, bg, Rn, Tn, gt_rgb, gt_depth, ea, eb)
    assert abs(float(loss) - float(ref[0])) <= 1e-6 * abs(float(ref[0]))
    assert rel_inf(g_rho.cpu().numpy(), ref[1].cpu().numpy()) <= 1e-5
    # map tensors are captured by address: a strided view or another dtype would be frozen as a hidden copy
    with pytest.raises(RuntimeError, match="contiguous float32"):
        tracking.TrackingGraph(ci, sc["means3D"].double(), *args[2:])
    with pytest.raises(RuntimeError, match="contiguous float32"):
        tracking.TrackingGraph(ci, torch.cat([sc["means3D"], sc["means3D"]], 1)[:, :3], *args[2:])
    torch.cuda.synchronize()
