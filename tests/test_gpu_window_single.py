"""MappingWindow on ONE GPU (world size 1; the multi-GPU exchange is covered by tests/test_window_exchange.py):
the window step -- local keyframes on two streams, fused mapping loss, per-Gaussian backward accumulated into the
64-byte window rows -- must equal the sums that `loss.backward()` forms over the window's keyframes through the
reference-facing autograd API (utils/slam_mapper.py:273-322,394).  Tolerance 1e-5 relative per tensor (fp32 atomics)."""
import numpy as np
import pytest
import torch

from conftest import rel_inf
from monogs_b200 import synth, window
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("fused,streams", [(True, 2), (False, 1)])
def test_window_step_equals_summed_autograd(built, fused, streams):
    K, P = 3, 30_000
    cams = synth.arc_cameras("tum", K)
    sc = {k: v.to(DEV) for k, v in synth.make_scene(P, synth.make_camera("tum"), seed=2).items()}
    targets = [tuple(t.to(DEV) for t in synth.make_targets(c, seed=i)) for i, c in enumerate(cams)]
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    w = window.MappingWindow(window.make_cuda_render_fn(cams, targets, bg, fused_loss=fused), n_streams=streams)
    res = w.step(sc, K)
    torch.cuda.synchronize()

    leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "colors", "opacities", "scales", "rotations")}
    total, g2, vis, rad = 0.0, torch.zeros(P, device=DEV), torch.zeros(P, device=DEV), torch.zeros(P, device=DEV)
    taus = []
    for k, cam in enumerate(cams):
        view, proj, proj_raw, campos = cam.matrices(DEV)
        s = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, proj, proj_raw, 0,
                                          campos, False, False)
        m2d = torch.zeros(P, 3, device=DEV, requires_grad=True)
        th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
        color, radii, depth, opacity, nt = GaussianRasterizer(s)(
            means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], colors_precomp=leaves["colors"],
            scales=leaves["scales"], rotations=leaves["rotations"], theta=th, rho=rh)
        loss, _, _ = window.mapping_loss_grad(color, depth, *targets[k])
        loss.backward()
        total += float(loss)
        g2 += m2d.grad[:, :2].norm(dim=1) * (radii > 0)
        vis += (radii > 0).float()
        rad = torch.maximum(rad, radii.float())
        taus.append(torch.cat([rh.grad.reshape(3), th.grad.reshape(3)]))
    n = lambda t: t.detach().cpu().numpy()
    for name in ("means3D", "colors", "opacities", "scales", "rotations"):
        assert rel_inf(n(res.grads[name]).reshape(n(leaves[name].grad).shape), n(leaves[name].grad)) <= 1e-5, name
    assert rel_inf(n(res.grad2d_norm), n(g2)) <= 1e-5
    assert torch.equal(res.vis_count, vis) and torch.equal(res.max_radii.float(), rad)
    assert abs(float(sum(r.loss for r in res.local)) - total) <= 1e-6 * abs(total)
    for r, t in zip(res.local, taus):
        assert rel_inf(n(r.grad_tau), n(t)) <= 1e-5
