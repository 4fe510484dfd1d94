"""Fused render + L1 loss (monogs_b200.fused_loss), pinned to the REFERENCE's own loss functions
(utils/slam_utils.py:58-146) two ways:
  * golden vectors tests/golden/loss.npz = get_loss_mapping / get_loss_tracking imported from /root/reference and
    differentiated by autograd on CPU (tests/golden/make_loss_golden.py): loss value, dL/dimage, dL/ddepth, exposure
    gradients of the CUDA loss kernels (gsr_l1_loss_forward / gsr_l1_loss_grad, the device code the fused backward
    uses per pixel) -- loss 2e-6 relative, gradient images and exposure gradients 1e-5 relative;
  * on the GPU box, the unmodified reference file itself (oracle/_ref/callers/utils/slam_utils.py) evaluated by torch
    autograd on top of the plain rasterizer node against the fused node: every gradient within 1e-4 relative.
"""
import types

import numpy as np
import pytest
import torch

import ref_callers
from conftest import ROOT, rel_inf
from monogs_b200 import synth
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from monogs_b200.fused_loss import render_l1_loss

pytestmark = pytest.mark.gpu
DEV = "cuda"


def ref_loss_mapping(image, depth, gt_rgb, gt_depth, gt_mask, a, b, lambda_depth=0.9):
    """get_loss_mapping, utils/slam_utils.py:101-146 (init=False, invert_depth=False)."""
    rgb = torch.exp(a) * image + b
    rgb = rgb.permute(1, 2, 0)
    gt = gt_rgb.permute(1, 2, 0)
    l1_rgb = torch.abs(rgb[gt_mask] - gt[gt_mask]).mean() if gt_mask is not None else torch.abs(rgb - gt).mean()
    dm = gt_depth > 0
    l1_depth = torch.abs(depth[dm] - gt_depth[dm]).mean()
    return lambda_depth * l1_rgb + (1 - lambda_depth) * l1_depth


def ref_loss_tracking(image, depth, opacity, gt_rgb, gt_depth, gt_mask, grad_mask, a, b):
    """get_loss_tracking, utils/slam_utils.py:58-98 (invert_depth=False)."""
    opacity_mask = opacity > 0.99
    rgb = torch.exp(a) * image + b
    rgb_mask = gt_mask * grad_mask * opacity_mask
    l1_rgb = (opacity * torch.abs(rgb * rgb_mask - gt_rgb * rgb_mask).mean()).mean()
    depth_mask = (gt_depth > 0) * opacity_mask
    l1_depth = torch.abs(depth[depth_mask] - gt_depth[depth_mask]).mean() if depth_mask.any() else 0.0
    return 0.5 * l1_rgb + l1_depth


def _reference_losses():
    """(get_loss_mapping, get_loss_tracking) of the reference's unmodified utils/slam_utils.py when it was shipped
    (oracle/_ref/callers), as functions of explicit tensors; else the formulas restated above."""
    if not ref_callers.available():
        return ref_loss_mapping, ref_loss_tracking
    ref_callers.verify_unmodified()
    S = ref_callers.load_stack("ours")

    def vp(gt_rgb, gt_depth, gt_mask, grad_mask, a, b):
        return types.SimpleNamespace(rgb=gt_rgb, depth=gt_depth[0], mask=gt_mask, grad_mask=grad_mask, exposure_a=a,
                                     exposure_b=b)

    def mapping(image, depth, gt_rgb, gt_depth, gt_mask, a, b):
        return S.get_loss_mapping(image, depth, vp(gt_rgb, gt_depth, gt_mask, None, a, b), init=False)

    def tracking(image, depth, opacity, gt_rgb, gt_depth, gt_mask, grad_mask, a, b):
        return S.get_loss_tracking(image, depth, opacity, vp(gt_rgb, gt_depth, gt_mask, grad_mask, a, b))

    return mapping, tracking


def _setup(P=12_000, seed=3):
    cam = synth.make_camera("tum_half", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=seed, device=DEV)
    sc["opacities"] = (sc["opacities"] * 0.5 + 0.5).contiguous()      # dense enough for opacity > 0.99 regions
    view, proj, proj_raw, campos = cam.matrices(DEV)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    s = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, proj, proj_raw, 0, campos,
                                      False, False)
    gt_rgb, gt_depth = [t.to(DEV) for t in synth.make_targets(cam, seed=seed)]
    g = torch.Generator().manual_seed(seed)
    mask = (torch.rand(cam.H, cam.W, generator=g) > 0.2).to(DEV)
    grad_mask = (torch.rand(1, cam.H, cam.W, generator=g) > 0.5).to(DEV)
    return cam, sc, s, gt_rgb, gt_depth, mask, grad_mask


def _leaves(sc):
    L = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "colors", "opacities", "scales", "rotations")}
    L["means2D"] = torch.zeros_like(sc["means3D"], requires_grad=True)
    L["theta"] = torch.zeros(3, device=DEV, requires_grad=True)
    L["rho"] = torch.zeros(3, device=DEV, requires_grad=True)
    L["a"] = torch.tensor([0.05], device=DEV, requires_grad=True)
    L["b"] = torch.tensor([-0.02], device=DEV, requires_grad=True)
    return L


@pytest.mark.parametrize("kind", ["mapping", "tracking"])
def test_fused_loss_matches_autograd(built, kind):
    cam, sc, s, gt_rgb, gt_depth, mask, grad_mask = _setup()
    A, B = _leaves(sc), _leaves(sc)
    # reference: plain rasterizer node + the callers' loss in torch
    color, radii, depth, opacity, nt = GaussianRasterizer(s)(
        means3D=A["means3D"], means2D=A["means2D"], colors_precomp=A["colors"], opacities=A["opacities"],
        scales=A["scales"], rotations=A["rotations"], theta=A["theta"], rho=A["rho"])
    loss_mapping, loss_tracking = _reference_losses()
    if kind == "mapping":
        loss_ref = loss_mapping(color, depth, gt_rgb, gt_depth, mask, A["a"], A["b"])
    else:
        loss_ref = loss_tracking(color, depth, opacity, gt_rgb, gt_depth, mask, grad_mask, A["a"], A["b"])
        assert int((opacity > 0.99).sum()) > 1000   # the opacity mask must actually select something
    (3.0 * loss_ref).backward()                      # upstream gradient != 1
    # fused
    loss, color2, radii2, depth2, opacity2, nt2, terms = render_l1_loss(
        s, B["means3D"], B["means2D"], B["opacities"], colors_precomp=B["colors"], scales=B["scales"],
        rotations=B["rotations"], theta=B["theta"], rho=B["rho"], gt_rgb=gt_rgb, gt_depth=gt_depth, mask=mask,
        grad_mask=grad_mask if kind == "tracking" else None, exposure_a=B["a"], exposure_b=B["b"], kind=kind,
        return_terms=True)
    (3.0 * loss).backward()
    assert torch.equal(color2, color.detach()) and torch.equal(radii2, radii) and torch.equal(nt2, nt)
    assert abs(float(loss) - float(loss_ref)) <= 2e-6 * abs(float(loss_ref))
    n = lambda t: t.detach().cpu().numpy()
    for k in ("means3D", "colors", "opacities", "scales", "rotations", "means2D", "theta", "rho", "a", "b"):
        assert B[k].grad is not None, k
        assert rel_inf(n(B[k].grad), n(A[k].grad)) <= 1e-4, (k, rel_inf(n(B[k].grad), n(A[k].grad)))


def test_fused_loss_no_masks_no_depth(built):
    cam, sc, s, gt_rgb, gt_depth, mask, grad_mask = _setup(P=5000, seed=9)
    A, B = _leaves(sc), _leaves(sc)
    color, radii, depth, opacity, nt = GaussianRasterizer(s)(
        means3D=A["means3D"], means2D=A["means2D"], colors_precomp=A["colors"], opacities=A["opacities"],
        scales=A["scales"], rotations=A["rotations"], theta=A["theta"], rho=A["rho"])
    (0.9 * torch.abs(color - gt_rgb).mean()).backward()
    loss = render_l1_loss(s, B["means3D"], B["means2D"], B["opacities"], colors_precomp=B["colors"],
                          scales=B["scales"], rotations=B["rotations"], theta=B["theta"], rho=B["rho"], gt_rgb=gt_rgb)[0]
    loss.backward()
    n = lambda t: t.detach().cpu().numpy()
    for k in ("means3D", "colors", "opacities", "theta", "rho"):
        assert rel_inf(n(B[k].grad), n(A[k].grad)) <= 1e-4, k
    with pytest.raises(ValueError):
        render_l1_loss(s, B["means3D"], B["means2D"], B["opacities"], colors_precomp=B["colors"], scales=B["scales"],
                       rotations=B["rotations"], gt_rgb=gt_rgb, kind="nope")


@pytest.fixture(scope="module")
def loss_golden():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_loss_golden", ROOT / "tests" / "golden" / "make_loss_golden.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, np.load(ROOT / "tests" / "golden" / "loss.npz")


@pytest.mark.parametrize("name", ["mapping", "mapping_init", "tracking", "tracking_sparse"])
def test_loss_kernels_match_reference_golden(built, loss_golden, name):
    """gsr_l1_loss_forward + gsr_l1_loss_grad against the reference functions' own outputs (see module docstring)."""
    from monogs_b200.fused_loss import l1_loss_images
    mod, g = loss_golden
    c = mod.case_inputs(name)
    d = lambda t: t.to(DEV)
    no_exposure = c["kind"] == "mapping" and c["init"]
    terms, d_color, d_depth = l1_loss_images(
        c["kind"], d(c["image"]), d(c["depth"]), d(c["opacity"]), d(c["gt_rgb"]), d(c["gt_depth"])[None],
        mask=d(c["mask"]), grad_mask=d(c["grad_mask"]) if c["kind"] == "tracking" else None,
        exposure_a=None if no_exposure else d(c["exposure_a"]), exposure_b=None if no_exposure else d(c["exposure_b"]),
        lambda_rgb=0.9, upstream=torch.tensor([c["upstream"]], device=DEV))
    terms = terms.cpu().numpy().astype(np.float64)
    want = float(g[f"{name}/loss"])
    assert abs(terms[0] - want) <= 2e-6 * abs(want), (terms[0], want)          # tolerance: 2e-6 relative
    assert rel_inf(d_color.cpu().numpy(), g[f"{name}/d_image"]) <= 1e-5          # tolerance: 1e-5 relative
    assert rel_inf(d_depth.cpu().numpy(), g[f"{name}/d_depth"]) <= 1e-5
    # where the reference's gradient image is exactly zero (masked-out pixels) ours is too
    assert ((d_color.cpu().numpy() == 0) == (g[f"{name}/d_image"] == 0)).all()
    assert ((d_depth.cpu().numpy() == 0) == (g[f"{name}/d_depth"] == 0)).all()
    if not no_exposure:
        up = c["upstream"]
        assert abs(up * terms[5] - float(g[f"{name}/d_a"][0])) <= 1e-5 * abs(float(g[f"{name}/d_a"][0]))
        assert abs(up * terms[6] - float(g[f"{name}/d_b"][0])) <= 1e-5 * abs(float(g[f"{name}/d_b"][0]))
