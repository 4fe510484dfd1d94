"""GPU tests of monogs_b200.render: gsr_camera_setup against golden vectors of the reference's camera code, and the
render() drop-in against a direct GaussianRasterizer call with the reference-style matrices."""
import types

import numpy as np
import pytest
import torch

from conftest import ROOT, rel_inf
from monogs_b200 import render as R
from monogs_b200 import synth
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLD = np.load(ROOT / "tests" / "golden" / "camera.npz")


class _Obj:
    def __init__(self, **kw):
        self.__dict__.update(kw)


def _intr(name):
    fx, fy, cx, cy, H, W = GOLD[f"{name}_intr"]
    return _Obj(fx=torch.tensor([fx], dtype=torch.float32, device=DEV), fy=torch.tensor([fy], dtype=torch.float32, device=DEV),
                cx=float(cx), cy=float(cy), height=int(H), width=int(W))


@pytest.mark.parametrize("name", ["tum", "replica"])
def test_camera_setup_matches_reference_camera_code(built, name):
    proj_T = torch.from_numpy(GOLD[f"{name}_proj_T"]).to(DEV)
    for k in range(3):
        pose = torch.eye(4, device=DEV)
        pose[:3, :3] = torch.from_numpy(GOLD[f"{name}_R{k}"]).to(DEV)
        pose[:3, 3] = torch.from_numpy(GOLD[f"{name}_T{k}"]).to(DEV)
        # strided views of a 4x4 pose, like CameraExtrinsics.__init__ (camera_utils.py:97-99)
        view, proj, campos = R.camera_matrices(pose[:3, :3], pose[:3, 3], proj_T)
        assert np.array_equal(view.cpu().numpy(), GOLD[f"{name}_view{k}"])
        assert rel_inf(proj.cpu().numpy(), GOLD[f"{name}_proj{k}"]) <= 1e-6
        assert np.abs(campos.cpu().numpy() - GOLD[f"{name}_campos{k}"]).max() <= 1e-5 * (1 + np.abs(GOLD[f"{name}_campos{k}"]).max())
        # the torch formulas of the reference evaluated on this device: bmm and general inverse
        ref_proj = view.unsqueeze(0).bmm(proj_T.unsqueeze(0)).squeeze(0)
        assert rel_inf(proj.cpu().numpy(), ref_proj.cpu().numpy()) <= 1e-6
        assert rel_inf(campos.cpu().numpy(), view.inverse()[3, :3].cpu().numpy()) <= 1e-5


def test_render_dropin_equals_direct_rasterizer_call(built):
    cam = synth.make_camera("tum", rot=(0.02, -0.01, 0.015), trans=(0.05, -0.02, 0.03))
    P = 50_000
    sc = synth.make_scene(P, cam, seed=3, device=DEV)
    ci = _intr("tum")
    w2c = cam.w2c.to(DEV)
    vp = _Obj(R=w2c[:3, :3], T=w2c[:3, 3], cam_rot_delta=torch.zeros(3, device=DEV, requires_grad=True),
              cam_trans_delta=torch.zeros(3, device=DEV, requires_grad=True))
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "rotations", "opacities", "colors")}
    iso = sc["scales"][:, :1].clone().requires_grad_(True)      # isotropic map: [P,1] scales (render() repeats them)
    pkg = R.render(vp, ci, leaves["means3D"], leaves["rotations"], iso, leaves["opacities"], leaves["colors"], bg)
    assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii", "depth", "opacity", "n_touched"}
    w = torch.rand(3, ci.height, ci.width, device=DEV)
    (pkg["render"] * w).sum().backward()

    # the reference's per-call path written out: properties + assembly + bmm + inverse, then the rasterizer
    tanfovx, tanfovy, proj_T = R.camera_constants(ci)
    view = torch.zeros(4, 4, device=DEV); view[:3, :3] = vp.R; view[:3, 3] = vp.T; view[3, 3] = 1.0
    view = view.transpose(0, 1)
    s = GaussianRasterizationSettings(image_height=ci.height, image_width=ci.width, tanfovx=tanfovx, tanfovy=tanfovy,
                                      bg=bg, scale_modifier=1.0, viewmatrix=view,
                                      projmatrix=view.unsqueeze(0).bmm(proj_T.unsqueeze(0)).squeeze(0),
                                      projmatrix_raw=proj_T, sh_degree=0, campos=view.inverse()[3, :3],
                                      prefiltered=False, debug=False)
    l2 = {k: sc[k].clone().requires_grad_(True) for k in leaves}
    iso2 = sc["scales"][:, :1].clone().requires_grad_(True)
    m2d = torch.zeros(P, 3, device=DEV, requires_grad=True)
    th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
    img, radii, depth, opa, nt = GaussianRasterizer(s)(means3D=l2["means3D"], means2D=m2d, opacities=l2["opacities"],
                                                       colors_precomp=l2["colors"], scales=iso2.repeat(1, 3),
                                                       rotations=l2["rotations"], theta=th, rho=rh)
    (img * w).sum().backward()
    n = lambda t: t.detach().cpu().numpy()
    assert torch.equal(pkg["radii"], radii) and torch.equal(pkg["n_touched"], nt)
    assert torch.equal(pkg["visibility_filter"], radii > 0)
    assert rel_inf(n(pkg["render"]), n(img)) <= 1e-5 and rel_inf(n(pkg["depth"]), n(depth)) <= 1e-5
    assert rel_inf(n(pkg["opacity"]), n(opa)) <= 1e-5
    for k in leaves:
        assert rel_inf(n(leaves[k].grad), n(l2[k].grad)) <= 1e-4, k
    assert rel_inf(n(iso.grad), n(iso2.grad)) <= 1e-4
    assert rel_inf(n(pkg["viewspace_points"].grad), n(m2d.grad)) <= 1e-4
    assert rel_inf(n(vp.cam_rot_delta.grad), n(th.grad)) <= 1e-4
    assert rel_inf(n(vp.cam_trans_delta.grad), n(rh.grad)) <= 1e-4
