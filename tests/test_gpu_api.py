"""GPU tests of the reference-facing Python surface: GaussianRasterizer autograd contract, the unchanged
MonoGS render() call shape, distCUDA2, and the stable radix-sort hook (cub::DeviceRadixSort::SortPairs)."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import ROOT, rel_inf
from monogs_b200 import _abi, synth
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from monogs_b200.simple_knn._C import distCUDA2
from oracle import cpu_oracle as co
from oracle import ref_cuda as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"
synth.INTRINSICS.setdefault("tiny", (160, 120, 133.85, 134.8, 80.03, 61.9))


def _settings(cam, bg):
    view, proj, proj_raw, campos = cam.matrices(DEV)
    return GaussianRasterizationSettings(image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx,
                                         tanfovy=cam.tanfovy, bg=bg, scale_modifier=1.0, viewmatrix=view,
                                         projmatrix=proj, projmatrix_raw=proj_raw, sh_degree=0, campos=campos,
                                         prefiltered=False, debug=False)


def test_autograd_contract_like_render(built):
    """Same call as gaussian_renderer/__init__.py:145-156 (keywords), grads for every leaf incl. theta/rho."""
    cam = synth.make_camera("tiny", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(4000, cam, seed=8, device=DEV)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "colors", "opacities", "rotations")}
    scaling = sc["scales"][:, :1].clone().requires_grad_(True)       # isotropic: repeat like render() :101-102
    scales = scaling.repeat(1, 3)
    screenspace = torch.zeros_like(leaves["means3D"], requires_grad=True) + 0
    screenspace.retain_grad()
    theta = torch.nn.Parameter(torch.zeros(3, device=DEV))           # camera_utils.py:114-119
    rho = torch.nn.Parameter(torch.zeros(3, device=DEV))
    out = GaussianRasterizer(raster_settings=_settings(cam, bg))(
        means3D=leaves["means3D"], means2D=screenspace, shs=None, colors_precomp=leaves["colors"],
        opacities=leaves["opacities"], scales=scales, rotations=leaves["rotations"], cov3D_precomp=None,
        theta=theta, rho=rho)
    assert len(out) == 5
    color, radii, depth, opacity, n_touched = out
    assert color.shape == (3, cam.H, cam.W) and depth.shape == (1, cam.H, cam.W) and opacity.shape == (1, cam.H, cam.W)
    assert radii.dtype == torch.int32 and n_touched.dtype == torch.int32 and radii.shape == (4000,)
    assert not radii.requires_grad and not n_touched.requires_grad
    g = torch.Generator().manual_seed(1)
    wc = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    wd = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    # opacity participates in the loss but its gradient is dropped, exactly like the reference (:116-138)
    ((color * wc).sum() + (depth * wd).sum() + 0.0 * opacity.sum()).backward()
    n = lambda t: t.detach().cpu().numpy()
    view, proj, proj_raw, campos = cam.matrices()
    f = co.forward(n(sc["means3D"]), n(sc["scales"]), n(sc["rotations"]), n(sc["opacities"]), n(sc["colors"]),
                   n(view), n(proj), n(campos), cam.W, cam.H, cam.tanfovx, cam.tanfovy, n(bg))
    b = co.backward(f, n(sc["means3D"]), n(sc["scales"]), n(sc["rotations"]), n(view), n(proj), n(proj_raw), cam.W,
                    cam.H, cam.tanfovx, cam.tanfovy, n(bg), n(wc), n(wd))
    assert rel_inf(n(leaves["means3D"].grad), b["means3D"]) <= 1e-4
    assert rel_inf(n(leaves["colors"].grad), b["colors"]) <= 1e-4
    assert rel_inf(n(leaves["opacities"].grad), b["opacity"]) <= 1e-4
    assert rel_inf(n(leaves["rotations"].grad), b["rotations"]) <= 1e-4
    assert rel_inf(n(scaling.grad), b["scales"].sum(1, keepdims=True)) <= 1e-4
    assert rel_inf(n(screenspace.grad), b["means2D"]) <= 1e-4 and screenspace.grad.shape == (4000, 3)
    assert theta.grad.shape == (3,) and rho.grad.shape == (3,)
    assert rel_inf(np.concatenate([n(rho.grad), n(theta.grad)]), b["tau_sum"]) <= 1e-4


def test_tracking_shape_only_pose_requires_grad(built):
    """Tracking (slam_tracker.py:138-170): no Gaussian input requires grad; only theta/rho (+ means2D leaf)."""
    cam = synth.make_camera("tiny")
    sc = synth.make_scene(3000, cam, seed=4, device=DEV)
    bg = torch.zeros(3, device=DEV)
    theta = torch.zeros(3, device=DEV, requires_grad=True)
    rho = torch.zeros(3, device=DEV, requires_grad=True)
    m2 = torch.zeros_like(sc["means3D"], requires_grad=True)
    color, radii, depth, opacity, nt = GaussianRasterizer(_settings(cam, bg))(
        means3D=sc["means3D"], means2D=m2, colors_precomp=sc["colors"], opacities=sc["opacities"],
        scales=sc["scales"], rotations=sc["rotations"], theta=theta, rho=rho)
    color.abs().mean().backward()
    assert theta.grad is not None and rho.grad is not None and float(theta.grad.abs().sum()) > 0
    # finite-difference check of the translational pose gradient (d/d rho_x of the loss), central difference
    def loss_at(dx):
        from monogs_b200.synth import Camera
        w2c = cam.w2c.clone(); w2c[0, 3] += dx
        c2 = Camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, w2c)
        col = GaussianRasterizer(_settings(c2, bg))(
            means3D=sc["means3D"], means2D=torch.zeros_like(sc["means3D"]), colors_precomp=sc["colors"],
            opacities=sc["opacities"], scales=sc["scales"], rotations=sc["rotations"])[0]
        return float(col.double().abs().mean())
    h = 2e-3
    fd = (loss_at(h) - loss_at(-h)) / (2 * h)
    assert abs(fd - float(rho.grad[0])) <= 0.05 * abs(fd) + 1e-5, (fd, float(rho.grad[0]))


def test_argument_validation(built):
    cam = synth.make_camera("tiny")
    sc = synth.make_scene(10, cam, seed=0, device=DEV)
    r = GaussianRasterizer(_settings(cam, torch.zeros(3, device=DEV)))
    m2 = torch.zeros_like(sc["means3D"])
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=sc["means3D"], means2D=m2, opacities=sc["opacities"], scales=sc["scales"], rotations=sc["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=sc["means3D"], means2D=m2, opacities=sc["opacities"], colors_precomp=sc["colors"], scales=sc["scales"])


def test_markvisible_module(built):
    cam = synth.make_camera("tiny")
    sc = synth.make_scene(1000, cam, seed=0, frac_behind=0.4, device=DEV)
    m = GaussianRasterizer(_settings(cam, torch.zeros(3, device=DEV))).markVisible(sc["means3D"])
    assert m.dtype == torch.bool and 0 < int(m.sum()) < 1000


# ---------------------------------------------------------------------------------------------------------------
def test_knn_golden_and_oracle(built, golden_mod):
    pts = golden_mod.knn_inputs().to(DEV)
    d = distCUDA2(pts).cpu().numpy()
    g = np.load(ROOT / "tests" / "golden" / "knn.npz")["dist2"]
    assert (d.view(np.uint32) == g.view(np.uint32)).all()
    for P, seed in ((1, 0), (2, 1), (3, 2), (4, 3), (33, 4), (1025, 5), (20_000, 6)):
        gen = torch.Generator().manual_seed(seed)
        p = torch.rand(P, 3, generator=gen) * torch.tensor([4.0, 2.0, 1.0])
        a = distCUDA2(p.to(DEV)).cpu().numpy()
        b = co.knn(p.numpy())
        assert (a.view(np.uint32) == b.view(np.uint32)).all(), P


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libref_dgr.so not shipped")
def test_knn_full_size_against_reference(built):
    gen = torch.Generator().manual_seed(3)
    P = 1_000_000                                                  # BASELINE config C5
    z = 0.5 + 5.5 * torch.rand(P, generator=gen)
    p = torch.stack([(torch.rand(P, generator=gen) - 0.5) * z, (torch.rand(P, generator=gen) - 0.5) * z, z], 1).to(DEV)
    a = distCUDA2(p)
    b = ref.distCUDA2(p)
    assert torch.equal(a.view(torch.int32), b.view(torch.int32))


def test_sort_pairs_u64_stable(built):
    lib = _abi.load()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    gen = torch.Generator().manual_seed(0)
    for n, end_bit in ((1, 41), (255, 43), (4096, 44), (4097, 45), (1_000_003, 43), (5_000_000, 45)):
        hi = torch.randint(0, 1 << (end_bit - 32), (n,), generator=gen, dtype=torch.int64)
        lo = torch.randint(0, 1 << 20, (n,), generator=gen, dtype=torch.int64) << 12  # many depth ties
        junk = torch.randint(0, 1 << 10, (n,), generator=gen, dtype=torch.int64) << end_bit  # bits above end_bit
        keys = ((hi << 32) | lo | junk).to(DEV)
        vals = torch.arange(n, dtype=torch.int32, device=DEV)
        ko, vo = torch.empty_like(keys), torch.empty_like(vals)
        ws = torch.empty(lib.gsr_binning_bytes(n), dtype=torch.uint8, device=DEV)
        rc = lib.gsr_sort_pairs_u64(st, n, keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), end_bit,
                                    ws.data_ptr())
        assert rc == 0
        mask = (1 << end_bit) - 1
        order = torch.sort(keys & mask, stable=True)[1]
        assert torch.equal(vo.long(), order) and torch.equal(ko, keys[order]), (n, end_bit)


def test_profile_hook_times_selected_stages_only(built):
    """gsr_profile_enable(1): every stage of a forward + backward is bracketed; gsr_profile_enable(1 | mask << 8): only
    the stages of the mask (bench.py brackets just the dominant stage inside its timed region)."""
    lib = _abi.load()
    cam = synth.make_camera("tum_half")
    sc = synth.make_scene(20_000, cam, seed=3, device=DEV)
    bg = torch.zeros(3, device=DEV)

    def step():
        leaves = [sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors", "scales", "rotations")]
        m3, op, col, s, r = leaves
        m2 = torch.zeros_like(m3, requires_grad=True)
        out = GaussianRasterizer(_settings(cam, bg))(means3D=m3, means2D=m2, opacities=op, colors_precomp=col, scales=s,
                                                     rotations=r)
        (out[0].sum() + out[2].sum()).backward()
        torch.cuda.synchronize()

    step()
    ms, n = (C.c_double * 9)(), (C.c_int * 9)()
    lib.gsr_profile_collect(ms, n)
    try:
        lib.gsr_profile_enable(1)
        step()
        lib.gsr_profile_enable(0)
        assert lib.gsr_profile_collect(ms, n) == 0
        timed = {i for i in range(9) if n[i] > 0}
        assert {0, 1, 2, 3, 4, 6, 7, 8} <= timed and all(ms[i] > 0 for i in timed)
        RENDER_BWD = _abi.STAGES.index("render_bwd")
        lib.gsr_profile_enable(1 | (1 << (8 + RENDER_BWD)))
        step()
        lib.gsr_profile_enable(0)
        assert lib.gsr_profile_collect(ms, n) == 0
        assert [i for i in range(9) if n[i] > 0] == [RENDER_BWD] and ms[RENDER_BWD] > 0
        step()                                            # disabled: nothing recorded
        assert lib.gsr_profile_collect(ms, n) == 0 and sum(n) == 0
    finally:
        lib.gsr_profile_enable(0)
