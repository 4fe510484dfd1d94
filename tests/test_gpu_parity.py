"""GPU parity tests proper (-m gpu): the sm_100a path, called through the C ABI (monogs_b200._abi via the
`_C` mirror of the reference's pybind module), against
  (1) the committed golden vectors = outputs of the unmodified reference CUDA code (tests/golden/*.npz),
  (2) the CPU oracle on fresh seeded inputs,
  (3) the reference CUDA library itself (oracle/_ref/libref_dgr.so) when it travelled to the box,
  (4) size-independent properties at BASELINE.json's full sizes.
Bars: bit-exact radii, sorted keys, point_list, tile ranges, n_touched; colour/depth/opacity <= 1e-5 relative;
gradients (incl. pose) <= 1e-4 relative.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import ROOT, rel_inf
from monogs_b200 import _abi, synth
from monogs_b200.diff_gaussian_rasterization import _C as ours
from oracle import cpu_oracle as co
from oracle import ref_cuda as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"
GRAD_NAMES = ["dmeans2D", "dcolors", "dopacity", "dmeans3D", "dcov3D", "dsh", "dscales", "drots", "dtau"]


def _fargs(cam, sc, bg, debug=True, sh=None, degree=0, cov=None, scale_modifier=1.0):
    view, proj, proj_raw, campos = cam.matrices(DEV)
    e = torch.empty(0, device=DEV)
    colors = e if sh is not None else sc["colors"]
    scales = e if cov is not None else sc["scales"]
    rots = e if cov is not None else sc["rotations"]
    return (bg, sc["means3D"], colors, sc["opacities"], scales, rots, scale_modifier, e if cov is None else cov,
            view, proj, proj_raw, cam.tanfovx, cam.tanfovy, cam.H, cam.W, e if sh is None else sh, degree, campos,
            False, debug)


def _bargs(fa, radii, R, geom, plist, img, dcol, ddep):
    (bg, m3, colors, _, scales, rots, smod, cov, view, proj, proj_raw, tx, ty, H, W, sh, degree, campos, _, _) = fa
    return (bg, m3, radii, colors, scales, rots, smod, cov, view, proj, proj_raw, tx, ty, dcol, ddep, sh, degree,
            campos, geom, R, plist, img, False)


def _image_state(img, W, H):
    lib = _abi.load()
    ft, nc, rg = C.c_void_p(), C.c_void_p(), C.c_void_p()
    assert lib.gsr_debug_image_views(img.data_ptr(), W, H, C.byref(ft), C.byref(nc), C.byref(rg)) == 0
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    v = lambda p, nbytes: img[p.value - img.data_ptr(): p.value - img.data_ptr() + nbytes]
    return (v(ft, 4 * W * H).view(torch.float32).cpu().numpy(), v(nc, 4 * W * H).view(torch.int32).cpu().numpy(),
            v(rg, 8 * tiles).view(torch.int32).cpu().numpy().reshape(-1, 2))


def _geom_state(geom, P, W, H):
    lib = _abi.load()
    pa, pb, pc, pt = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    assert lib.gsr_debug_geom_views(geom.data_ptr(), P, W, H, C.byref(pa), C.byref(pb), C.byref(pc), C.byref(pt)) == 0
    v = lambda p, n: geom[p.value - geom.data_ptr(): p.value - geom.data_ptr() + 4 * n]
    assert pb.value - pa.value == 16 and pc.value - pa.value == 32      # one packed 48-byte record per Gaussian
    rec = v(pa, 12 * P).view(torch.float32).view(P, 12).cpu().numpy()
    ra, rb = rec[:, 0:4], rec[:, 4:8]
    tt = v(pt, P).view(torch.int32).cpu().numpy()
    return ra, rb, tt


def _run_ours(cam, sc, bg, dcol, ddep, **kw):
    fa = _fargs(cam, sc, bg, **kw)
    R, col, radii, geom, plist, img, dep, opa, nt = ours.rasterize_gaussians(*fa)
    binning = ours.rasterize_gaussians.last_binning
    P = sc["means3D"].shape[0]
    keys = ours.debug_sorted_keys(geom, binning, plist, P, R, cam.W, cam.H) if R > 0 else torch.empty(0)
    ba = _bargs(fa, radii, R, geom, plist, img, dcol, ddep)
    grads = ours.rasterize_gaussians_backward(*ba)
    grads2 = ours.rasterize_gaussians_backward(*ba)     # second run: our own run-to-run spread (fp32 atomics)
    _, tau = ours.rasterize_gaussians_backward_fused(*ba)
    torch.cuda.synchronize()
    return dict(R=R, color=col, radii=radii, geom=geom, plist=plist, img=img, depth=dep, opacity=opa, n_touched=nt,
                keys=keys, grads=grads, grads2=grads2, tau=tau, fa=fa)


# ----------------------------------------------------------------------------------------------------------------
# (1) golden vectors of the unmodified reference
# ----------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["iso_small", "aniso_layered"])
def test_against_reference_golden(built, golden_mod, name):
    cam, sc, bg, dcol, ddep = golden_mod.case_inputs(name, DEV)
    g = np.load(ROOT / "tests" / "golden" / f"{name}.npz")
    o = _run_ours(cam, sc, bg, dcol, ddep)
    P, R = sc["means3D"].shape[0], int(g["num_rendered"])
    assert o["R"] == R
    assert (o["radii"].cpu().numpy() == g["radii"]).all()
    assert (o["n_touched"].cpu().numpy() == g["n_touched"]).all()
    assert (o["keys"].cpu().numpy().view(np.uint64) == g["keys"]).all()
    assert (o["plist"].view(torch.int32)[:R].cpu().numpy() == g["point_list"]).all()
    final_T, n_contrib, ranges = _image_state(o["img"], cam.W, cam.H)
    assert (ranges == g["ranges"]).all()
    assert (n_contrib == g["n_contrib"]).all()
    assert (final_T.view(np.uint32) == g["final_T"].view(np.uint32)).all()
    ra, rb, tt = _geom_state(o["geom"], P, cam.W, cam.H)
    vis = g["radii"] > 0
    assert (tt == g["tiles_touched"]).all()
    assert (ra[vis, :2].view(np.uint32) == g["means2D"][vis].view(np.uint32)).all()
    assert (ra[vis, 2].view(np.uint32) == g["depths"][vis].view(np.uint32)).all()
    assert (rb[vis].view(np.uint32) == g["conic_opacity"][vis].view(np.uint32)).all()
    for k in ("color", "depth", "opacity"):
        assert rel_inf(o[k].cpu().numpy(), g[k]) <= 1e-5, k  # tolerance: 1e-5 relative
    for n, t in zip(GRAD_NAMES, o["grads"]):
        if n in g.files:
            assert rel_inf(t.cpu().numpy(), g[n]) <= 1e-4, (n, rel_inf(t.cpu().numpy(), g[n]))  # 1e-4 relative
    assert rel_inf(o["tau"].cpu().numpy(), g["dtau"].astype(np.float64).sum(0)) <= 1e-4


# ----------------------------------------------------------------------------------------------------------------
# (2) CPU oracle on fresh seeded inputs (sizes the oracle finishes in seconds)
# ----------------------------------------------------------------------------------------------------------------
synth.INTRINSICS.setdefault("tiny", (160, 120, 133.85, 134.8, 80.03, 61.9))
ORACLE_CASES = [
    # P, kind, seed, anisotropic, layered
    (10_000, "tum_half", 0, False, False),   # BASELINE config C1
    (6_000, "tiny", 1, True, False),
    (4_000, "tiny", 2, False, True),
    (1, "tiny", 3, False, False),            # single Gaussian
    (257, "tiny", 4, True, True),            # ragged: one over a 256 block
]


@pytest.mark.parametrize("P,kind,seed,aniso,layered", ORACLE_CASES)
def test_against_cpu_oracle(built, P, kind, seed, aniso, layered):
    cam = synth.make_camera(kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=seed, anisotropic=aniso, layered=layered, frac_behind=0.05 if P > 20 else 0.0,
                          device=DEV)
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    g = torch.Generator().manual_seed(seed + 7)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    o = _run_ours(cam, sc, bg, dcol, ddep)
    view, proj, proj_raw, campos = cam.matrices()
    n = lambda t: t.detach().cpu().numpy()
    f = co.forward(n(sc["means3D"]), n(sc["scales"]), n(sc["rotations"]), n(sc["opacities"]), n(sc["colors"]),
                   n(view), n(proj), n(campos), cam.W, cam.H, cam.tanfovx, cam.tanfovy, n(bg))
    b = co.backward(f, n(sc["means3D"]), n(sc["scales"]), n(sc["rotations"]), n(view), n(proj), n(proj_raw), cam.W,
                    cam.H, cam.tanfovx, cam.tanfovy, n(bg), n(dcol), n(ddep))
    # the g++ oracle rounds the anisotropic covariance differently in a few last bits -> a radius can flip
    exact = not aniso
    if exact:
        assert o["R"] == f["num_rendered"]
        assert (n(o["radii"]) == f["radii"]).all()
        assert (n(o["keys"]).view(np.uint64) == f["keys"]).all()
        assert (n(o["plist"].view(torch.int32)[:o["R"]]).view(np.uint32) == f["point_list"]).all()
        _, _, ranges = _image_state(o["img"], cam.W, cam.H)
        assert (ranges.view(np.uint32) == f["ranges"]).all()
    else:
        assert (n(o["radii"]) != f["radii"]).mean() <= 2e-3
    assert (n(o["n_touched"]) != f["n_touched"]).mean() <= 2e-3  # host expf vs device: threshold flips only
    for k in ("color", "depth", "opacity"):
        assert rel_inf(n(o[k]), f[k]) <= 1e-5, k
    if exact:
        pairs = [("means2D", 0), ("colors", 1), ("opacity", 2), ("means3D", 3), ("cov3D", 4), ("scales", 6),
                 ("rotations", 7), ("tau", 8)]
        for k, i in pairs:
            assert rel_inf(n(o["grads"][i]), b[k]) <= 1e-4, (k, rel_inf(n(o["grads"][i]), b[k]))
        assert rel_inf(n(o["tau"]), b["tau_sum"]) <= 1e-4


# ----------------------------------------------------------------------------------------------------------------
# (3) the reference CUDA library itself, same process, same inputs
# ----------------------------------------------------------------------------------------------------------------
needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libref_dgr.so not shipped")
REF_CASES = [
    (100_000, "tum", 1, False, False),       # BASELINE config C2 shape
    (50_000, "tum", 2, True, False),
    (60_000, "tum", 3, False, True),
    (300_000, "replica", 4, False, False),
    (200_000, "hd", 5, False, False),        # 8160 tiles > 4096: the instance radix-sort binning path (C5 shape)
    (3_000, "tum_half", 6, False, False),    # fewer Gaussians than binning CTAs
    # BASELINE.json's real sizes (the cases tools/compare_gpu.py ran by hand in round 1)
    (1_000_000, "tum", 0, False, False),     # HEADLINE: 640x480, 1 M Gaussians
    (500_000, "replica", 0, False, False),   # C3: 1200x680, 500 k
    (2_000_000, "replica", 2, True, True),   # C4: 1200x680, 2 M, anisotropic + layered (long per-tile lists)
    (4_000_000, "hd", 0, False, False),      # C5: 1920x1080, 4 M (8160 tiles -> instance radix-sort binning)
]


def _compare_with_reference(cam, sc, bg, dcol, ddep, grad_tol=1e-4, well_conditioned=None):
    """well_conditioned: optional bool[P]; the gradients that pass through the 2D-covariance inverse (means3D, cov3D,
    scales, rotations, tau) are then compared on those Gaussians only (the renderer's own outputs -- mean2D, colour,
    opacity gradients -- always on all)."""
    P = sc["means3D"].shape[0]
    o = _run_ours(cam, sc, bg, dcol, ddep)
    fa = o["fa"]
    Rr, col_r, rad_r, _, _, _, dep_r, opa_r, nt_r = ref.rasterize_gaussians(*fa)
    torch.cuda.synchronize()
    st = ref.last_state(P, Rr, cam.W, cam.H, DEV)
    gr = ref.rasterize_gaussians_backward(*_bargs(fa, rad_r, Rr, None, None, None, dcol, ddep))
    gr2 = ref.rasterize_gaussians_backward(*_bargs(fa, rad_r, Rr, None, None, None, dcol, ddep))
    torch.cuda.synchronize()
    assert o["R"] == Rr
    assert torch.equal(o["radii"], rad_r)
    assert torch.equal(o["n_touched"], nt_r)
    assert torch.equal(o["keys"], st["point_list_keys"])
    assert torch.equal(o["plist"].view(torch.int32)[:Rr], st["point_list"])
    final_T, n_contrib, ranges = _image_state(o["img"], cam.W, cam.H)
    assert (ranges.ravel() == st["ranges"].cpu().numpy()).all()
    assert (n_contrib == st["n_contrib"].cpu().numpy()).all()
    assert (final_T.view(np.uint32) == st["accum_alpha"].cpu().numpy().view(np.uint32)).all()
    assert torch.equal(o["color"].view(torch.int32), col_r.view(torch.int32))        # image BITS, not just 1e-5
    n = lambda t: t.detach().cpu().numpy()
    for a, r, k in ((o["color"], col_r, "color"), (o["depth"], dep_r, "depth"), (o["opacity"], opa_r, "opacity")):
        assert rel_inf(n(a), n(r)) <= 1e-5, k
    # Bar: the north star's 1e-4 relative per tensor. Both implementations accumulate fp32 partial gradients with
    # atomics in arbitrary order; on the cancellation-dominated rotation / scale gradients of the anisotropic layered
    # 2 M scene the reference disagrees with its own second run by 1-3e-5 and this library by 1-2e-5 (one global
    # atomic per (tile, Gaussian) like the reference), the two with each other by 3-6e-5. Should a box ever show a
    # larger run-to-run spread, the bar follows 3 x the measured spread of the two -- never beyond 5e-4.
    for name, a, a2, r, r2 in zip(GRAD_NAMES, o["grads"], o["grads2"], gr, gr2):
        if a.numel():
            if well_conditioned is not None and name in ("dmeans3D", "dcov3D", "dscales", "drots", "dtau"):
                a, a2, r, r2 = a[well_conditioned], a2[well_conditioned], r[well_conditioned], r2[well_conditioned]
            spread = rel_inf(n(r2), n(r)) + rel_inf(n(a2), n(a))
            err = rel_inf(n(a), n(r))
            assert err <= max(grad_tol, 3 * spread), (name, err, spread)
            assert err <= 5 * grad_tol, (name, err)   # and never more than 5e-4, whatever the spread
    if well_conditioned is None:
        assert rel_inf(n(o["tau"]), n(gr[8]).astype(np.float64).sum(0)) <= max(grad_tol, 1e-4)
    return o


@needs_ref
@pytest.mark.parametrize("P,kind,seed,aniso,layered", REF_CASES)
def test_against_reference_cuda(built, P, kind, seed, aniso, layered):
    cam = synth.make_camera(kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=seed, anisotropic=aniso, layered=layered, device=DEV)
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    g = torch.Generator().manual_seed(seed + 7)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    _compare_with_reference(cam, sc, bg, dcol, ddep)


@needs_ref
def test_huge_and_near_degenerate_gaussians_against_reference(built):
    """Splats the per-rectangle cull of the renderers must never get wrong: footprints of hundreds to thousands of
    pixels, needle-shaped (near-singular 2D covariance, conic terms that cancel), faint (255 * opacity barely above
    1) and fully opaque ones, mixed into an ordinary scene. Every integer output and the image bits must still equal
    the reference's: a pair culled although the reference blends it would change final_T / n_contrib / colour bits."""
    P = 30_000
    cam = synth.make_camera("tum_half", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=77, anisotropic=True, device=DEV)
    g = torch.Generator().manual_seed(78)
    k = 600
    idx = torch.randperm(P, generator=g)[:k].to(DEV)
    z = sc["means3D"][idx, 2].abs().clamp_min(0.3)
    shape = torch.ones(k, 3)
    shape[:200] = torch.tensor([3.0, 1e-3, 1e-3])                      # needles: ~1000 px long, ~0.3 px wide
    shape[200:400] = torch.tensor([2.0, 2.0, 1e-3])                    # sheets seen at random angles
    shape[400:] = torch.tensor([1.5, 1.5, 1.5])                        # blobs covering the whole image
    sc["scales"][idx] = (shape.to(DEV) * z[:, None] * (0.2 + torch.rand(k, 1, generator=g).to(DEV))).contiguous()
    op = sc["opacities"].clone()
    op[idx[0:k:3]] = 1.0                                               # fully opaque
    op[idx[1:k:3]] = (1.0 / 255.0) * (1.0 + 0.05 * torch.rand(len(idx[1:k:3]), 1, generator=g).to(DEV))  # barely visible
    sc["opacities"] = op.contiguous()
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    # The needles and sheets have 2D covariances of aspect ratio ~3000: the gradient THROUGH their inverse is
    # ill-conditioned in both implementations (the reference disagrees with its own second run by 30 % on them:
    # atomic order), so those gradients are compared on all other Gaussians; the renderer's own gradient outputs
    # (mean2D, colour, opacity) are compared on every Gaussian, and every integer / image-bit bar stays exact.
    well = torch.ones(P, dtype=torch.bool, device=DEV)
    well[idx[:400]] = False
    o = _compare_with_reference(cam, sc, bg, dcol, ddep, well_conditioned=well)
    assert int(o["radii"].max()) > 500


@needs_ref
@pytest.mark.parametrize("case", ["planes", "one_depth", "far_and_near", "tiny"])
def test_depth_order_with_equal_and_crowded_depths_against_reference(built, case):
    """The depth order of the Gaussians must be the reference's -- depth bits, ties by Gaussian id (its stable sort of
    keys emitted in id order) -- also in degenerate scenes: tens of thousands of EQUAL depths (five planes, or every
    Gaussian at one depth), depths of 2e4 .. 3e6 m and just above the 0.2 m near plane with ties there too, and a
    map of 37 Gaussians. Identity camera, so view depth = z exactly."""
    cam = synth.make_camera("tum_half")
    P = {"planes": 120_000, "one_depth": 60_000, "far_and_near": 40_000, "tiny": 37}[case]
    sc = synth.make_scene(P, cam, seed=41, device=DEV)
    g = torch.Generator().manual_seed(42)
    m = sc["means3D"].clone()
    z = m[:, 2].clone()
    front = z > 0.2
    if case == "planes":          # five planes of ~20 k Gaussians each + 30 short runs + the rest random
        sel = torch.rand(P, generator=g).to(DEV)
        planes = torch.tensor([0.75, 1.5, 1.5000001, 2.25, 4.0], device=DEV)
        pick = (sel * 6).long().clamp_max(5)
        znew = torch.where(pick < 5, planes[pick.clamp_max(4)], z)
        runs = (sel * 4000).long() % 130
        znew = torch.where((pick == 5) & (runs < 30), 3.0 + 0.001 * runs.float(), znew)
    elif case == "one_depth":
        znew = torch.full_like(z, 2.0)
    elif case == "far_and_near":  # beyond both ends of the binned range, with ties there too
        sel = torch.rand(P, generator=g).to(DEV)
        znew = torch.where(sel < 0.3, 20000.0 + 5000.0 * (sel * 1000).floor() / 300.0, z)
        znew = torch.where(sel > 0.9, torch.full_like(z, 0.2000001), znew)
        znew = torch.where((sel > 0.8) & (sel <= 0.9), torch.full_like(z, 3.0e6), znew)
    else:
        znew = z
    scale = torch.where(front, znew / z.clamp_min(1e-6), torch.ones_like(z))
    m[:, 0] = torch.where(front, m[:, 0] * scale, m[:, 0])      # keep the pixel, move along the ray
    m[:, 1] = torch.where(front, m[:, 1] * scale, m[:, 1])
    m[:, 2] = torch.where(front, znew, z)
    sc["means3D"] = m.contiguous()
    sc["scales"] = (sc["scales"] * scale[:, None].clamp(1e-3, 1e3)).contiguous()
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    o = _compare_with_reference(cam, sc, bg, dcol, ddep)
    if case != "tiny":
        keys = o["keys"].cpu().numpy().astype(np.int64)
        depth_bits = keys & 0xFFFFFFFF
        tile = keys >> 32
        same = (tile[1:] == tile[:-1]) & (depth_bits[1:] == depth_bits[:-1])
        assert same.sum() > 1000                                   # the scene really has ties inside tiles ...
        pl = o["plist"].view(torch.int32).cpu().numpy()[:len(keys)].astype(np.int64)
        assert (pl[1:][same] > pl[:-1][same]).all()                # ... and they are in id order


@needs_ref
def test_sh_and_cov3d_precomp_paths_against_reference(built):
    """shs (degree 3) instead of precomputed colours, and cov3D_precomp instead of scale/rotation."""
    P = 20_000
    cam = synth.make_camera("tum_half", rot=(0.01, 0.0, -0.01), trans=(0.0, 0.02, 0.0))
    sc = synth.make_scene(P, cam, seed=9, anisotropic=True, device=DEV)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    g = torch.Generator().manual_seed(99)
    sh = (torch.randn(P, 16, 3, generator=g) * 0.3).to(DEV)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    n = lambda t: t.detach().cpu().numpy()
    for kw in (dict(sh=sh, degree=3), dict(sh=sh, degree=1), dict(cov=None)):
        if "cov" in kw:
            # cov3D from the reference's own forward state
            fa0 = _fargs(cam, sc, bg)
            R0 = ref.rasterize_gaussians(*fa0)[0]
            cov = ref.last_state(P, R0, cam.W, cam.H, DEV)["cov3D"].view(P, 6).clone()
            cov[~torch.isfinite(cov).all(1)] = 0
            cov[(ref.rasterize_gaussians(*fa0)[2] <= 0)] = 1e-4 * torch.tensor([1, 0, 0, 1, 0, 1.0], device=DEV)
            kw = dict(cov=cov)
        o = _run_ours(cam, sc, bg, dcol, ddep, **kw)
        fa = o["fa"]
        Rr, col_r, rad_r, _, _, _, dep_r, opa_r, nt_r = ref.rasterize_gaussians(*fa)
        gr = ref.rasterize_gaussians_backward(*_bargs(fa, rad_r, Rr, None, None, None, dcol, ddep))
        torch.cuda.synchronize()
        assert o["R"] == Rr and torch.equal(o["radii"], rad_r)
        assert (o["n_touched"] != nt_r).float().mean().item() <= 1e-4
        assert rel_inf(n(o["color"]), n(col_r)) <= 1e-5
        for name, a, r in zip(GRAD_NAMES, o["grads"], gr):
            if a.numel():
                assert rel_inf(n(a), n(r)) <= 1e-4, (name, kw.keys(), rel_inf(n(a), n(r)))


# ----------------------------------------------------------------------------------------------------------------
# (4) properties at full size + edge cases
# ----------------------------------------------------------------------------------------------------------------
def test_full_size_properties_headline(built):
    """640x480, 1 M Gaussians: sortedness of the (tile, depth) keys, ranges partition [0,R), image linearity in
    colour, backward determinism bound."""
    P = 1_000_000
    cam = synth.make_camera("tum", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=0, device=DEV)
    bg = torch.zeros(3, device=DEV)
    g = torch.Generator().manual_seed(7)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    o = _run_ours(cam, sc, bg, dcol, ddep)
    R = o["R"]
    keys = o["keys"]
    assert R > P and bool((keys[1:] >= keys[:-1]).all())                     # sorted
    pl = o["plist"].view(torch.int32)[:R].long()
    assert int(pl.min()) >= 0 and int(pl.max()) < P
    same = keys[1:] == keys[:-1]
    assert bool((pl[1:][same] > pl[:-1][same]).all())                        # stable: ties in id order
    _, n_contrib, ranges = _image_state(o["img"], cam.W, cam.H)
    nz = ranges[ranges[:, 1] > ranges[:, 0]]
    assert nz[0, 0] == 0 and nz[-1, 1] == R and (nz[1:, 0] == nz[:-1, 1]).all()  # ranges partition [0, R)
    tiles_of_keys = (keys >> 32).cpu().numpy()
    assert (np.bincount(tiles_of_keys, minlength=len(ranges)) == (ranges[:, 1] - ranges[:, 0])).all()
    tt = _geom_state(o["geom"], P, cam.W, cam.H)[2]
    assert int(tt.astype(np.int64).sum()) == R                                # checksum of tile counts
    # linearity in colour: render(2c) == 2 render(c) exactly (power-of-two scaling commutes with rounding)
    sc2 = dict(sc); sc2["colors"] = sc["colors"] * 2
    col2 = ours.rasterize_gaussians(*_fargs(cam, sc2, bg, debug=False))[1]
    assert torch.equal(col2, o["color"] * 2)
    # opacity in [0, 1], depth >= 0
    assert float(o["opacity"].min()) >= 0 and float(o["opacity"].max()) <= 1.0 + 1e-6
    # sum of colour gradients == sum over pixels of dL/dpix * (alpha T) -- checked through a second backward
    g2 = ours.rasterize_gaussians_backward(*_bargs(o["fa"], o["radii"], R, o["geom"], o["plist"], o["img"], dcol,
                                                   ddep))
    for a, b in zip(o["grads"], g2):
        if a.numel():
            assert rel_inf(b.cpu().numpy(), a.cpu().numpy()) <= 1e-4  # atomics: order-dependent fp32 sums


def test_edge_cases(built):
    cam = synth.make_camera("tiny")
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    e = torch.empty(0, device=DEV)
    view, proj, proj_raw, campos = cam.matrices(DEV)
    # P == 0 -> zero images like the reference (rasterize_points.cu:69-73,85)
    z = torch.zeros(0, 3, device=DEV)
    R, col, radii, *_rest = ours.rasterize_gaussians(bg, z, z, torch.zeros(0, 1, device=DEV), z,
                                                     torch.zeros(0, 4, device=DEV), 1.0, e, view, proj, proj_raw,
                                                     cam.tanfovx, cam.tanfovy, cam.H, cam.W, e, 0, campos, False, False)
    assert R == 0 and col.shape == (3, cam.H, cam.W) and float(col.abs().max()) == 0 and radii.numel() == 0
    # everything behind the camera -> R == 0, image == background, backward gives zeros
    sc = synth.make_scene(500, cam, seed=1, device=DEV)
    sc["means3D"][:, 2] = -sc["means3D"][:, 2].abs() - 1
    fa = _fargs(cam, sc, bg, debug=False)
    R, col, radii, geom, plist, img, dep, opa, nt = ours.rasterize_gaussians(*fa)
    assert R == 0 and int(radii.abs().sum()) == 0 and int(nt.sum()) == 0
    assert torch.allclose(col[:, 5, 5], bg) and float(opa.max()) == 0
    dcol = torch.ones(3, cam.H, cam.W, device=DEV)
    out = ours.rasterize_gaussians_backward(*_bargs(fa, radii, R, geom, plist, img, dcol, None))
    assert all(float(t.abs().max()) == 0 for t in out if t.numel())
    # bad shape raises like the reference's AT_ERROR
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        ours.rasterize_gaussians(bg, torch.zeros(5, 4, device=DEV), *fa[2:])
    # image not a multiple of the tile size (ragged tiles)
    synth.INTRINSICS["odd"] = (150, 97, 120.0, 121.0, 75.3, 48.1)
    camo = synth.make_camera("odd")
    sco = synth.make_scene(3000, camo, seed=3, device=DEV)
    o = _run_ours(camo, sco, bg, torch.ones(3, 97, 150, device=DEV) / 1e4, torch.ones(1, 97, 150, device=DEV) / 1e4)
    n = lambda t: t.detach().cpu().numpy()
    vo, po, pro, co_ = camo.matrices()
    f = co.forward(n(sco["means3D"]), n(sco["scales"]), n(sco["rotations"]), n(sco["opacities"]), n(sco["colors"]),
                   n(vo), n(po), n(co_), 150, 97, camo.tanfovx, camo.tanfovy, n(bg))
    assert o["R"] == f["num_rendered"] and (n(o["radii"]) == f["radii"]).all()
    assert rel_inf(n(o["color"]), f["color"]) <= 1e-5


def test_mark_visible(built):
    cam = synth.make_camera("tiny", rot=(0.1, 0.2, 0.0))
    sc = synth.make_scene(5000, cam, seed=2, frac_behind=0.3, device=DEV)
    view, proj, _, _ = cam.matrices(DEV)
    m = ours.mark_visible(sc["means3D"], view, proj)
    want = co.mark_visible(sc["means3D"].cpu().numpy(), view.cpu().numpy())
    assert m.dtype == torch.bool and (m.cpu().numpy() == want).all() and 0 < int(m.sum()) < 5000


def test_single_call_forward_equals_two_phase(built):
    """gsr_forward (instance kernels enqueued for a remembered capacity, count read on the device) must give the
    bits of gsr_forward_plan + gsr_forward_render, also through the GSR_ECAPACITY retry."""
    cam = synth.make_camera("tum_half", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(30_000, cam, seed=21, device=DEV)
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    key = (30_000, cam.W, cam.H, torch.device(DEV, torch.cuda.current_device()).index)
    ours._capacity.pop(key, None)
    fa = _fargs(cam, sc, bg, debug=False)
    first = ours.rasterize_gaussians(*fa)                 # two-phase (no capacity known yet)
    assert key in ours._capacity or any(k[:3] == key[:3] for k in ours._capacity)
    second = ours.rasterize_gaussians(*fa)                # single call
    for k in list(ours._capacity):
        if k[:3] == key[:3]:
            ours._capacity[k] = 1000                      # far too small -> GSR_ECAPACITY -> retry
    third = ours.rasterize_gaussians(*fa)
    R = first[0]
    for other in (second, third):
        assert other[0] == R
        for i in (1, 2, 6, 7, 8):                         # color, radii, depth, opacity, n_touched
            assert torch.equal(first[i], other[i]), i
        assert torch.equal(first[4].view(torch.int32)[:R], other[4].view(torch.int32)[:R])  # point_list
        _, nc0, rg0 = _image_state(first[5], cam.W, cam.H)
        _, nc1, rg1 = _image_state(other[5], cam.W, cam.H)
        assert (nc0 == nc1).all() and (rg0 == rg1).all()
    # and the backward consumes the capacity-sized point list
    dcol = torch.ones(3, cam.H, cam.W, device=DEV) / (cam.H * cam.W)
    g1 = ours.rasterize_gaussians_backward(*_bargs(fa, first[2], R, first[3], first[4], first[5], dcol, None))
    g2 = ours.rasterize_gaussians_backward(*_bargs(fa, second[2], R, second[3], second[4], second[5], dcol, None))
    for a, b in zip(g1, g2):
        if a.numel():
            assert rel_inf(b.cpu().numpy(), a.cpu().numpy()) <= 1e-5


def test_window_rows_backward_equals_plain_backward(built):
    """gsr_backward_render + gsr_backward_rows (chunked, write then accumulate) == gsr_backward column for column."""
    from monogs_b200 import window
    cam = synth.make_camera("tum_half", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    P = 20_000
    sc = synth.make_scene(P, cam, seed=31, anisotropic=True, device=DEV)
    bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
    g = torch.Generator().manual_seed(5)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
    fa = _fargs(cam, sc, bg, debug=False)
    R, col, radii, geom, plist, img, dep, opa, nt = ours.rasterize_gaussians(*fa)
    plain = ours.rasterize_gaussians_backward(*_bargs(fa, radii, R, geom, plist, img, dcol, ddep))
    view, proj, proj_raw, campos = cam.matrices(DEV)
    rows = torch.full((P, 16), 7.0, device=DEV)       # garbage: accumulate=False must overwrite it
    chunks = window.window_chunks(P, 4)
    assert len(chunks) == 4
    seen = []
    tau, m2d = ours.rasterize_gaussians_backward_rows(
        bg, sc["means3D"], radii, sc["scales"], sc["rotations"], 1.0, view, proj, proj_raw, cam.tanfovx, cam.tanfovy,
        dcol, ddep, geom, R, plist, img, rows, False, chunks=chunks, chunk_done=lambda a, b: seen.append((a, b)),
        want_means2D=True)
    assert seen == chunks
    n = lambda t: t.detach().cpu().numpy()
    sl = window.column_slices()
    vis = (radii > 0)
    # two separate backward passes: the float atomics of the tile renderer land in a different order each time,
    # so the two results agree to accumulation noise (the gradient bar of the north star is 1e-4)
    for name, i in (("means3D", 3), ("colors", 1), ("opacities", 2), ("scales", 6), ("rotations", 7)):
        assert rel_inf(n(rows[:, sl[name]]), n(plain[i])) <= 5e-5, name
    assert rel_inf(n(m2d), n(plain[0])) <= 5e-5
    assert torch.equal(rows[:, 15], vis.float())
    assert rel_inf(n(rows[:, 14]), n(plain[0][:, :2].norm(dim=1) * vis)) <= 5e-5
    assert rel_inf(n(tau), n(plain[8]).astype(np.float64).sum(0)) <= 1e-4
    # second keyframe added on top: rows == 2 x single (same keyframe twice)
    once = rows.clone()
    ours.rasterize_gaussians_backward_rows(
        bg, sc["means3D"], radii, sc["scales"], sc["rotations"], 1.0, view, proj, proj_raw, cam.tanfovx, cam.tanfovy,
        dcol, ddep, geom, R, plist, img, rows, True, chunks=[(0, P)])
    assert rel_inf(n(rows), 2 * n(once)) <= 1e-5
    # GSR_ROWS_TOUCHED_ONLY: into zero-filled rows only the Gaussians with a gradient are written (then added);
    # columns 0..14 must equal the dense form, untouched rows stay exactly zero, column 15 is not maintained
    sparse = torch.zeros(P, 16, device=DEV)
    tau_s, _ = ours.rasterize_gaussians_backward_rows(
        bg, sc["means3D"], radii, sc["scales"], sc["rotations"], 1.0, view, proj, proj_raw, cam.tanfovx, cam.tanfovy,
        dcol, ddep, geom, R, plist, img, sparse, False, chunks=[(0, P)], touched_only=True)
    assert rel_inf(n(sparse[:, :15]), n(once[:, :15])) <= 5e-5
    untouched = (once[:, :15] == 0).all(1)
    assert int(untouched.sum()) > 0 and float(sparse[untouched].abs().max()) == 0.0
    assert rel_inf(n(tau_s), n(tau)) <= 1e-5
    ours.rasterize_gaussians_backward_rows(
        bg, sc["means3D"], radii, sc["scales"], sc["rotations"], 1.0, view, proj, proj_raw, cam.tanfovx, cam.tanfovy,
        dcol, ddep, geom, R, plist, img, sparse, True, chunks=[(0, P)], touched_only=True)
    assert rel_inf(n(sparse[:, :15]), 2 * n(once[:, :15])) <= 5e-5


def test_mapping_window_cuda_backend_single_gpu(built):
    """MappingWindow with the product backend on one GPU: 3 keyframes summed == sum of 3 plain backward calls."""
    from monogs_b200 import window
    cams = synth.arc_cameras("tum_half", 3)
    P = 15_000
    sc = synth.make_scene(P, cams[1], seed=41, device=DEV)
    bg = torch.tensor([0.0, 0.0, 0.0], device=DEV)
    targets = [tuple(t.to(DEV) for t in synth.make_targets(c, seed=i)) for i, c in enumerate(cams)]
    params = dict(sc)
    w = window.MappingWindow(window.make_cuda_render_fn(cams, targets, bg))
    res = w.step(params, 3)
    assert [r.index for r in res.local] == [0, 1, 2] and all(r.grad_tau.shape == (6,) for r in res.local)
    want = {k: torch.zeros_like(v) for k, v in res.grads.items()}
    vis_count = torch.zeros(P, device=DEV)
    max_r = torch.zeros(P, dtype=torch.int32, device=DEV)
    for cam, (gi, gd) in zip(cams, targets):
        fa = _fargs(cam, sc, bg, debug=False)
        R, col, radii, geom, plist, img, dep, opa, nt = ours.rasterize_gaussians(*fa)
        loss, dcol, ddep = window.mapping_loss_grad(col, dep, gi, gd)
        o = ours.rasterize_gaussians_backward(*_bargs(fa, radii, R, geom, plist, img, dcol.contiguous(),
                                                      ddep.contiguous()))
        for name, i in (("means3D", 3), ("colors", 1), ("opacities", 2), ("scales", 6), ("rotations", 7)):
            want[name] += o[i]
        vis_count += (radii > 0).float()
        max_r = torch.maximum(max_r, radii)
    n = lambda t: t.detach().cpu().numpy()
    for name in want:
        assert rel_inf(n(res.grads[name]), n(want[name])) <= 1e-4, name
    assert torch.equal(res.vis_count, vis_count) and torch.equal(res.max_radii, max_r)


# ----------------------------------------------------------------------------------------------------------------
# (5) the two binning implementations produce the same lists: GSR_BINNING=radix (read once per process, so a
#     child process) against the default sort-free path at a size both support
# ----------------------------------------------------------------------------------------------------------------
_BINNING_CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
import test_gpu_parity as T
from monogs_b200 import synth
cam = synth.make_camera("tum", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
sc = synth.make_scene(120_000, cam, seed=11, device="cuda")
bg = torch.tensor([0.1, 0.2, 0.3], device="cuda")
z3, z1 = torch.zeros(3, cam.H, cam.W, device="cuda"), torch.zeros(1, cam.H, cam.W, device="cuda")
o = T._run_ours(cam, sc, bg, z3, z1)
_, n_contrib, ranges = T._image_state(o["img"], cam.W, cam.H)
np.savez(sys.argv[2], R=o["R"], plist=o["plist"].cpu().numpy()[:o["R"]], keys=o["keys"].cpu().numpy(), ranges=ranges,
         n_contrib=n_contrib, color=o["color"].cpu().numpy(), n_touched=o["n_touched"].cpu().numpy())
"""


def test_radix_and_sort_free_binning_agree(built, tmp_path):
    import os, subprocess, sys
    outs = {}
    for mode in ("radix", "single", "twolevel", "scatter"):     # "scatter" (any other value) = the automatic choice
        env = dict(os.environ, GSR_BINNING=mode)
        f = tmp_path / f"{mode}.npz"
        r = subprocess.run([sys.executable, "-c", _BINNING_CHILD, str(ROOT), str(f)], env=env, capture_output=True,
                           text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[mode] = np.load(f)
    a = outs["radix"]
    for other in ("single", "twolevel", "scatter"):
        b = outs[other]
        assert int(a["R"]) == int(b["R"]) > 0
        for k in ("plist", "keys", "ranges", "n_contrib", "n_touched", "color"):
            assert np.array_equal(a[k], b[k]), (other, k)


def test_binning_implementations_agree_in_process_and_are_deterministic(built):
    """Random scenes / image shapes: the two-level (mode 3), the single-level (mode 2) sort-free binning and the
    automatic choice (mode 0) equal the instance radix sort (mode 1) bit for bit, and 10 repeats of each give identical lists (a shared-memory race in
    the scatter kernels would show up as run-to-run variation)."""
    lib = _abi.load()
    cases = [(20_000, "tum_half", 21, False, False), (80_000, "tum", 22, True, True), (150_000, "replica", 23, False, False),
             (257, "tum_half", 24, False, False), (40_000, "tum", 25, False, True), (300_000, "hd", 26, True, True),
             (3_000, "hd", 27, False, False)]
    try:
        for P, kind, seed, aniso, layered in cases:
            cam = synth.make_camera(kind, rot=(0.003, 0.004, -0.002), trans=(0.01, 0.02, -0.01))
            sc = synth.make_scene(P, cam, seed=seed, anisotropic=aniso, layered=layered, device=DEV)
            bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
            fa = _fargs(cam, sc, bg, debug=False)
            outs = {}
            modes = (1, 3, 0, 2) if cam.W * cam.H <= 4096 * 256 else (1, 3, 0)
            for mode in modes:
                assert lib.gsr_debug_set_binning(mode) == 0
                reps = []
                for _ in range(10 if mode != 1 else 1):
                    R, col, radii, geom, plist, img, dep, opa, nt = ours.rasterize_gaussians(*fa)
                    # (sorted keys are compared in the two-phase tests above: the key hook follows gsr_forward_render)
                    _, n_contrib, ranges = _image_state(img, cam.W, cam.H)
                    torch.cuda.synchronize()
                    reps.append((R, plist[:R].clone(), None, ranges.copy(), n_contrib.copy(), col.clone()))
                for r in reps[1:]:
                    assert r[0] == reps[0][0] and torch.equal(r[1], reps[0][1]) and torch.equal(r[5], reps[0][5])
                outs[mode] = reps[0]
            a = outs[1]
            for mode in modes[1:]:
                b = outs[mode]
                assert a[0] == b[0] > 0, mode
                assert torch.equal(a[1], b[1]) and torch.equal(a[5], b[5]), mode
                assert (a[3] == b[3]).all() and (a[4] == b[4]).all(), mode
    finally:
        lib.gsr_debug_set_binning(0)
