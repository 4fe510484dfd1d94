"""Gaussian insertion (gaussian_model.py:121-319 create_viewpoint_pcd).
CPU: the numpy oracle against the golden vectors = outputs of the reference's own function
(tests/golden/make_insert_golden.py).  GPU: monogs_b200.insert (C ABI kernels) against the same golden vectors and
against the oracle at a larger image; medians by radix select against torch.median.
Tolerances: indices / ids / counts exact; positions and features 1e-6 relative to the scene scale (matmul order);
log-scales 1e-5 absolute (they go through the kNN distances of those positions)."""
import types

import numpy as np
import pytest
import torch

from conftest import ROOT
from oracle import insert_oracle

GOLD = np.load(ROOT / "tests" / "golden" / "insert.npz")
CASES = ("init", "extend", "extend_aniso")
NAMES = ("points_3d", "features", "scales", "rots", "opacities", "points_ids")


def _case(c):
    inp = {k.split("/")[-1]: GOLD[k] for k in GOLD.files if k.startswith(c + "/in/")}
    out = {k.split("/")[-1]: GOLD[k] for k in GOLD.files if k.startswith(c + "/out/")}
    return inp, out, int(GOLD[c + "/seed"]), bool(GOLD[c + "/init"]), bool(GOLD[c + "/isotropic"])


def _draw(seed, n_selected, init):
    torch.manual_seed(seed)
    keep = int(n_selected * (1.0 / (32 if init else 64)))
    return torch.randperm(n_selected)[:keep]


def _check(got, out, atol_scale=1e-5):
    scale = np.abs(out["points_3d"]).max()
    assert got[0].shape == out["points_3d"].shape
    assert np.abs(got[0] - out["points_3d"]).max() <= 1e-6 * scale
    assert np.abs(got[1] - out["features"]).max() <= 1e-6
    assert np.abs(got[2] - out["scales"]).max() <= atol_scale
    assert (got[3] == out["rots"]).all() and (got[4] == out["opacities"]).all()
    assert (got[5] == out["points_ids"]).all()


@pytest.mark.parametrize("c", CASES)
def test_oracle_matches_reference_golden(c):
    inp, out, seed, init, iso = _case(c)
    sel, _ = insert_oracle.select_mask(inp["depth"], inp["render_depth"], inp["render_opacity"], init)
    pick = _draw(seed, len(sel), init).numpy()
    assert len(pick) == out["points_3d"].shape[0]
    got = insert_oracle.create_viewpoint_pcd(inp["rgb"], inp["depth"], inp["segmentation"], inp["R"], inp["T"],
                                             float(inp["fx"]), float(inp["fy"]), float(inp["cx"]), float(inp["cy"]),
                                             pick, inp["render_depth"], inp["render_opacity"], inp["exposure_a"],
                                             inp["exposure_b"], init, iso)
    _check(got, out)


def _viewpoint(inp, dev):
    t = lambda a: torch.from_numpy(np.asarray(a)).to(dev)
    vp = types.SimpleNamespace(rgb=t(inp["rgb"]), depth=t(inp["depth"]), segmentation=t(inp["segmentation"]),
                               R=t(inp["R"]), T=t(inp["T"]), exposure_a=t(inp["exposure_a"]),
                               exposure_b=t(inp["exposure_b"]))
    intr = types.SimpleNamespace(fx=float(inp["fx"]), fy=float(inp["fy"]), cx=float(inp["cx"]), cy=float(inp["cy"]))
    return vp, intr


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES)
def test_cuda_matches_reference_golden(c):
    from monogs_b200 import insert
    inp, out, seed, init, iso = _case(c)
    vp, intr = _viewpoint(inp, "cuda")
    torch.manual_seed(seed)  # the same host randperm draw as the reference run that made the golden file
    got = insert.create_viewpoint_pcd(vp, intr, None if init else torch.from_numpy(inp["render_depth"]).cuda(),
                                      None if init else torch.from_numpy(inp["render_opacity"]).cuda(), init=init,
                                      isotropic=iso)
    _check([g.cpu().numpy() for g in got], out)


@pytest.mark.gpu
def test_cuda_matches_oracle_full_size():
    """640x480 (the TUM resolution), both branches, vs the numpy oracle on the same draw."""
    from monogs_b200 import insert
    import importlib.util
    spec = importlib.util.spec_from_file_location("mk", ROOT / "tests" / "golden" / "make_insert_golden.py")
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    for init in (True, False):
        inp = {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in mk.make_inputs(11, 640, 480).items()}
        sel, med_err = insert_oracle.select_mask(inp["depth"], inp["render_depth"], inp["render_opacity"], init)
        pick = _draw(5, len(sel), init)
        vp, intr = _viewpoint(inp, "cuda")
        got = insert.create_viewpoint_pcd(vp, intr, torch.from_numpy(inp["render_depth"]).cuda(),
                                          torch.from_numpy(inp["render_opacity"]).cuda(), init=init,
                                          random_indices=pick)
        ref = insert_oracle.create_viewpoint_pcd(inp["rgb"], inp["depth"], inp["segmentation"], inp["R"], inp["T"],
                                                 inp["fx"], inp["fy"], inp["cx"], inp["cy"], pick.numpy(),
                                                 inp["render_depth"], inp["render_opacity"], inp["exposure_a"],
                                                 inp["exposure_b"], init, True)
        _check([g.cpu().numpy() for g in got], dict(zip(NAMES, ref)), atol_scale=2e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 255, 4097, 307200, 1000003])
def test_radix_select_median_matches_torch(n):
    from monogs_b200 import insert
    g = torch.Generator().manual_seed(n)
    a = (torch.randn(n, generator=g) * 3).cuda()
    b = torch.randn(n, generator=g).cuda()
    a[::7] = 0.0                                      # many ties (invalid-depth zeros)
    assert float(insert.median_lower(a)) == float(torch.median(a))
    assert float(insert.median_lower(a, b)) == float(torch.median((a - b).abs()))


def test_insert_rejects_cpu_tensors():
    from monogs_b200 import insert
    inp, *_ = _case("init")
    vp, intr = _viewpoint(inp, "cpu")
    with pytest.raises(RuntimeError):
        insert.create_viewpoint_pcd(vp, intr, init=True)
