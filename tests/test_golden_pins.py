"""CPU checks of the golden vectors that pin the loss / statistics rows (a17, f1, f4) to the reference's code.

* tests/golden/loss.npz is re-derived here from the shipped, sha256-verified copy of the reference's
  utils/slam_utils.py (oracle/_ref/callers, built by `make -C oracle callers`) -- the file must reproduce its own
  golden outputs bit for bit;
* tests/golden/stats.npz (produced by executing the reference's Mapper.optimize_map) is checked against the
  formulas the C ABI header states for gsr_densify_stats / gsr_visibility_prune, so header, golden and reference
  say the same thing.
"""
import importlib.util
import types

import numpy as np
import pytest
import torch

import ref_callers
from conftest import ROOT


def _load(name):
    spec = importlib.util.spec_from_file_location(name, ROOT / "tests" / "golden" / f"{name}.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.skipif(not ref_callers.available(), reason="oracle/_ref/callers not built")
@pytest.mark.parametrize("name", ["mapping", "mapping_init", "tracking", "tracking_sparse"])
def test_loss_golden_reproduced_by_the_shipped_reference_file(name):
    assert ref_callers.verify_unmodified() == 7
    spec = importlib.util.spec_from_file_location("shipped_slam_utils", ref_callers.CALLERS / "utils" / "slam_utils.py")
    slam_utils = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(slam_utils)
    mod = _load("make_loss_golden")
    g = np.load(ROOT / "tests" / "golden" / "loss.npz")
    c = mod.case_inputs(name)
    image = c["image"].clone().requires_grad_(True)
    depth = c["depth"].clone().requires_grad_(True)
    opacity = c["opacity"].clone().requires_grad_(True)
    vp = types.SimpleNamespace(rgb=c["gt_rgb"], depth=c["gt_depth"], mask=c["mask"], grad_mask=c["grad_mask"],
                               exposure_a=c["exposure_a"].clone().requires_grad_(True),
                               exposure_b=c["exposure_b"].clone().requires_grad_(True))
    if c["kind"] == "mapping":
        loss = slam_utils.get_loss_mapping(image, depth, vp, init=c["init"], invert_depth=False)
    else:
        loss = slam_utils.get_loss_tracking(image, depth, opacity, vp, invert_depth=False)
    (c["upstream"] * loss).backward()
    assert float(loss.detach()) == float(g[f"{name}/loss"])
    assert np.array_equal(image.grad.numpy(), g[f"{name}/d_image"])
    assert np.array_equal(depth.grad.numpy(), g[f"{name}/d_depth"])


def test_stats_golden_obeys_the_header_formulas():
    mod = _load("make_stats_golden")
    g = np.load(ROOT / "tests" / "golden" / "stats.npz")
    # gsr_densify_stats, applied keyframe after keyframe
    c = mod.case_inputs(1)
    max_r, accum, denom = c["max_radii"].clone(), c["accum"].clone(), c["denom"].clone()
    for radii, grad in zip(c["radii"], c["grads"]):
        vis = radii > 0
        max_r = torch.where(vis, torch.maximum(max_r, radii.float()), max_r)
        accum = accum + vis[:, None] * grad[:, :2].norm(dim=-1, keepdim=True)
        denom = denom + vis[:, None].float()
    assert np.array_equal(max_r.numpy(), g["stats/max_radii_2d"])
    assert np.array_equal(denom.numpy(), g["stats/denom"])
    np.testing.assert_allclose(accum.numpy(), g["stats/xyz_gradient_accum"], rtol=1e-6)
    # gsr_visibility_prune
    for case, seed in (("prune_first", 2), ("prune_later", 3)):
        c = mod.case_inputs(seed)
        nr_obs = sum((t > 0).int() for t in c["n_touched"])
        to_prune = (nr_obs <= 3) & (c["kf_idx"] >= int(g[f"{case}/min_kf"]))
        assert np.array_equal(nr_obs.numpy(), g[f"{case}/nr_obs"])
        assert np.array_equal(to_prune.numpy(), g[f"{case}/to_prune"])
