"""GPU parity of monogs_b200.stats, pinned to the reference's own code: tests/golden/stats.npz holds what
Mapper.optimize_map (utils/slam_mapper.py:399-460) and GaussianModel.add_densification_stats
(gaussian_splatting/scene/gaussian_model.py:888-892) themselves produced on seeded inputs
(tests/golden/make_stats_golden.py executes them on CPU); the formula tests below them stay as a second check at
other sizes."""
import importlib.util

import numpy as np
import pytest
import torch

from conftest import ROOT
from monogs_b200 import stats

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_add_densification_stats_matches_reference_formulas(built):
    g = torch.Generator().manual_seed(5)
    P = 100_003
    radii = torch.randint(-2, 40, (P,), generator=g, dtype=torch.int32).clamp_min(0).to(DEV)
    grad = torch.randn(P, 3, generator=g).to(DEV) * 1e-3
    max_r = (torch.rand(P, generator=g) * 30).to(DEV)
    accum = torch.rand(P, 1, generator=g).to(DEV)
    denom = torch.randint(0, 9, (P, 1), generator=g).float().to(DEV)
    # the reference: boolean-mask updates
    vis = radii > 0
    r_max, r_acc, r_den = max_r.clone(), accum.clone(), denom.clone()
    r_max[vis] = torch.max(r_max[vis], radii[vis])
    r_acc[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
    r_den[vis] += 1
    stats.add_densification_stats(radii, grad, max_r, accum, denom)
    assert torch.equal(max_r, r_max) and torch.equal(denom, r_den)
    assert torch.allclose(accum, r_acc, rtol=1e-6, atol=0)
    assert torch.equal(accum[~vis], r_acc[~vis])
    # radii-only form (no gradient statistics)
    m2 = torch.zeros(P, device=DEV)
    stats.add_densification_stats(radii, None, m2, None, None)
    assert torch.equal(m2, radii.float())


def test_visibility_prune_mask_matches_reference_formulas(built):
    g = torch.Generator().manual_seed(6)
    P, K = 70_001, 8
    touched = [(torch.randint(0, 3, (P,), generator=g, dtype=torch.int32) *
                torch.randint(0, 2, (P,), generator=g, dtype=torch.int32)).to(DEV) for _ in range(K)]
    kf_idx = torch.randint(0, 20, (P,), generator=g, dtype=torch.int32).to(DEV)
    window = [19, 17, 14, 11, 9, 6, 3, 0]
    min_kf = sorted(window, reverse=True)[2]
    nr_obs, to_prune = stats.visibility_prune_mask(touched, kf_idx, min_kf, coviz=3)
    ref_obs = torch.zeros(P, dtype=torch.int32, device=DEV)
    for t in touched:
        ref_obs += (t > 0).int()
    ref_prune = torch.logical_and(ref_obs <= 3, kf_idx >= min_kf)
    assert torch.equal(nr_obs, ref_obs)
    assert to_prune.dtype == torch.bool and torch.equal(to_prune, ref_prune)
    with pytest.raises(ValueError):
        stats.visibility_prune_mask([], kf_idx, 0)


@pytest.fixture(scope="module")
def stats_golden():
    spec = importlib.util.spec_from_file_location("make_stats_golden",
                                                  ROOT / "tests" / "golden" / "make_stats_golden.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, np.load(ROOT / "tests" / "golden" / "stats.npz")


def test_densification_stats_match_reference_golden(built, stats_golden):
    """8 keyframes of one window iteration through gsr_densify_stats == the reference mapper's own updates."""
    mod, g = stats_golden
    c = mod.case_inputs(1)
    max_r, accum, denom = c["max_radii"].to(DEV), c["accum"].to(DEV), c["denom"].to(DEV)
    for radii, grad in zip(c["radii"], c["grads"]):
        stats.add_densification_stats(radii.to(DEV), grad.to(DEV), max_r, accum, denom)
    assert np.array_equal(max_r.cpu().numpy(), g["stats/max_radii_2d"])
    assert np.array_equal(denom.cpu().numpy(), g["stats/denom"])
    # 8 sequential fp32 additions of sqrt(x^2 + y^2): torch.norm and the kernel may round the norm differently
    np.testing.assert_allclose(accum.cpu().numpy(), g["stats/xyz_gradient_accum"], rtol=1e-6, atol=0)
    # occlusion-aware visibility (slam_mapper.py:399-404) = n_touched > 0
    vis = np.stack([(t > 0).numpy() for t in c["n_touched"]]).astype(np.uint8)
    assert np.array_equal(vis, g["stats/occ_visibility"])


@pytest.mark.parametrize("case,seed", [("prune_first", 2), ("prune_later", 3)])
def test_visibility_prune_matches_reference_golden(built, stats_golden, case, seed):
    mod, g = stats_golden
    c = mod.case_inputs(seed)
    nr_obs, to_prune = stats.visibility_prune_mask([t.to(DEV) for t in c["n_touched"]], c["kf_idx"].to(DEV),
                                                   int(g[f"{case}/min_kf"]), coviz=3)
    assert np.array_equal(nr_obs.cpu().numpy(), g[f"{case}/nr_obs"])
    assert np.array_equal(to_prune.cpu().numpy(), g[f"{case}/to_prune"])
