"""GPU parity of monogs_b200.stats against the reference's torch formulas (utils/slam_mapper.py:399-460,
gaussian_splatting/scene/gaussian_model.py:888-892), written out here."""
import pytest
import torch

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
