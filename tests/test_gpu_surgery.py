"""Optimizer surgery (monogs_b200.optimizer_surgery: gsr_rows_append / gsr_rows_compact behind the C ABI) against
golden vectors the REFERENCE's GaussianModel methods produced (tests/golden/make_surgery_golden.py runs
training_setup / densification_postfix / prune_points of gaussian_splatting/scene/gaussian_model.py:398-778 on CPU).
Pure data movement: every tensor must match bit for bit."""
import importlib.util
import types

import numpy as np
import pytest
import torch
from torch import nn

from conftest import ROOT
from monogs_b200 import optimizer_surgery as surgery

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _golden_mod():
    spec = importlib.util.spec_from_file_location("make_surgery_golden", ROOT / "tests" / "golden" / "make_surgery_golden.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class Model:
    """The state GaussianModel.__init__ + training_setup leave behind (gaussian_model.py:34-70, 398-437); the four
    surgery methods are this package's."""

    def __init__(self, args, device):
        self.device = device
        e = lambda: torch.empty(0, device=device)
        self._xyz, self._features_dc, self._scaling, self._rotation, self._opacity = e(), e(), e(), e(), e()
        self._obj_prob, self.max_radii_2d, self.xyz_gradient_accum = e(), e(), e()
        self.kf_idx = torch.empty(0).int()
        self.nr_obs = torch.empty(0).int()
        self.denom = torch.zeros((0, 1), device=device)
        groups = [{"params": [self._xyz], "lr": args.position_lr_init * 6.0, "name": "xyz"},
                  {"params": [self._features_dc], "lr": args.feature_lr, "name": "f_dc"},
                  {"params": [self._opacity], "lr": args.opacity_lr, "name": "opacity"},
                  {"params": [self._scaling], "lr": args.scaling_lr * 6.0, "name": "scaling"},
                  {"params": [self._rotation], "lr": args.rotation_lr, "name": "rotation"}]
        self.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)

    @property
    def get_xyz(self):
        return self._xyz


def test_surgery_matches_reference_golden(built):
    mod = _golden_mod()
    g = np.load(ROOT / "tests" / "golden" / "surgery.npz")
    model = surgery.install(Model(mod.training_args(), DEV))
    out = mod.scenario(model, DEV)
    assert set(out) == set(g.files)
    for k in g.files:
        a, b = np.asarray(out[k]), g[k]
        assert a.shape == b.shape and a.dtype == b.dtype, (k, a.shape, b.shape, a.dtype, b.dtype)
        assert np.array_equal(a, b), k
    # the Adam state keeps working on the new parameter objects: one step changes exactly the rows with a gradient
    for grp in model.optimizer.param_groups:
        p = grp["params"][0]
        p.grad = torch.zeros_like(p)
        p.grad[:10] = 1.0
    before = model._xyz.detach().clone()
    model.optimizer.step()
    assert not torch.equal(model._xyz[:10], before[:10])


def test_append_is_in_place_and_compaction_equals_boolean_indexing(built):
    """The semantics in torch terms at a larger size: append == torch.cat (old rows not moved while capacity lasts),
    compaction == t[mask] for every array, for ragged sizes and empty / full masks."""
    mod = _golden_mod()
    gen = torch.Generator().manual_seed(3)
    model = surgery.install(Model(mod.training_args(), DEV))
    P0 = 100_003

    def fresh(n, kf):
        r = lambda c: nn.Parameter(torch.randn(n, c, generator=gen).to(DEV))
        obj = torch.zeros(n, mod.NR_OBJECTS, device=DEV)
        return dict(new_xyz=r(3), new_features_dc=r(3), new_opacities=r(1), new_scaling=r(3), new_rotation=r(4),
                    new_obj_prob=obj, new_kf_idxs=torch.ones(n).int() * kf, new_nr_obs=torch.zeros(n).int())

    first = fresh(P0, 0)
    model.densification_postfix(**first)
    mod.install_adam_state(model, DEV)
    ptr = model._xyz.data_ptr()
    expect = {k: getattr(model, a).detach().clone() for k, a in mod.ATTRS.items()}
    more = fresh(2_500, 1)
    model.densification_postfix(**more)
    assert model._xyz.shape[0] == P0 + 2_500
    assert model._xyz.data_ptr() == ptr                                  # capacity 1.5x: the old rows were not moved
    for k, a, key in (("xyz", "_xyz", "new_xyz"), ("rotation", "_rotation", "new_rotation")):
        assert torch.equal(getattr(model, a), torch.cat((expect[k], more[key].detach())))
    st = model.optimizer.state[model._xyz]
    assert float(st["exp_avg"][P0:].abs().max()) == 0 and float(st["exp_avg_sq"][P0:].abs().max()) == 0
    assert float(st["step"]) == 7.0
    for frac in (0.37, 0.0, 1.0):
        n = model._xyz.shape[0]
        mask = torch.rand(n, generator=gen).to(DEV) < frac
        mod.set_stats(model, DEV, 2)
        want = {a: getattr(model, a).detach()[~mask].clone() for a in
                ("_xyz", "_features_dc", "_opacity", "_scaling", "_rotation", "_obj_prob", "xyz_gradient_accum",
                 "denom", "max_radii_2d")}
        want_m = model.optimizer.state[model._rotation]["exp_avg_sq"][~mask].clone()
        want_kf = model.kf_idx[(~mask).cpu()].clone()
        model.prune_points(mask)
        for a, w in want.items():
            assert torch.equal(getattr(model, a).detach(), w), (a, frac)
        assert torch.equal(model.optimizer.state[model._rotation]["exp_avg_sq"], want_m)
        assert torch.equal(model.kf_idx, want_kf) and model.nr_obs.shape[0] == model._xyz.shape[0]
        if model._xyz.shape[0] == 0:
            break
