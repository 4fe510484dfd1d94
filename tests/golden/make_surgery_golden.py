"""Golden vectors for the optimizer surgery (row f3, second half): produced by EXECUTING the reference's own
GaussianModel.training_setup / densification_postfix / cat_tensors_to_optimizer / prune_points / _prune_optimizer
(/root/reference/gaussian_splatting/scene/gaussian_model.py:398-449, 642-778) on CPU tensors in this container.
`open3d`, `plyfile` and `simple_knn` (imported by that module, unused on these paths) are stubbed.

    python tests/golden/make_surgery_golden.py        -> tests/golden/surgery.npz

Scenario (the order MonoGS runs them in: slam.py:87-88, then keyframe insertions and prunes):
  setup      empty model, training_setup (Adam over 5 one-parameter groups, no state yet)
  A          first insertion of 1200 Gaussians (no Adam state: only the parameters grow)
  --         Adam state installed (step = 7, seeded moments), statistics set
  B          insertion of 700 more (moments extended by zero rows, step kept, statistics reset to zero)
  --         statistics set again
  C          prune_points of a seeded 30 % mask (parameters, moments, statistics, _obj_prob, kf_idx, nr_obs compacted)
  D          insertion of 3000 (outgrows a 4096-row capacity)
`scenario(model, ops)` below is shared with tests/test_gpu_surgery.py, which runs the same steps through
monogs_b200.optimizer_surgery on the GPU and compares every tensor with these files bit for bit.
"""
import sys
import types
from pathlib import Path

import numpy as np
import torch
from torch import nn

ROOT = Path(__file__).resolve().parents[2]
REF = Path("/root/reference")
NR_OBJECTS = 5
GROUP_COLS = dict(xyz=3, f_dc=3, opacity=1, scaling=3, rotation=4)
ATTRS = dict(xyz="_xyz", f_dc="_features_dc", opacity="_opacity", scaling="_scaling", rotation="_rotation")


def rows(first, n, cols, tag, device):
    """Distinct, exactly representable values: row id * 16 + column + tag / 64."""
    r = torch.arange(first, first + n, dtype=torch.float32).view(n, 1) * 16.0
    c = torch.arange(cols, dtype=torch.float32).view(1, cols)
    return (r + c + tag / 64.0).to(device)


def new_gaussians(first, n, kf, device):
    g = torch.Generator().manual_seed(1000 + first)
    obj = torch.zeros(n, NR_OBJECTS)
    obj[torch.arange(n), torch.randint(0, NR_OBJECTS, (n,), generator=g)] = 1
    return dict(new_xyz=nn.Parameter(rows(first, n, 3, 1, device)), new_features_dc=nn.Parameter(rows(first, n, 3, 2, device)),
                new_opacities=nn.Parameter(rows(first, n, 1, 3, device)), new_scaling=nn.Parameter(rows(first, n, 3, 4, device)),
                new_rotation=nn.Parameter(rows(first, n, 4, 5, device)),
                new_obj_prob=nn.Parameter(obj.to(device).requires_grad_(False)),
                new_kf_idxs=torch.ones(n).int() * kf, new_nr_obs=torch.zeros(n).int())


def install_adam_state(model, device):
    for g in model.optimizer.param_groups:
        p = g["params"][0]
        n, c = p.shape[0], GROUP_COLS[g["name"]]
        model.optimizer.state[p] = {"step": torch.tensor(7.0), "exp_avg": rows(0, n, c, 10, device) * -1.0,
                                    "exp_avg_sq": rows(0, n, c, 20, device)}


def set_stats(model, device, tag):
    n = model._xyz.shape[0]
    model.xyz_gradient_accum = rows(0, n, 1, 30 + tag, device)
    model.denom = rows(0, n, 1, 40 + tag, device)
    model.max_radii_2d = rows(0, n, 1, 50 + tag, device).view(n)
    model.nr_obs = (torch.arange(n) % 7).int()


def snapshot(model, out, name):
    for g in model.optimizer.param_groups:
        p = g["params"][0]
        assert p is getattr(model, ATTRS[g["name"]]) and isinstance(p, nn.Parameter) and p.requires_grad
        out[f"{name}/{g['name']}"] = p.detach().cpu().numpy()
        st = model.optimizer.state.get(p, None)
        if st is not None:
            out[f"{name}/{g['name']}/step"] = np.float32(float(st["step"]))
            out[f"{name}/{g['name']}/exp_avg"] = st["exp_avg"].detach().cpu().numpy()
            out[f"{name}/{g['name']}/exp_avg_sq"] = st["exp_avg_sq"].detach().cpu().numpy()
    assert len(model.optimizer.state) in (0, 5)
    for k in ("_obj_prob", "xyz_gradient_accum", "denom", "max_radii_2d", "kf_idx", "nr_obs"):
        out[f"{name}/{k}"] = getattr(model, k).detach().cpu().numpy()


def prune_mask(n, device):
    g = torch.Generator().manual_seed(77)
    return (torch.rand(n, generator=g) < 0.3).to(device)


def scenario(model, device):
    """Runs the steps on `model` (whose surgery methods are the implementation under test) -> dict of snapshots."""
    out = {}
    model.densification_postfix(**new_gaussians(0, 1200, 0, device))
    snapshot(model, out, "A")
    install_adam_state(model, device)
    set_stats(model, device, 0)
    model.densification_postfix(**new_gaussians(1200, 700, 3, device))
    snapshot(model, out, "B")
    set_stats(model, device, 1)
    model.prune_points(prune_mask(model._xyz.shape[0], device))
    snapshot(model, out, "C")
    model.densification_postfix(**new_gaussians(5000, 3000, 9, device))
    snapshot(model, out, "D")
    return out


def training_args():
    return types.SimpleNamespace(percent_dense=0.01, position_lr_init=0.0016, position_lr_final=0.0000016,
                                 position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025,
                                 opacity_lr=0.05, scaling_lr=0.001, rotation_lr=0.001)


def stub_modules():
    for name, attrs in (("open3d", {}), ("plyfile", dict(PlyData=object, PlyElement=object)), ("simple_knn", {}),
                        ("simple_knn._C", dict(distCUDA2=None))):
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules.setdefault(name, m)


def main():
    sys.path.insert(0, str(REF))
    stub_modules()
    from gaussian_splatting.scene.gaussian_model import GaussianModel   # the reference's own code
    model = GaussianModel(types.SimpleNamespace(isotropic=False), NR_OBJECTS, "cpu")
    model.init_lr(6.0)
    model.training_setup(training_args())
    out = scenario(model, "cpu")
    for k, v in out.items():
        print(k, v.shape if hasattr(v, "shape") else v)
    np.savez_compressed(ROOT / "tests" / "golden" / "surgery.npz", **out)


if __name__ == "__main__":
    main()
