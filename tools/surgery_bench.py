"""tools/surgery_bench.py -- keyframe-insertion append and prune of the map state: monogs_b200.optimizer_surgery against the
reference's op sequence (gaussian_model.py:642-778: torch.cat per tensor / boolean-mask indexing per tensor) written
with the same torch ops on the same GPU. Wall clock incl. host syncs, median of 20. One JSON line."""
import json, sys, time, types
from pathlib import Path
import torch
from torch import nn
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from monogs_b200 import optimizer_surgery as surgery
from test_gpu_surgery import Model, _golden_mod

DEV = "cuda"
mod = _golden_mod()


def ref_cat(model, d, obj, kf=None, obs=None):    # cat_tensors_to_optimizer + densification_postfix, reference ops
    for g in model.optimizer.param_groups:
        ext = d[g["name"]]
        p = g["params"][0]
        st = model.optimizer.state.get(p, None)
        if st is not None:
            st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)), dim=0)
            st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
            del model.optimizer.state[p]
            g["params"][0] = nn.Parameter(torch.cat((p, ext), dim=0).requires_grad_(True))
            model.optimizer.state[g["params"][0]] = st
        else:
            g["params"][0] = nn.Parameter(torch.cat((p, ext), dim=0).requires_grad_(True))
        setattr(model, mod.ATTRS[g["name"]], g["params"][0])
    model._obj_prob = torch.cat((model._obj_prob, obj))
    n = model._xyz.shape[0]
    model.xyz_gradient_accum = torch.zeros((n, 1), device=DEV); model.denom = torch.zeros((n, 1), device=DEV)
    model.max_radii_2d = torch.zeros(n, device=DEV)
    if kf is not None:
        model.kf_idx = torch.cat((model.kf_idx, kf)).int()
        model.nr_obs = torch.cat((model.nr_obs, obs)).int()


def ref_prune(model, mask):                       # prune_points, reference ops
    valid = ~mask
    for g in model.optimizer.param_groups:
        p = g["params"][0]
        st = model.optimizer.state.get(p, None)
        if st is not None:
            st["exp_avg"] = st["exp_avg"][valid]; st["exp_avg_sq"] = st["exp_avg_sq"][valid]
            del model.optimizer.state[p]
            g["params"][0] = nn.Parameter(p[valid].requires_grad_(True))
            model.optimizer.state[g["params"][0]] = st
        else:
            g["params"][0] = nn.Parameter(p[valid].requires_grad_(True))
        setattr(model, mod.ATTRS[g["name"]], g["params"][0])
    model.xyz_gradient_accum = model.xyz_gradient_accum[valid]; model._obj_prob = model._obj_prob[valid]
    model.denom = model.denom[valid]; model.max_radii_2d = model.max_radii_2d[valid]
    model.kf_idx = model.kf_idx[valid.cpu()]; model.nr_obs = model.nr_obs[valid.cpu()]


def build(P, ours):
    m = Model(mod.training_args(), DEV)
    if ours:
        surgery.install(m)
    g = torch.Generator().manual_seed(0)
    d = dict(xyz=3, f_dc=3, opacity=1, scaling=3, rotation=4)
    first = {k: nn.Parameter(torch.randn(P, c, generator=g).to(DEV)) for k, c in d.items()}
    obj = torch.zeros(P, mod.NR_OBJECTS, device=DEV)
    if ours:
        m.densification_postfix(first["xyz"], first["f_dc"], first["opacity"], first["scaling"], first["rotation"], obj,
                                torch.zeros(P).int(), torch.zeros(P).int())
    else:
        ref_cat(m, first, obj, torch.zeros(P).int(), torch.zeros(P).int())
    mod.install_adam_state(m, DEV)
    return m


def med(f, n=20):
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); f(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return sorted(ts)[len(ts) // 2] * 1e3


def main():
    P, K = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 20_000
    g = torch.Generator().manual_seed(1)
    d = dict(xyz=3, f_dc=3, opacity=1, scaling=3, rotation=4)
    ext = {k: nn.Parameter(torch.randn(K, c, generator=g).to(DEV)) for k, c in d.items()}
    obj = torch.zeros(K, mod.NR_OBJECTS, device=DEV)
    res = {"P": P, "K": K}
    for who in ("ours", "reference_ops"):
        m = build(P, who == "ours")
        mask = (torch.rand(m._xyz.shape[0] + K, generator=g) < 0.02).to(DEV)

        def cycle():   # one insertion of K followed by a prune back to ~P: the map size stays put across repeats
            if who == "ours":
                m.densification_postfix(ext["xyz"], ext["f_dc"], ext["opacity"], ext["scaling"], ext["rotation"], obj,
                                        torch.zeros(K).int(), torch.zeros(K).int())
            else:
                ref_cat(m, ext, obj, torch.zeros(K).int(), torch.zeros(K).int())
        def prune():
            n = m._xyz.shape[0]
            mk = torch.zeros(n, dtype=torch.bool, device=DEV); mk[n - K:] = True     # drop what was just inserted
            if who == "ours":
                m.prune_points(mk)
            else:
                ref_prune(m, mk)
        ta, tp = [], []
        for _ in range(12):
            ta.append(med(cycle, 1)); tp.append(med(prune, 1))
        res[who] = {"append_ms": sorted(ta)[len(ta) // 2], "prune_ms": sorted(tp)[len(tp) // 2]}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
