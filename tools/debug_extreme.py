"""tools/debug_extreme.py -- where do ours and the reference differ on the huge / near-degenerate scene?"""
import sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1])); sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
import test_gpu_parity as T
from monogs_b200 import synth
from oracle import ref_cuda as ref
DEV = "cuda"
P = 30_000
cam = synth.make_camera("tum_half", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
sc = synth.make_scene(P, cam, seed=77, anisotropic=True, device=DEV)
g = torch.Generator().manual_seed(78)
k = 600
idx = torch.randperm(P, generator=g)[:k].to(DEV)
z = sc["means3D"][idx, 2].abs().clamp_min(0.3)
shape = torch.ones(k, 3)
shape[:200] = torch.tensor([3.0, 1e-3, 1e-3]); shape[200:400] = torch.tensor([2.0, 2.0, 1e-3]); shape[400:] = torch.tensor([1.5, 1.5, 1.5])
sc["scales"][idx] = (shape.to(DEV) * z[:, None] * (0.2 + torch.rand(k, 1, generator=g).to(DEV))).contiguous()
op = sc["opacities"].clone()
op[idx[0:k:3]] = 1.0
op[idx[1:k:3]] = (1.0 / 255.0) * (1.0 + 0.05 * torch.rand(len(idx[1:k:3]), 1, generator=g).to(DEV))
sc["opacities"] = op.contiguous()
bg = torch.tensor([0.1, 0.2, 0.3], device=DEV)
dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(DEV)
o = T._run_ours(cam, sc, bg, dcol, ddep)
fa = o["fa"]
Rr, col_r, rad_r, *_ = ref.rasterize_gaussians(*fa)
gr = ref.rasterize_gaussians_backward(*T._bargs(fa, rad_r, Rr, None, None, None, dcol, ddep))
gr2 = ref.rasterize_gaussians_backward(*T._bargs(fa, rad_r, Rr, None, None, None, dcol, ddep))
special = torch.zeros(P, dtype=torch.int32, device=DEV); special[idx[:200]] = 1; special[idx[200:400]] = 2; special[idx[400:]] = 3
for name, a, r, r2 in zip(T.GRAD_NAMES, o["grads"], gr, gr2):
    if not a.numel(): continue
    a2 = a.reshape(P, -1).double(); r_ = r.reshape(P, -1).double(); rr2 = r2.reshape(P, -1).double()
    d = (a2 - r_).abs().max(1).values; s = (rr2 - r_).abs().max(1).values
    top = torch.topk(d, 5).indices
    print(name, "max|ref|", float(r_.abs().max()), "max diff", float(d.max()), "ref self spread", float(s.max()))
    for i in top.tolist():
        print("   id", i, "class", int(special[i]), "radius", int(rad_r[i]), "opac", float(sc["opacities"][i]), "scales", sc["scales"][i].tolist(),
              "ours", a2[i].tolist()[:4], "ref", r_[i].tolist()[:4], "ref2", rr2[i].tolist()[:4])
