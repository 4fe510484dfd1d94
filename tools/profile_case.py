"""tools/profile_case.py -- a few fwd+bwd iterations of one synthetic case (for ncu launch lists / captures)."""
import argparse, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth
from monogs_b200.diff_gaussian_rasterization import _C as ours

ap = argparse.ArgumentParser()
ap.add_argument("--P", type=int, default=1_000_000)
ap.add_argument("--kind", default="tum")
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--impl", default="ours")
a = ap.parse_args()
dev = "cuda"
cam = synth.make_camera(a.kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
sc = synth.make_scene(a.P, cam, seed=0, device=dev)
view, proj, proj_raw, campos = cam.matrices(dev)
bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
e = torch.empty(0, device=dev)
g = torch.Generator().manual_seed(7)
dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(dev)
ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(dev)
mod = ours
if a.impl == "ref":
    from oracle import ref_cuda as mod
for it in range(a.iters):
    R, col, radii, geom, plist, img, dep, opa, nt = mod.rasterize_gaussians(
        bg, sc["means3D"], sc["colors"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, proj, proj_raw,
        cam.tanfovx, cam.tanfovy, cam.H, cam.W, e, 0, campos, False, False)
    args = (bg, sc["means3D"], radii, sc["colors"], sc["scales"], sc["rotations"], 1.0, e, view, proj, proj_raw,
            cam.tanfovx, cam.tanfovy, dcol, ddep, e, 0, campos, geom, R, plist, img, False)
    if a.impl == "ref":
        mod.rasterize_gaussians_backward(*args)
    else:
        mod.rasterize_gaussians_backward_fused(*args)
torch.cuda.synchronize()
print("done R", R)
