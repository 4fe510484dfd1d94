"""tools/grad_noise.py -- run-to-run and ours-vs-reference spread of the gradients on the cancellation-dominated scene
(2 M anisotropic layered Gaussians, 1200x680): three backward runs each of ours and of the reference CUDA build."""
import sys, json
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth
from monogs_b200.diff_gaussian_rasterization import _C as ours
from oracle import ref_cuda as ref

def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))

P, kind, seed = (int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000), "replica", 2
dev = "cuda"
cam = synth.make_camera(kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
sc = synth.make_scene(P, cam, seed=seed, anisotropic=True, layered=True, device=dev)
view, proj, proj_raw, campos = cam.matrices(dev)
bg = torch.tensor([0.1, 0.2, 0.3], device=dev); e = torch.empty(0, device=dev)
fargs = (bg, sc["means3D"], sc["colors"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, proj, proj_raw,
         cam.tanfovx, cam.tanfovy, cam.H, cam.W, e, 0, campos, False, False)
g = torch.Generator(device="cpu").manual_seed(seed + 7)
dcol = torch.randn(3, cam.H, cam.W, generator=g).to(dev) / (cam.H * cam.W)
ddep = torch.randn(1, cam.H, cam.W, generator=g).to(dev) / (cam.H * cam.W)
def bargs(radii, R, geom, plist, img):
    return (bg, sc["means3D"], radii, sc["colors"], sc["scales"], sc["rotations"], 1.0, e, view, proj, proj_raw,
            cam.tanfovx, cam.tanfovy, dcol, ddep, e, 0, campos, geom, R, plist, img, False)
Ro, _, rad_o, geom, plist, img, *_ = ours.rasterize_gaussians(*fargs)
Rr, _, rad_r, *_ = ref.rasterize_gaussians(*fargs)
torch.cuda.synchronize()
go = [ours.rasterize_gaussians_backward(*bargs(rad_o, Ro, geom, plist, img)) for _ in range(3)]
gr = [ref.rasterize_gaussians_backward(*bargs(rad_r, Rr, None, None, None)) for _ in range(3)]
torch.cuda.synchronize()
names = ["dmeans2D", "dcolors", "dopacity", "dmeans3D", "dcov3D", "dsh", "dscales", "drots", "dtau_pg"]
out = {}
for i, n in enumerate(names):
    if not go[0][i].numel(): continue
    out[n] = {"ours_vs_ref": [round(rel(go[a][i], gr[b][i]), 9) for a in range(3) for b in range(3)][:5],
              "ref_vs_ref": [round(rel(gr[1][i], gr[0][i]), 9), round(rel(gr[2][i], gr[0][i]), 9)],
              "ours_vs_ours": [round(rel(go[1][i], go[0][i]), 9), round(rel(go[2][i], go[0][i]), 9)]}
print(json.dumps(out))
