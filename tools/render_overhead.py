"""tools/render_overhead.py -- what the cached-camera fast path of monogs_b200.render buys in a tracking-style loop
(config C2: 100 k Gaussians, 640x480, render + L1 loss + backward, pose gradients): wall-clock iterations/s with
  (a) the reference wrapper's per-call host work written out (two FoV `.cpu().item()` reads, projection matrix rebuilt
      from one-element tensor ops, 4x4 assembly, general inverse, bmm -- gaussian_renderer/__init__.py:63-84,
      camera_utils.py:30-49,171-178,224-231) in front of the same rasterizer, and
  (b) monogs_b200.render.render."""
import json, math, sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import render as R
from monogs_b200 import synth
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


class Obj:
    def __init__(self, **kw):
        self.__dict__.update(kw)


def reference_style_render(vp, ci, sc, bg):
    fovx = 2 * torch.atan(ci.width / (2 * ci.fx)).cpu().item()
    fovy = 2 * torch.atan(ci.height / (2 * ci.fy)).cpu().item()
    proj_T = R.projection_matrix_T(ci.fx, ci.fy, ci.cx, ci.cy, ci.width, ci.height)   # rebuilt per call
    w2c = torch.zeros((4, 4), device=vp.R.device); w2c[:3, :3] = vp.R; w2c[:3, 3] = vp.T; w2c[3, 3] = 1.0
    view = w2c.transpose(0, 1)
    full = view.unsqueeze(0).bmm(proj_T.unsqueeze(0)).squeeze(0)
    campos = view.inverse()[3, :3]
    m2d = torch.zeros_like(sc["means3D"], requires_grad=True) + 0
    m2d.retain_grad()
    s = GaussianRasterizationSettings(image_height=ci.height, image_width=ci.width, tanfovx=math.tan(fovx * 0.5),
                                      tanfovy=math.tan(fovy * 0.5), bg=bg, scale_modifier=1.0, viewmatrix=view,
                                      projmatrix=full, projmatrix_raw=proj_T, sh_degree=0, campos=campos,
                                      prefiltered=False, debug=False)
    out = GaussianRasterizer(s)(means3D=sc["means3D"], means2D=m2d, opacities=sc["opacities"],
                                colors_precomp=sc["colors"], scales=sc["scales"], rotations=sc["rotations"],
                                theta=vp.cam_rot_delta, rho=vp.cam_trans_delta)
    return out[0], out[2]


def main(P=100_000, iters=300):
    dev = "cuda"
    cam = synth.make_camera("tum", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=0, device=dev)
    ci = Obj(fx=torch.tensor([cam.fx], device=dev), fy=torch.tensor([cam.fy], device=dev), cx=cam.cx, cy=cam.cy,
             height=cam.H, width=cam.W)
    w2c = cam.w2c.to(dev)
    vp = Obj(R=w2c[:3, :3], T=w2c[:3, 3], cam_rot_delta=torch.zeros(3, device=dev, requires_grad=True),
             cam_trans_delta=torch.zeros(3, device=dev, requires_grad=True))
    bg = torch.zeros(3, device=dev)
    gt = torch.rand(3, cam.H, cam.W, device=dev)
    gtd = torch.rand(1, cam.H, cam.W, device=dev) * 5

    def step_ref():
        img, depth = reference_style_render(vp, ci, sc, bg)
        ((img - gt).abs().mean() + 0.1 * (depth - gtd).abs().mean()).backward()

    def step_fast():
        pkg = R.render(vp, ci, sc["means3D"], sc["rotations"], sc["scales"], sc["opacities"], sc["colors"], bg)
        ((pkg["render"] - gt).abs().mean() + 0.1 * (pkg["depth"] - gtd).abs().mean()).backward()

    res = {"P": P, "iters": iters}
    for name, fn in (("reference_style_host_path", step_ref), ("fast_path", step_fast)):
        for _ in range(20): fn()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(iters): fn()
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        res[name + "_it_s"] = iters / dt
        res[name + "_ms"] = 1e3 * dt / iters
    res["speedup"] = res["fast_path_it_s"] / res["reference_style_host_path_it_s"]
    print(json.dumps(res))


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 100_000)
