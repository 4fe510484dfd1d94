"""tools/tracking_bench.py -- tracking-style loop at config C2 (100 k Gaussians, 640x480): iterations per second of
 (a) the CUDA-graph iteration (monogs_b200.tracking.TrackingGraph: camera setup + forward + tracking loss + backward
     to the pose / exposure gradients in one graph launch),
 (b) the eager fused path (render fast-path matrices + fused_loss.render_l1_loss + backward),
 (c) the reference-style loop (render() drop-in + the reference's torch loss + autograd backward).
Wall clock over 200 iterations after warm-up, one host sync at the end (like the tracker, which only reads the loss
for logging), pose perturbed every iteration."""
import json, sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import fused_loss, render as R, synth, tracking
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings

DEV = "cuda"

class Obj:
    def __init__(self, **kw): self.__dict__.update(kw)

def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
    n = 200
    cam = synth.make_camera("tum")
    sc = synth.make_scene(P, cam, seed=0, device=DEV)
    ci = Obj(fx=torch.tensor([535.4], device=DEV), fy=torch.tensor([539.2], device=DEV), cx=320.1, cy=247.6,
             height=cam.H, width=cam.W)
    bg = torch.zeros(3, device=DEV)
    gt_rgb, gt_depth = synth.make_targets(cam, seed=2, device=DEV)
    ea, eb = torch.zeros(1, device=DEV, requires_grad=True), torch.zeros(1, device=DEV, requires_grad=True)
    w2c = cam.w2c.to(DEV)
    R0, T0 = w2c[:3, :3].contiguous(), w2c[:3, 3].contiguous()
    dT = torch.tensor([1e-4, -1e-4, 5e-5], device=DEV)
    out = {"gaussians": P, "image": [cam.W, cam.H], "iterations": n}

    tg = tracking.TrackingGraph(ci, sc["means3D"], sc["rotations"], sc["scales"], sc["opacities"], sc["colors"], bg,
                                gt_rgb, gt_depth, R0=R0, T0=T0)
    def run_graph():
        T = T0.clone()
        for _ in range(n):
            loss, g_rho, g_theta, _ = tg.step(R0, T, ea, eb)
            T = T - 1e-3 * g_rho                      # stand-in for the pose optimiser step (device only)
        return loss
    def run_eager():
        T = T0.clone()
        tanfovx, tanfovy, proj_T = R.camera_constants(ci)
        for _ in range(n):
            view, proj, campos = R.camera_matrices(R0, T, proj_T)
            s = GaussianRasterizationSettings(cam.H, cam.W, tanfovx, tanfovy, bg, 1.0, view, proj, proj_T, 0, campos,
                                              False, False)
            th, rh = torch.zeros(3, device=DEV, requires_grad=True), torch.zeros(3, device=DEV, requires_grad=True)
            loss, *_ = fused_loss.render_l1_loss(s, sc["means3D"], torch.zeros_like(sc["means3D"]), sc["opacities"],
                                                 colors_precomp=sc["colors"], scales=sc["scales"],
                                                 rotations=sc["rotations"], theta=th, rho=rh, gt_rgb=gt_rgb,
                                                 gt_depth=gt_depth, exposure_a=ea, exposure_b=eb, kind="tracking")
            loss.backward()
            T = T - 1e-3 * rh.grad.reshape(3)
        return loss
    def run_reference_style():
        T = T0.clone()
        for _ in range(n):
            vp = Obj(R=R0, T=T, cam_rot_delta=torch.zeros(3, device=DEV, requires_grad=True),
                     cam_trans_delta=torch.zeros(3, device=DEV, requires_grad=True))
            pkg = R.render(vp, ci, sc["means3D"], sc["rotations"], sc["scales"], sc["opacities"], sc["colors"], bg)
            img, depth, opacity = pkg["render"], pkg["depth"], pkg["opacity"]
            # get_loss_tracking, RGB-D branch (utils/slam_utils.py:58-98)
            image_ab = torch.exp(ea) * img + eb
            om = (opacity > 0.99)
            l1 = (opacity * torch.abs(image_ab * om - gt_rgb * om)).mean()
            dm = (gt_depth > 0) & om
            l1d = torch.abs(depth * dm - gt_depth * dm).sum() / dm.sum().clamp_min(1)
            loss = 0.9 * l1 + 0.1 * l1d
            loss.backward()
            T = T - 1e-3 * vp.cam_trans_delta.grad
        return loss
    for name, fn in (("graph", run_graph), ("eager_fused", run_eager), ("reference_style", run_reference_style)):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        out[name + "_it_s"] = n / (time.perf_counter() - t0)
    print(json.dumps(out))

if __name__ == "__main__":
    main()
