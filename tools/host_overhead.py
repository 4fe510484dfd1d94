"""tools/host_overhead.py -- wall clock vs device time of the public API (GaussianRasterizer + backward) with
device-resident inputs, and a cProfile of the host side (host overhead per render call)."""
import cProfile, io, pstats, sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth
from monogs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

def main(P=100_000, iters=300):
    dev = "cuda"
    cam = synth.make_camera("tum", rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=0, device=dev)
    view, proj, proj_raw, campos = cam.matrices(dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    s = GaussianRasterizationSettings(image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
                                      bg=bg, scale_modifier=1.0, viewmatrix=view, projmatrix=proj,
                                      projmatrix_raw=proj_raw, sh_degree=0, campos=campos, prefiltered=False,
                                      debug=False)
    rast = GaussianRasterizer(s)
    theta = torch.zeros(1, 3, device=dev, requires_grad=True)
    rho = torch.zeros(1, 3, device=dev, requires_grad=True)
    m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
    w = torch.rand(3, cam.H, cam.W, device=dev)
    def step():
        color, radii, depth, opacity, n_touched = rast(means3D=sc["means3D"], means2D=m2d, opacities=sc["opacities"],
                                                       colors_precomp=sc["colors"], scales=sc["scales"],
                                                       rotations=sc["rotations"], theta=theta, rho=rho)
        ((color * w).sum() + depth.sum()).backward()
    for _ in range(20): step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(iters): step()
    b.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"P={P}: wall {1e3*(t1-t0)/iters:.3f} ms/iter, device span {a.elapsed_time(b)/iters:.3f} ms/iter")
    # host-only cost: enqueue without waiting
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(iters): step()
    t1 = time.perf_counter(); torch.cuda.synchronize()
    print(f"host enqueue time {1e3*(t1-t0)/iters:.3f} ms/iter")
    pr = cProfile.Profile(); pr.enable()
    for _ in range(iters): step()
    pr.disable(); torch.cuda.synchronize()
    st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("tottime").print_stats(22); print(st.getvalue()[:5000])

if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 100_000)
