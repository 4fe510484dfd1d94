"""tests/golden/make_golden.py -- generate the golden vectors with the UNMODIFIED reference CUDA code.

Run on a GPU box (needs oracle/_ref/libref_dgr.so, built by `make -C oracle ref` where /root/reference
exists):   python tests/golden/make_golden.py gpurun_out/golden
The .npz files it writes are committed under tests/golden/; they pin the CPU oracle (tests/test_oracle_golden.py)
and are compared with our CUDA path (tests/test_gpu_parity.py).  Inputs are NOT stored: they are regenerated
from monogs_b200.synth with the seeds recorded in each file.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from monogs_b200 import synth  # noqa: E402
from oracle import ref_cuda as ref  # noqa: E402

CASES = {
    # name: (P, camera kind, seed, anisotropic, layered, camera rot, camera trans)
    "iso_small": (3000, "tiny", 11, False, False, (0.004, -0.006, 0.003), (0.02, -0.01, 0.015)),
    "aniso_layered": (2500, "tiny", 12, True, True, (-0.01, 0.02, 0.005), (-0.03, 0.02, 0.05)),
}
synth.INTRINSICS.setdefault("tiny", (160, 120, 133.85, 134.8, 80.03, 61.9))


def case_inputs(name, device="cpu"):
    P, kind, seed, aniso, layered, rot, trans = CASES[name]
    cam = synth.make_camera(kind, rot=rot, trans=trans)
    sc = synth.make_scene(P, cam, seed=seed, anisotropic=aniso, layered=layered, device=device)
    g = torch.Generator().manual_seed(seed + 7)
    dcol = (torch.randn(3, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(device)
    ddep = (torch.randn(1, cam.H, cam.W, generator=g) / (cam.H * cam.W)).to(device)
    bg = torch.tensor([0.1, 0.2, 0.3], device=device)
    return cam, sc, bg, dcol, ddep


def knn_inputs(seed=21, P=5000):
    g = torch.Generator().manual_seed(seed)
    pts = torch.randn(P, 3, generator=g) * torch.tensor([2.0, 1.0, 0.5])
    pts[: P // 10] = pts[P // 10: 2 * (P // 10)]  # exact duplicates -> zero distances
    return pts.contiguous()


def main(out_dir):
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    dev = "cuda"
    for name in CASES:
        cam, sc, bg, dcol, ddep = case_inputs(name, dev)
        view, proj, proj_raw, campos = cam.matrices(dev)
        e = torch.empty(0, device=dev)
        R, col, radii, _, _, _, dep, opa, nt = ref.rasterize_gaussians(
            bg, sc["means3D"], sc["colors"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, proj,
            proj_raw, cam.tanfovx, cam.tanfovy, cam.H, cam.W, e, 0, campos, False, True)
        torch.cuda.synchronize()
        P = sc["means3D"].shape[0]
        st = ref.last_state(P, R, cam.W, cam.H, dev)
        grads = ref.rasterize_gaussians_backward(
            bg, sc["means3D"], radii, sc["colors"], sc["scales"], sc["rotations"], 1.0, e, view, proj, proj_raw,
            cam.tanfovx, cam.tanfovy, dcol, ddep, e, 0, campos, None, R, None, None, True)
        torch.cuda.synchronize()
        vis = (radii > 0).cpu().numpy()
        names = ["dmeans2D", "dcolors", "dopacity", "dmeans3D", "dcov3D", "dsh", "dscales", "drots", "dtau"]
        d = dict(num_rendered=np.int64(R), radii=radii.cpu().numpy(), n_touched=nt.cpu().numpy(),
                 color=col.cpu().numpy(), depth=dep.cpu().numpy(), opacity=opa.cpu().numpy(),
                 keys=st["point_list_keys"].cpu().numpy().view(np.uint64), point_list=st["point_list"].cpu().numpy(),
                 ranges=st["ranges"].cpu().numpy().reshape(-1, 2), n_contrib=st["n_contrib"].cpu().numpy(),
                 final_T=st["accum_alpha"].cpu().numpy(), tiles_touched=st["tiles_touched"].cpu().numpy(),
                 means2D=st["means2D"].cpu().numpy().reshape(P, 2) * vis[:, None],
                 depths=st["depths"].cpu().numpy() * vis,
                 conic_opacity=st["conic_opacity"].cpu().numpy().reshape(P, 4) * vis[:, None])
        for n, t in zip(names, grads):
            if t.numel():
                d[n] = t.cpu().numpy()
        np.savez_compressed(out / f"{name}.npz", **d)
        print(name, "R", R, "V", int(vis.sum()))
    pts = knn_inputs().to(dev)
    np.savez_compressed(out / "knn.npz", dist2=ref.distCUDA2(pts).cpu().numpy())
    print("knn done")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
