"""tools/compare_gpu.py -- run our rasterizer and the reference CUDA rasterizer (oracle/_ref) on the same
synthetic scene on a B200, compare every output, and time both. Development tool (uses oracle/: test infra)."""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth  # noqa: E402
from monogs_b200.diff_gaussian_rasterization import _C as ours  # noqa: E402
from oracle import ref_cuda as ref  # noqa: E402


def rel(a, b, floor=1e-12):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / max(float(b.abs().max()), floor))


def run_pair(P, kind, seed, anisotropic, layered, timing_iters, log):
    dev = "cuda"
    cam = synth.make_camera(kind, rot=(0.004, -0.006, 0.003), trans=(0.02, -0.01, 0.015))
    sc = synth.make_scene(P, cam, seed=seed, anisotropic=anisotropic, layered=layered, device=dev)
    view, proj, proj_raw, campos = cam.matrices(dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    e = torch.empty(0, device=dev)
    fargs = (bg, sc["means3D"], sc["colors"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, proj,
             proj_raw, cam.tanfovx, cam.tanfovy, cam.H, cam.W, e, 0, campos, False, True)
    g = torch.Generator(device="cpu").manual_seed(seed + 7)
    dcol = torch.randn(3, cam.H, cam.W, generator=g).to(dev) / (cam.H * cam.W)
    ddep = torch.randn(1, cam.H, cam.W, generator=g).to(dev) / (cam.H * cam.W)

    def bargs(radii, R, geom, plist, img):
        return (bg, sc["means3D"], radii, sc["colors"], sc["scales"], sc["rotations"], 1.0, e, view, proj, proj_raw,
                cam.tanfovx, cam.tanfovy, dcol, ddep, e, 0, campos, geom, R, plist, img, False)

    res = {"P": P, "kind": kind, "seed": seed, "anisotropic": anisotropic, "layered": layered}
    # ---- ours ----
    Ro, col_o, rad_o, geom, plist, img, dep_o, opa_o, nt_o = ours.rasterize_gaussians(*fargs)
    torch.cuda.synchronize()
    go = ours.rasterize_gaussians_backward(*bargs(rad_o, Ro, geom, plist, img))
    _, tau_o = ours.rasterize_gaussians_backward_fused(*bargs(rad_o, Ro, geom, plist, img))
    torch.cuda.synchronize()
    # ---- reference ----
    Rr, col_r, rad_r, _, _, _, dep_r, opa_r, nt_r = ref.rasterize_gaussians(*fargs)
    torch.cuda.synchronize()
    st = ref.last_state(P, Rr, cam.W, cam.H, dev)
    gr = ref.rasterize_gaussians_backward(*bargs(rad_r, Rr, None, None, None))
    torch.cuda.synchronize()

    res["R"] = [Ro, Rr]
    res["V"] = int((rad_r > 0).sum())
    res["radii_mismatch"] = int((rad_o != rad_r).sum())
    res["n_touched_mismatch"] = int((nt_o != nt_r).sum())
    if Ro == Rr:
        pl_o = plist.view(torch.int32)[:Ro]
        res["point_list_mismatch"] = int((pl_o != st["point_list"]).sum())
        import ctypes as C
        from monogs_b200 import _abi
        lib = _abi.load()
        keys_o = ours.debug_sorted_keys(geom, ours.rasterize_gaussians.last_binning, plist, P, Ro, cam.W, cam.H)
        res["keys_mismatch"] = int((keys_o != st["point_list_keys"]).sum())
        rp = C.c_void_p(); ncp = C.c_void_p(); ftp = C.c_void_p()
        lib.gsr_debug_image_views(img.data_ptr(), cam.W, cam.H, C.byref(ftp), C.byref(ncp), C.byref(rp))
        tiles = ((cam.W + 15) // 16) * ((cam.H + 15) // 16)
        o = rp.value - img.data_ptr()
        res["ranges_mismatch"] = int((img[o:o + 8 * tiles].view(torch.int32) != st["ranges"]).sum())
        o = ncp.value - img.data_ptr()
        res["n_contrib_mismatch"] = int((img[o:o + 4 * cam.W * cam.H].view(torch.int32) != st["n_contrib"]).sum())
        o = ftp.value - img.data_ptr()
        res["final_T_bits_mismatch"] = int((img[o:o + 4 * cam.W * cam.H].view(torch.int32) !=
                                            st["accum_alpha"].view(torch.int32)).sum())
    # per-Gaussian geometry bits (visible Gaussians only: the reference leaves the others uninitialised)
    import ctypes as C2
    from monogs_b200 import _abi as _abi2
    lib2 = _abi2.load()
    pa, pb, pc, pt = C2.c_void_p(), C2.c_void_p(), C2.c_void_p(), C2.c_void_p()
    lib2.gsr_debug_geom_views(geom.data_ptr(), P, cam.W, cam.H, C2.byref(pa), C2.byref(pb), C2.byref(pc), C2.byref(pt))
    def gview(ptr, n):
        o = ptr.value - geom.data_ptr()
        return geom[o:o + 4 * n].view(torch.int32)
    vis = rad_r > 0
    rec = gview(pa, 12 * P).view(P, 12)   # one packed 48-byte record per Gaussian
    ra = rec[:, 0:4]; rb = rec[:, 4:8]
    res["means2D_bits_mismatch"] = int((ra[:, :2][vis] != st["means2D"].view(torch.int32).view(P, 2)[vis]).sum())
    res["depth_bits_mismatch"] = int((ra[:, 2][vis] != st["depths"].view(torch.int32)[vis]).sum())
    res["conic_bits_mismatch"] = int((rb[vis] != st["conic_opacity"].view(torch.int32).view(P, 4)[vis]).sum())
    res["tiles_touched_mismatch"] = int((gview(pt, P) != st["tiles_touched"]).sum())
    res["color_bits_mismatch"] = int((col_o.view(torch.int32) != col_r.view(torch.int32)).sum())
    res["color_rel"] = rel(col_o, col_r)
    res["depth_rel"] = rel(dep_o, dep_r)
    res["opacity_rel"] = rel(opa_o, opa_r)
    names = ["dmeans2D", "dcolors", "dopacity", "dmeans3D", "dcov3D", "dsh", "dscales", "drots", "dtau_pg"]
    for n, a, b in zip(names, go, gr):
        if a.numel():
            res["grad_rel_" + n] = rel(a, b)
    tau_r = gr[8].double().sum(0)
    res["tau_ours"] = tau_o.tolist()
    res["tau_ref"] = tau_r.tolist()
    res["tau_rel"] = float((tau_o.double() - tau_r).abs().max() / max(float(tau_r.abs().max()), 1e-30))
    # reference run-to-run spread (atomics): second backward
    gr2 = ref.rasterize_gaussians_backward(*bargs(rad_r, Rr, None, None, None))
    res["ref_self_rel_dmeans3D"] = rel(gr2[3], gr[3])

    # ---- timing ----
    def timeit(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(n):
            fn()
        t1.record(); torch.cuda.synchronize()
        return t0.elapsed_time(t1) / n

    if timing_iters > 0:
        fa = fargs[:-1] + (False,)
        res["ms_fwd_ours"] = timeit(lambda: ours.rasterize_gaussians(*fa), timing_iters)
        res["ms_fwd_ref"] = timeit(lambda: ref.rasterize_gaussians(*fa), timing_iters)
        ours_state = ours.rasterize_gaussians(*fa)
        ba = bargs(ours_state[2], ours_state[0], ours_state[3], ours_state[4], ours_state[5])
        res["ms_bwd_ours"] = timeit(lambda: ours.rasterize_gaussians_backward_fused(*ba), timing_iters)
        rs = ref.rasterize_gaussians(*fa)
        bar = bargs(rs[2], rs[0], None, None, None)
        res["ms_bwd_ref"] = timeit(lambda: ref.rasterize_gaussians_backward(*bar), timing_iters)
        res["speedup_fwd_bwd"] = (res["ms_fwd_ref"] + res["ms_bwd_ref"]) / (res["ms_fwd_ours"] + res["ms_bwd_ours"])
    log(json.dumps(res))
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="small")
    ap.add_argument("--out", default="gpurun_out/compare.jsonl")
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    f = open(a.out, "a")

    def log(s):
        print(s, flush=True)
        f.write(s + "\n"); f.flush()

    cases = {
        "small": [(10_000, "tum_half", 0, False, False), (100_000, "tum", 1, False, False),
                  (50_000, "tum", 2, True, False), (60_000, "tum", 3, False, True)],
        "headline": [(1_000_000, "tum", 0, False, False)],
        "big": [(500_000, "replica", 0, False, False), (4_000_000, "hd", 0, False, False),
                (4_000_000, "tum", 1, False, False), (2_000_000, "replica", 2, True, True)],
    }
    for name in a.cases.split(","):
        for (P, kind, seed, aniso, lay) in cases[name]:
            run_pair(P, kind, seed, aniso, lay, a.iters, log)


if __name__ == "__main__":
    main()
