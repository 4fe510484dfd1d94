"""tools/insert_bench.py -- keyframe insertion (create_viewpoint_pcd, gaussian_model.py:121-319) on a GPU:
monogs_b200.insert (C ABI kernels + our distCUDA2) vs the reference's op sequence run with torch on the same GPU
(restated op for op below, distCUDA2 = the reference CUDA kNN from oracle/_ref when built, else ours).
Wall clock including the host synchronisations both versions contain, median of 20 after 5 warm-ups.
`--downsample 1` keeps every selected pixel (the reference's `TODO: compute dist2 with all existing points` regime,
:293): the kNN then dominates."""
import argparse
import json
import sys
import time
import types
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import insert  # noqa: E402
from monogs_b200.simple_knn._C import distCUDA2 as ours_knn  # noqa: E402


def reference_ops(vp, intr, render_depth, render_opacity, init, knn, factor):
    """The reference's sequence of framework ops (gaussian_model.py:131-319), on vp.rgb's device."""
    dev = vp.rgb.device
    gt_rgb = vp.rgb if init else torch.clamp(torch.exp(vp.exposure_a) * vp.rgb + vp.exposure_b, 0.0, 1.0)
    gt_rgb = gt_rgb.permute(1, 2, 0)
    gt_depth = vp.depth.unsqueeze(-1)
    gt_seg = vp.segmentation.unsqueeze(-1)
    if not init:
        render_opacity = render_opacity.permute(2, 1, 0).flatten()
        render_depth = render_depth.permute(2, 1, 0).flatten()
    K = torch.tensor([[intr.fx, 0, intr.cx], [0, intr.fy, intr.cy], [0, 0, 1]], dtype=torch.float32, device=dev)
    K_inv = torch.linalg.inv(K)
    c2w = torch.linalg.inv(insert.w2c_from_RT(vp.R, vp.T))
    point_size = min(0.05, 0.01 * torch.median(gt_depth))
    gt_rgb, gt_depth, gt_seg = gt_rgb.permute(1, 0, 2), gt_depth.permute(1, 0, 2), gt_seg.permute(1, 0, 2)
    points_depth, points_ids, points_rgb = gt_depth.flatten(), gt_seg.flatten(), gt_rgb.reshape(-1, 3)
    mask = points_depth >= 1e-3
    if not init:
        med = torch.abs(points_depth - render_depth).median()
        cond = torch.logical_and(points_depth < render_depth, torch.abs(points_depth - render_depth) > 50 * med)
        mask = torch.logical_and(mask, torch.logical_or(render_opacity < 0.5, cond))
    W, H = gt_rgb.shape[0], gt_rgb.shape[1]
    px, py = torch.meshgrid(torch.arange(W, device=dev), torch.arange(H, device=dev), indexing="ij")
    p2d = (torch.stack([px, py], dim=-1).type(torch.int32).float() + 0.5).reshape(-1, 2)
    p2d, points_depth, points_rgb, points_ids = p2d[mask], points_depth[mask], points_rgb[mask], points_ids[mask]
    n = p2d.shape[0]
    idx = torch.randperm(n)[:int(n * (1.0 / factor))]
    p2d, points_depth, points_rgb, points_ids = p2d[idx], points_depth[idx], points_rgb[idx], points_ids[idx]
    aug = torch.cat((p2d, torch.ones((p2d.shape[0], 1), dtype=p2d.dtype, device=dev)), dim=1)[..., None]
    cam = (K_inv @ aug).squeeze(-1) * points_depth[..., None]
    world = (c2w[:3, :3] @ cam.T).T + c2w[:3, -1][None, ...]
    dist2 = torch.clamp_min(knn(world.contiguous()), 0.0000001) * point_size
    scales = torch.log(torch.sqrt(dist2))[..., None]
    rots = torch.zeros((world.shape[0], 4), device=dev)
    rots[:, 0] = 1
    opac = torch.log(0.5 * torch.ones((world.shape[0], 1), device=dev) / (1 - 0.5))
    return world, points_rgb, scales, rots, opac, points_ids


def timed(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", default="1200x680")
    ap.add_argument("--downsample", type=int, default=0, help="0 = the reference's 32 / 64")
    a = ap.parse_args()
    W, H = map(int, a.size.split("x"))
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "mk", Path(__file__).resolve().parents[1] / "tests" / "golden" / "make_insert_golden.py")
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    inp = mk.make_inputs(3, W, H)
    cu = lambda t: t.cuda() if torch.is_tensor(t) else t
    vp = types.SimpleNamespace(rgb=cu(inp["rgb"]), depth=cu(inp["depth"]), segmentation=cu(inp["segmentation"]),
                               R=cu(inp["R"]), T=cu(inp["T"]), exposure_a=cu(inp["exposure_a"]),
                               exposure_b=cu(inp["exposure_b"]))
    intr = types.SimpleNamespace(fx=inp["fx"] * W / 160, fy=inp["fy"] * H / 120, cx=inp["cx"] * W / 160,
                                 cy=inp["cy"] * H / 120)
    rd, ro = cu(inp["render_depth"]), cu(inp["render_opacity"])
    ref_knn, knn_name = ours_knn, "ours (oracle/_ref not built)"
    try:
        from oracle import ref_cuda
        if ref_cuda.available():
            ref_knn, knn_name = ref_cuda.distCUDA2, "reference CUDA (oracle/_ref)"
    except Exception:  # noqa: BLE001
        pass
    out = {"image": [W, H], "reference_knn": knn_name}
    for init in (True, False):
        factor = a.downsample or (32 if init else 64)
        old = (insert.DOWNSAMPLE_INIT, insert.DOWNSAMPLE)
        insert.DOWNSAMPLE_INIT = insert.DOWNSAMPLE = factor
        ours = lambda: insert.create_viewpoint_pcd(vp, intr, rd, ro, init=init)
        ref = lambda: reference_ops(vp, intr, rd, ro, init, ref_knn, factor)
        k = ours()[0].shape[0]
        ours_dev = lambda: insert.create_viewpoint_pcd(vp, intr, rd, ro, init=init, rng="device")
        out["init" if init else "extend"] = {"points": int(k), "downsample": factor, "ours_ms": timed(ours),
                                             "ours_device_rng_ms": timed(ours_dev), "reference_ops_ms": timed(ref)}
        insert.DOWNSAMPLE_INIT, insert.DOWNSAMPLE = old
    for v in out.values():
        if isinstance(v, dict):
            v["speedup"] = v["reference_ops_ms"] / v["ours_ms"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
