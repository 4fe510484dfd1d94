"""Window-row exchange check on >= 2 GPUs (run under torchrun, NCCL):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/xchg_check.py [--gaussians 200000] [--keyframes 5] [--out gpurun_out/xchg_check.json]

Runs the sharded mapping window (monogs_b200.window.MappingWindow + the sm_100a rasterizer) once per exchange
mode -- our sparse in-switch kernel (gsr_rows_exchange), torch's dense multimem allreduce, ncclAllReduce -- on the
same scene and compares gradients, densification statistics, visibility counts and radii of the sparse mode with
the NCCL result; then times the three exchanges on the rows of one keyframe of the headline-sized map.
Rank 0 prints one JSON line (and writes --out)."""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def run_mode(mode, cams, targets, bg, sc, n_kf, n_streams=2):
    from monogs_b200 import window
    os.environ["GSR_XCHG"] = mode
    w = window.MappingWindow(window.make_cuda_render_fn(cams, targets, bg), n_streams=n_streams)
    res = None
    for _ in range(3):  # the first step decides / warms up; later steps must give the same sums
        res = w.step(sc, n_kf)
    torch.cuda.synchronize()
    out = {k: v.clone() for k, v in res.grads.items()}
    out["grad2d_norm"] = res.grad2d_norm.clone()
    out["vis_count"] = res.vis_count.clone()
    out["max_radii"] = res.max_radii.clone().float()
    return out, w.exchange, w


def snapshot(res):
    out = {k: v.clone() for k, v in res.grads.items()}
    out["grad2d_norm"] = res.grad2d_norm.clone()
    out["vis_count"] = res.vis_count.clone()
    out["max_radii"] = res.max_radii.clone().float()
    return out


def exchange_trace_us():
    """Phase durations (us) of the last gsr_rows_exchange on this device: marks multicast, barrier 1, row sums, barrier 2."""
    import ctypes as C
    from monogs_b200 import _abi
    t = (C.c_uint64 * 5)()
    torch.cuda.synchronize()
    if _abi.load().gsr_debug_exchange_trace(t) != 0 or t[0] == 0:
        return None
    names = ("marks_multicast", "barrier1", "row_sums", "barrier2")
    return {n: round((t[i + 1] - t[i]) / 1e3, 2) for i, n in enumerate(names)} | {"total": round((t[4] - t[0]) / 1e3, 2)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=200000)
    ap.add_argument("--keyframes", type=int, default=5)
    ap.add_argument("--kind", default="tum")
    ap.add_argument("--out", default="")
    ap.add_argument("--opacity-scale", type=float, default=1.0,
                    help="multiply the opacities (a translucent map: most Gaussians receive a gradient = low occlusion)")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    from monogs_b200 import synth

    cams = synth.arc_cameras(a.kind, a.keyframes)
    sc = {k: v.to(dev) for k, v in synth.make_scene(a.gaussians, synth.make_camera(a.kind), seed=1).items()}
    if a.opacity_scale != 1.0:
        sc["opacities"] = (sc["opacities"] * a.opacity_scale).contiguous()
    targets = [tuple(t.to(dev) for t in synth.make_targets(c, seed=i)) for i, c in enumerate(cams)]
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)

    report = {"world": world, "gaussians": a.gaussians, "keyframes": a.keyframes, "kind": a.kind,
              "opacity_scale": a.opacity_scale, "modes": {}}
    # reference result: ncclAllReduce, every local keyframe on ONE stream; the modes below also run the local
    # keyframes round-robin on two streams (MappingWindow n_streams=2), so that path is checked as well
    ref, _, w_ref = run_mode("all_reduce", cams, targets, bg, sc, a.keyframes, n_streams=1)
    worst = 0.0
    w_sparse = None
    for mode in ("sparse", "multimem"):
        try:
            got, ex, w_mode = run_mode(mode, cams, targets, bg, sc, a.keyframes)
            if mode == "sparse":
                w_sparse = w_mode
                report["sparse_phases_us"] = exchange_trace_us()
                report["sparse_report"] = {k: v for k, v in ex.report.items() if k != "symmetric_memory_error"}
        except Exception as e:  # noqa: BLE001
            report["modes"][mode] = {"error": repr(e)[:300]}
            continue
        errs = {}
        for k, r in ref.items():
            scale = float(r.abs().max().clamp_min(1e-30))
            errs[k] = float((got[k].reshape(r.shape) - r).abs().max()) / scale
        st = int(ex._status.item()) if getattr(ex, "_status", None) is not None else None
        report["modes"][mode] = {"used_mode": ex.mode, "rel_err": errs, "status": st}
        if ex.mode == mode:
            worst = max(worst, max(errs.values()))
    # rows that carry a gradient on SOME rank (what the sparse kernel moves) as a fraction of the map
    nz = (ref["means3D"].abs().sum(1) + ref["colors"].abs().sum(1)) > 0
    report["touched_fraction_union"] = float(nz.float().mean())
    # the map grows (keyframe insertion / densification): first past the buffer's capacity (one re-allocation and
    # rendezvous, mode kept), then inside the new capacity (same block) -- both must still equal the NCCL sums
    if w_sparse is not None:
        g = torch.Generator().manual_seed(99)
        grow = {}
        for extra in (max(4096, a.gaussians // 20), max(8192, a.gaussians // 16)):
            more = synth.make_scene(extra, synth.make_camera(a.kind), seed=7)
            sc2 = {k: torch.cat([sc[k], more[k].to(dev)]).contiguous() for k in sc}
            cap_before = w_sparse.exchange._cap
            os.environ["GSR_XCHG"] = "sparse"
            got2 = snapshot(w_sparse.step(sc2, a.keyframes))
            os.environ["GSR_XCHG"] = "all_reduce"
            ref2 = snapshot(w_ref.step(sc2, a.keyframes))
            errs = {k: float((got2[k].reshape(r.shape) - r).abs().max()) / float(r.abs().max().clamp_min(1e-30))
                    for k, r in ref2.items()}
            grow[str(sc2["means3D"].shape[0])] = {"mode": w_sparse.exchange.mode, "capacity": w_sparse.exchange._cap,
                                                  "reallocated": w_sparse.exchange._cap != cap_before,
                                                  "rel_err": max(errs.values())}
            worst = max(worst, max(errs.values()))
        report["grow"] = grow
    t = torch.tensor([worst], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    report["worst_rel_err_all_ranks"] = float(t)
    report["pass"] = bool(float(t) <= 1e-5)
    if rank == 0:
        line = json.dumps(report)
        print(line, flush=True)
        if a.out:
            Path(a.out).parent.mkdir(parents=True, exist_ok=True)
            Path(a.out).write_text(line + "\n")
    dist.barrier()
    dist.destroy_process_group()
    if not report["pass"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
