"""tools/window_bench.py -- one mapping-window iteration (K keyframes, fwd + mapping loss + bwd into window rows)
through monogs_b200.window.MappingWindow on ONE GPU (config C3's shape), with the local keyframes on 1 vs 2 streams.
Median of 10 windows after 3 warm-ups, CUDA events."""
import json, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth, window

def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "replica"
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 500_000
    K = int(sys.argv[3]) if len(sys.argv) > 3 else 8
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cams = synth.arc_cameras(kind, K)
    sc = {k: v.to(dev) for k, v in synth.make_scene(P, synth.make_camera(kind), seed=0).items()}
    targets = [tuple(t.to(dev) for t in synth.make_targets(c, seed=i)) for i, c in enumerate(cams)]
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    out = {"kind": kind, "gaussians": P, "keyframes": K, "ms_per_window": {}}
    ref = None
    for S in (1, 2, 3):
        w = window.MappingWindow(window.make_cuda_render_fn(cams, targets, bg), n_streams=S)
        for _ in range(3):
            res = w.step(sc, K)
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); res = w.step(sc, K); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        out["ms_per_window"][str(S)] = sorted(ts)[len(ts) // 2]
        g = torch.cat([v.reshape(P, -1) for v in res.grads.values()], 1).clone()
        if ref is None:
            ref = g
        else:
            out[f"rel_err_streams{S}_vs_1"] = float((g - ref).abs().max() / ref.abs().max())
    print(json.dumps(out))

if __name__ == "__main__":
    main()
