"""tools/window_bench.py -- one mapping-window iteration (K keyframes, fwd + mapping loss + bwd into window rows)
through monogs_b200.window.MappingWindow on ONE GPU (config C3's shape), with the local keyframes on 1 vs 2 streams and the mapping loss by framework ops vs fused (gsr_l1_loss_forward +
gsr_backward_render_l1); gradients / loss of every variant are compared with the first.
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
    for fused in (False, True):
        for S in (1, 2):
            w = window.MappingWindow(window.make_cuda_render_fn(cams, targets, bg, fused_loss=fused), n_streams=S)
            for _ in range(3):
                res = w.step(sc, K)
            torch.cuda.synchronize()
            ts = []
            for _ in range(10):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); res = w.step(sc, K); b.record(); torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            tag = ("fused_loss" if fused else "torch_loss") + f"_streams{S}"
            out["ms_per_window"][tag] = sorted(ts)[len(ts) // 2]
            g = torch.cat([v.reshape(P, -1) for v in res.grads.values()], 1).clone()
            loss = float(sum(r.loss for r in res.local))
            if ref is None:
                ref, ref_loss = g, loss
            else:
                out[f"rel_err_{tag}"] = float((g - ref).abs().max() / ref.abs().max())
                out[f"loss_rel_err_{tag}"] = abs(loss - ref_loss) / abs(ref_loss)
    print(json.dumps(out))

if __name__ == "__main__":
    main()
