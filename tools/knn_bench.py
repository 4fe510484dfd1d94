"""tools/knn_bench.py -- distCUDA2 on 1 M back-projected points (config C5's kNN leg): ours vs the reference CUDA
(oracle/_ref), CUDA events, median of 20 after 5 warm-ups; checks bit-equality of the two results."""
import json, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from monogs_b200 import synth
from monogs_b200.simple_knn._C import distCUDA2

def timed(fn, n=20, warm=5):
    for _ in range(warm): fn()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]

def main():
    P = 1_000_000
    cam = synth.make_camera("tum")
    pts = synth.make_scene(P, cam, seed=0, device="cuda")["means3D"].contiguous()
    out = {"points": P, "ours_ms": timed(lambda: distCUDA2(pts))}
    ours = distCUDA2(pts)
    try:
        from oracle import ref_cuda
        if ref_cuda.available():
            out["reference_ms"] = timed(lambda: ref_cuda.distCUDA2(pts))
            ref = ref_cuda.distCUDA2(pts)
            out["bit_equal"] = bool(torch.equal(ours, ref))
            out["speedup"] = out["reference_ms"] / out["ours_ms"]
    except Exception as e:  # reference build absent
        out["reference"] = f"unavailable: {e}"
    print(json.dumps(out))

if __name__ == "__main__":
    main()
