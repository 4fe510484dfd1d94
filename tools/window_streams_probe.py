"""tools/window_streams_probe.py -- how much of a GPU the one-keyframe pipeline leaves idle: K keyframes of a
mapping window (fwd + bwd each, own workspaces, through the C ABI like bench.py's device leg) issued round-robin on
S CUDA streams of ONE GPU.  Prints ms per window for S = 1, 2, 3, 4, 8."""
import json, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    K = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    steppers = []
    for k in range(K):
        wl = bench.Workload(name, k, K, dev)          # camera k of the K-view arc, same Gaussians
        if steppers:                                   # share the scene tensors (one map)
            wl.scene = steppers[0].wl.scene
        steppers.append(bench.OursDeviceStep(wl, 1))
    for s in steppers:
        s.step(); s.step()
    torch.cuda.synchronize()
    out = {"workload": name, "keyframes": K, "ms_per_window": {}}
    for S in (1, 2, 3, 4, 8):
        streams = [torch.cuda.Stream() for _ in range(S)]
        def window():
            cur = torch.cuda.current_stream()
            for st in streams:
                st.wait_stream(cur)
            for k, s in enumerate(steppers):
                with torch.cuda.stream(streams[k % S]):
                    s.step()
            for st in streams:
                cur.wait_stream(st)
        for _ in range(3):
            window()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        a.record()
        for _ in range(n):
            window()
        b.record()
        torch.cuda.synchronize()
        out["ms_per_window"][str(S)] = a.elapsed_time(b) / n
    print(json.dumps(out))

if __name__ == "__main__":
    main()
