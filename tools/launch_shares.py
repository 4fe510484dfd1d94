"""tools/launch_shares.py launches.csv -- per-kernel count, mean duration and share of the summed kernel time
from `ncu --metrics gpu__time_duration.sum --csv` (the launch list of tools/gpu_round.sh launches)."""
import collections, csv, sys

def main(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 5]
    h = next(i for i, r in enumerate(rows) if r[0] == "ID")
    hdr = rows[h]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[h + 1:]:
        v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1e-3)
        agg.setdefault(r[ki], []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none, command: python bench.py --steps 2 --warmup 1 "
          "--no-e2e --no-cpu-baseline --no-window-c4")
    print("# (cold-cache, serialised launches: compare SHARES with bench.py's live CUDA-event stage times, not absolutes)")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k[:70]:70s} n={len(v):3d} mean_us={sum(v)/len(v):8.1f} share={100*sum(v)/tot:5.1f}%")

if __name__ == "__main__":
    main(sys.argv[1])
