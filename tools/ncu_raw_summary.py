"""tools/ncu_raw_summary.py -- one line per captured launch from `ncu -i X.ncu-rep --page raw --csv`."""
import csv, subprocess, sys
WANT = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "Minst"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"), ("launch__grid_size", "grid")]
def conv(v, unit, tag):
    v = float(v.replace(",", ""))
    if tag in ("rdMB", "wrMB"):
        v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}[unit]
    if tag == "us":
        v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1)
    if tag == "Minst":
        v *= 1e-6
    return v
def main(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"{'kernel':34s} " + " ".join(f"{t:>8s}" for _, t in WANT))
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("gsr::", "").replace("void ", "")[:34]
        vals = []
        for m, t in WANT:
            vals.append(f"{conv(r[idx[m]], units[idx[m]], t):8.1f}" if m in idx and r[idx[m]] else "     n/a")
        print(f"{name:34s} " + " ".join(vals))
if __name__ == "__main__":
    main(sys.argv[1])
