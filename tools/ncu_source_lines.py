"""tools/ncu_source_lines.py REP KERNEL_SUBSTR [N] -- per CUDA source line (summed over the captured launches): stall samples and executed warp
instructions, from `ncu -i REP --page source --csv --print-source cuda,sass` (needs -lineinfo, --import-source on)."""
import csv, subprocess, sys

def main(rep, kern, top=25):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True, errors="replace").stdout
    rows = list(csv.reader(out.splitlines()))
    fn, hdr, acc, seen = None, None, {}, False
    for r in rows:
        if r and r[0] == "Function Name":
            fn = r[1]; hdr = None; continue
        if r and r[0] == "Line No":
            hdr = r; continue
        if not hdr or fn is None or kern not in fn or len(r) != len(hdr): continue
        if r[0] == "": continue     # SASS row
        seen = True
        i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
        key = (int(r[0]), r[1].strip()[:90])
        a = acc.setdefault(key, [0, 0]); a[0] += int(r[i_s] or 0); a[1] += int(r[i_i] or 0)
    ts, ti = sum(a[0] for a in acc.values()) or 1, sum(a[1] for a in acc.values()) or 1
    print(f"{kern}: samples {ts} warp-inst {ti}")
    for (ln, src), (s, i) in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{ln:5d} samp {100*s/ts:5.1f}% inst {100*i/ti:5.1f}% | {src}")

if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 25)
