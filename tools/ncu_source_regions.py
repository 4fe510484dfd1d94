"""tools/ncu_source_regions.py -- summarise `ncu --page source --csv` output: contiguous SASS regions with similar
execution counts, their share of executed warp-instructions and of stall samples, and the top stall reasons."""
import csv, sys

def blocks(path):
    rows = list(csv.reader(open(path)))
    out, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "data": []}
            out.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None and len(r) == len(cur["hdr"]):
            cur["data"].append(r)
    return out

def main(path, which=0, thresh=0.01):
    b = blocks(path)[which]
    hdr, data = b["hdr"], b["data"]
    ia, isamp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[ia]) for r in data); tots = sum(int(r[isamp]) for r in data)
    print(b["name"][:80]); print("total warp-inst", tot, "samples", tots, "sass lines", len(data))
    regions, cur = [], None
    for i, r in enumerate(data):
        n = int(r[ia])
        if cur and abs(n - cur["n"]) <= 0.15 * max(n, cur["n"], 1):
            cur["b"] = i; cur["inst"] += n; cur["samp"] += int(r[isamp])
            for c in stall_cols: cur["st"][c] = cur["st"].get(c, 0) + int(r[c] or 0)
        else:
            if cur: regions.append(cur)
            cur = {"a": i, "b": i, "n": n, "inst": n, "samp": int(r[isamp]), "st": {c: int(r[c] or 0) for c in stall_cols}}
    regions.append(cur)
    for g in regions:
        if g["inst"] > tot * thresh or g["samp"] > tots * thresh:
            top = sorted(g["st"].items(), key=lambda kv: -kv[1])[:3]
            tops = " ".join(f"{hdr[c][6:]}={v}" for c, v in top if v)
            print(f"sass[{g['a']:4d}:{g['b']:4d}] exec/inst~{g['n']:9d} inst {g['inst']/tot*100:5.1f}% samples {g['samp']/max(tots,1)*100:5.1f}% | {tops} | {data[g['a']][isrc].strip()[:44]}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, float(sys.argv[3]) if len(sys.argv) > 3 else 0.01)
