"""Summarise `ncu -i X.ncu-rep --page source --csv --print-source sass`: per kernel, the instructions with the
most stall samples and the most executions, plus totals per stall reason.  usage: ncu_source_hot.py file.csv [topN]"""
import csv
import sys
from collections import defaultdict


def main(path, top=25):
    rows = list(csv.reader(open(path)))
    kernels, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1][:90], "hdr": None, "rows": []}
            kernels.append(cur)
        elif cur is not None and cur["hdr"] is None and r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and cur["hdr"] is not None and r:
            cur["rows"].append(r)
    for k in kernels:
        h = k["hdr"]
        ix = {n: i for i, n in enumerate(h)}
        samp, ex = ix["# Samples"], ix["Instructions Executed"]
        stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
        tot = defaultdict(int)
        total_s = total_e = 0
        for r in k["rows"]:
            total_s += int(r[samp] or 0)
            total_e += int(r[ex] or 0)
            for i in stall_cols:
                tot[h[i]] += int(r[i] or 0)
        print("=" * 100)
        print(k["name"], "| samples", total_s, "| warp-instr executed", total_e)
        print("stall totals:", ", ".join(f"{n[6:]}={v}" for n, v in sorted(tot.items(), key=lambda x: -x[1]) if v))
        print("-- top by samples (idx, samples, executed, sass, main stalls)")
        order = sorted(range(len(k["rows"])), key=lambda j: -int(k["rows"][j][samp] or 0))[:top]
        for j in sorted(order):
            r = k["rows"][j]
            st = sorted(((int(r[i] or 0), h[i][6:]) for i in stall_cols), reverse=True)[:3]
            print(f"{j:5d} {int(r[samp] or 0):6d} {int(r[ex] or 0):9d}  {r[1].strip()[:60]:60s} "
                  + " ".join(f"{n}={v}" for v, n in st if v))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
