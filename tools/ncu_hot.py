#!/usr/bin/env python
"""Hot instructions of one kernel from `ncu --page source --csv` (SASS view): samples, executions, top stall reasons.
    ncu -i X.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv ; python tools/ncu_hot.py src.csv [N]"""
import csv
import sys


def main(path, topn=40):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ix = {c: i for i, c in enumerate(hdr)}
    data = [r for r in rows[h + 1:] if len(r) >= len(hdr) and r[0].startswith("0x")]
    tot = sum(int(r[ix["# Samples"]]) for r in data)
    ninst = sum(int(r[ix["Instructions Executed"]]) for r in data)
    print("samples", tot, "sass lines", len(data), "warp instructions", ninst)
    stalls = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
    agg = {c: sum(int(r[ix[c]]) for r in data) for c in stalls}
    print("stall totals:", sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
    order = sorted(range(len(data)), key=lambda i: -int(data[i][ix["# Samples"]]))[:topn]
    for i in order:
        r = data[i]
        s = sorted(((int(r[ix[c]]), c[6:]) for c in stalls if int(r[ix[c]]) > 0), reverse=True)[:3]
        print(f"{i:5d} {r[ix['# Samples']]:>7s} {r[ix['Instructions Executed']]:>10s}  {r[ix['Source']].strip()[:64]:64s} {s}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
