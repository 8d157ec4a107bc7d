#!/usr/bin/env python
"""Per-launch DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) of the kernels in an `ncu --page raw --csv`
export, as the JSON table bench.py reads for `roofline.traffic`:

    ncu -i X.ncu-rep --page raw --csv > X_raw.csv
    python tools/ncu_traffic.py X_raw.csv C4:auto [C4:nodes ...] > profiles/r2_ncu_traffic.json

Keys are "<workload>:<algorithm>:<kernel name>"; values are bytes per launch (mean over the captured launches)."""
import csv
import json
import re
import sys


def main(path, prefixes):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        name = re.sub(r"<.*", "", re.sub(r"^(void )?(<unnamed>::)?", "", r[idx["Kernel Name"]])).split("(")[0]
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[idx[k]].replace(",", "")) * scale[units[idx[k]]]
        acc.setdefault(name, []).append(tot)
    out = {"_source": path, "_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full)"}
    for name, v in sorted(acc.items()):
        for p in prefixes:
            out[f"{p}:{name}"] = sum(v) / len(v)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:] or ["C4:auto"])
