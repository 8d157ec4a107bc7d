#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` export: one block per kernel launch with the metrics the roofline uses.
    ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; python tools/ncu_summary.py X_raw.csv"""
import csv
import re
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp16.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "smsp__cycles_active.avg", "sm__inst_executed_pipe_tc.sum", "sm__inst_executed_pipe_uniform.sum",
]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = rows[0]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        print("==", r[idx["Kernel Name"]][:90])
        for k in KEYS:
            if k in idx:
                print(f"   {k:75s} {r[idx[k]]}")
        st = [(float(r[i].replace(',', '')), h) for h, i in idx.items()
              if re.match(r"smsp__average_warps?_issue_stalled_.*_per_issue_active\.ratio|smsp__average_warp_latency_issue_stalled_.*\.ratio", h) and r[i] not in ("", "n/a")]
        st.sort(reverse=True)
        for v, h in st[:6]:
            print(f"   stall {h.split('issue_stalled_')[1].split('.')[0]:40s} {v:.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
