#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` dump: one line per captured kernel with the metrics the roofline needs."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
want = [("Kernel Name", "kernel", 44), ("gpu__time_duration.sum", "us", 9), ("dram__bytes_read.sum", "rdMB", 8),
        ("dram__bytes_write.sum", "wrMB", 8), ("dram__bytes.sum.per_second", "dramTB/s", 8),
        ("sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active", "hmma%", 6),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "hmmaCyc%", 8),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 6),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 6),
        ("l1tex__data_bank_conflicts_pipe_lsu.sum", "bankconf", 9),
        ("launch__grid_size", "grid", 7), ("launch__block_size", "blk", 4), ("launch__registers_per_thread", "regs", 4),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stallLS", 7),
        ("lts__t_sector_hit_rate.pct", "L2hit%", 6)]
idx = [(hdr.index(m), n, w) for m, n, w in want if m in hdr]
print(" ".join(f"{n:>{w}s}" for _, n, w in idx))
for r in data:
    out = []
    for i, n, w in idx:
        v = r[i]
        if n == "kernel":
            v = v.replace("void ", "").replace("cd::", "").replace("<unnamed>::", "")[:w]
        else:
            try:
                v = f"{float(v):.1f}" if n != "dramTB/s" else f"{float(v):.2f}"
            except ValueError:
                pass
        out.append(f"{v:>{w}s}")
    print(" ".join(out))
