#!/usr/bin/env python
"""Summarise an ncu --set full report (raw page CSV) into the handful of metrics DESIGN.md / bench.py quote.
usage: python tools/ncu_summary.py report.ncu-rep > profiles/rNN_<what>.txt"""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__cycles_elapsed.max",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed.sum", "smsp__cycles_active.avg"]
def col(h):
    h = h.split(".TriageCompute.")[-1]
    return h in KEYS
cols = [i for i, h in enumerate(hdr) if col(h)]
for d in data:
    print("== %s  grid %s block %s" % (d[hdr.index("Kernel Name")][:90], d[hdr.index("Grid Size")], d[hdr.index("Block Size")]))
    for i in cols:
        print("   %-86s %s %s" % (hdr[i].split(".TriageCompute.")[-1], d[i], units[i]))
