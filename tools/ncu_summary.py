#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): per launch - kernel, grid, duration, tensor-pipe %, DRAM bytes read / written,
DRAM throughput %, L2 throughput %, registers.  Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.txt"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {n: i for i, n in enumerate(hdr)}
want = [("gpu__time_duration.sum", "dur"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"), ("launch__registers_per_thread", "regs"),
        ("launch__grid_size", "grid"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%")]
print("# %s" % rep)
print("%-4s %-46s %8s %s" % ("id", "kernel", "grid", " ".join("%12s" % w[1] for w in want if w[1] != "grid")))
for r in data:
    name = r[col["Kernel Name"]][:46]
    vals = []
    for m, short in want:
        if short == "grid":
            continue
        v = r[col[m]] if m in col else ""
        u = units[col[m]] if m in col else ""
        vals.append("%12s" % ((v + u[:3]) if v else "-"))
    print("%-4s %-46s %8s %s" % (r[col["ID"]], name, r[col["launch__grid_size"]] if "launch__grid_size" in col else "", " ".join(vals)))
