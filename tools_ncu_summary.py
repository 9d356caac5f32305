#!/usr/bin/env python
"""Summarise an ncu --set full report (raw page CSV) into the handful of metrics DESIGN.md / bench quote."""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__cycles_active",
        "gpu__dram_throughput.avg.pct", "sm__pipe_tensor_cycles_active", "sm__inst_executed_pipe_tensor",
        "sm__warps_active.avg.pct", "launch__registers_per_thread", "lts__t_bytes.sum", "lts__throughput.avg.pct",
        "l1tex__throughput.avg.pct", "sm__throughput.avg.pct", "launch__shared_mem_per_block_dynamic",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "lts__t_sectors_srcunit_tex_op_read.sum",
        "sm__pipe_tensor_subpipe", "tensor"]
cols = [i for i, h in enumerate(hdr) if any(k in h for k in keys)]
for d in data:
    print("== %s grid %s block %s" % (d[hdr.index("Kernel Name")][:70], d[hdr.index("Grid Size")], d[hdr.index("Block Size")]))
    for i in cols:
        print("   %-90s %s %s" % (hdr[i], d[i], units[i]))
