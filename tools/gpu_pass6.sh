#!/bin/bash
# GPU pass 6: full ncu capture of the fused resize+metric kernels (after the plain run of the same command).
mkdir -p gpurun_out
N=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fused_kernel" -c 4 -o gpurun_out/p6_metrics -f \
  python tools/bench_metrics.py > gpurun_out/p6_ncu_metrics.log 2>&1; echo "ncu metrics rc=$?"
ls -la gpurun_out/p6*.ncu-rep
