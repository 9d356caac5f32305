#!/bin/bash
# GPU pass 5: ncu launch list of the default bench command + full captures of the dominant kernels (after the plain runs).
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/p5_bench_inf.json 2> gpurun_out/p5_bench_inf.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p5_launches_inf.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p5_ncu_inf.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gru_seq|tapgemm_tc" -c 4 -o gpurun_out/p5_fwd -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p5_ncu_full.log 2>&1; echo "ncu full rc=$?"
N=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sim_cc_nss|auc_judd|resize_bicubic" -c 6 -o gpurun_out/p5_metrics -f \
  python tools/bench_metrics.py > gpurun_out/p5_ncu_metrics.log 2>&1; echo "ncu metrics rc=$?"
ls -la gpurun_out/*.ncu-rep
