#!/bin/bash
# GPU pass 2: training bench with backward stage timing, metric-kernel bench, ncu launch list of a training step,
# full ncu capture of the wgrad kernel and the metric kernels.
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_train3.json 2> gpurun_out/bench_train3.err; echo "bench train rc=$?"
timeout 600 python tools/bench_metrics.py > gpurun_out/bench_metrics.json 2> gpurun_out/bench_metrics.err; echo "metrics rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_train.csv \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/ncu_train.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wgrad_tc -c 3 -o gpurun_out/wgrad_tc -f \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/ncu_full_wgrad.log 2>&1; echo "ncu wgrad rc=$?"
N=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sim_cc_nss|auc_judd|resize_bicubic|auc_sampled" -c 8 -o gpurun_out/metrics_k -f \
  python tools/bench_metrics.py > gpurun_out/ncu_full_metrics.log 2>&1; echo "ncu metrics rc=$?"
cat gpurun_out/bench_train3.json gpurun_out/bench_metrics.json
