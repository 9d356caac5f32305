#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_metrics.py -m gpu -q -x > gpurun_out/r02p25_metrics_tests.log 2>&1; echo "metrics tests rc=$?"; tail -5 gpurun_out/r02p25_metrics_tests.log | cut -c1-400
for mode in smooth noise; do
  RGP_SIM_FUSED_KERNEL=old timeout 300 python tools/ab_sim_fused.py gpurun_out/r02p25_old_$mode.npy $mode 2>&1 | tail -1
  timeout 300 python tools/ab_sim_fused.py gpurun_out/r02p25_lean_$mode.npy $mode 2>&1 | tail -1
  python - <<P
import numpy as np
a=np.load("gpurun_out/r02p25_old_$mode.npy"); b=np.load("gpurun_out/r02p25_lean_$mode.npy")
print("$mode", "max|d| sim cc nss:", np.nanmax(np.abs(a[:,:3]-b[:,:3]),0), "min/max/nfix equal:", np.array_equal(a[:,3:6],b[:,3:6]), "sums rel:", np.max(np.abs(a[:,6:]-b[:,6:])/np.abs(a[:,6:])))
P
done
timeout 600 python tools/bench_metrics.py > gpurun_out/r02p25_bench_metrics.json 2> gpurun_out/r02p25_bench_metrics.err; echo "bench_metrics rc=$?"; python -c "
import json;d=json.load(open('gpurun_out/r02p25_bench_metrics.json'));print({k:round(v['ms'],3) for k,v in d['kernels'].items()});print({k:round(v['ms'],3) for k,v in d['fused']['kernels'].items()});print(d['available_metrics_set'])"
