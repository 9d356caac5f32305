#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_baseline_configs.py -m gpu -q -x -k "small_batch" > gpurun_out/r02p16_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02p16_pytest.log | cut -c1-300
timeout 300 python - <<'PY' > gpurun_out/r02p16_refbatch.log 2>&1
import sys, json
sys.argv=["bench.py"]
import bench, torch
args = bench.parse()
H = bench.Harness(torch.device("cuda",0), 0, 0, 1)
print(json.dumps(bench.bench_reference_batch(args, H)))
PY
echo "refbatch rc=$?"; tail -2 gpurun_out/r02p16_refbatch.log | cut -c1-900
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02p16_pytest_all.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/r02p16_pytest_all.log | cut -c1-300
