#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_metrics.py -m gpu -q -x > gpurun_out/r02p12_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02p12_pytest.log | cut -c1-300
timeout 300 python - <<'PY' > gpurun_out/r02p12_cfg5.log 2>&1
import sys, json
sys.argv=["bench.py"]
import bench, torch
args = bench.parse()
H = bench.Harness(torch.device("cuda",0), 0, 0, 1)
print(json.dumps(bench.bench_config5(args, H)))
PY
echo "cfg5 rc=$?"; tail -2 gpurun_out/r02p12_cfg5.log | cut -c1-2500
