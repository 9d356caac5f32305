#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_baseline_configs.py -m gpu -q -x -k "persistent_convlstm or config4" > gpurun_out/r02p10_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02p10_pytest.log | cut -c1-300
timeout 300 python - <<'PY' > gpurun_out/r02p10_cfg4.log 2>&1
import sys, json
sys.argv=["bench.py"]
import bench, torch
args = bench.parse()
H = bench.Harness(torch.device("cuda",0), 0, 0, 1)
print(json.dumps(bench.bench_config4(args, H)))
PY
echo "cfg4 rc=$?"; tail -2 gpurun_out/r02p10_cfg4.log | cut -c1-1800
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_models.py -m gpu -q > gpurun_out/r02p10_pytest2.log 2>&1; echo "pytest2 rc=$?"; tail -4 gpurun_out/r02p10_pytest2.log | cut -c1-300
