#!/bin/bash
# round 2, pass 5: persistent ConvLSTM kernel (X1): parity, then config-4 timing at 8 clips
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 420 python -m pytest tests/test_gpu_baseline_configs.py -m gpu -q -x -k "persistent_convlstm or config4" > gpurun_out/r02p5_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r02p5_pytest.log | cut -c1-300
timeout 300 python - <<'PY' > gpurun_out/r02p5_cfg4.log 2>&1
import sys, json, argparse
sys.argv=["bench.py"]
import bench, torch
args = bench.parse()
H = bench.Harness(torch.device("cuda",0), 0, 0, 1)
print(json.dumps(bench.bench_config4(args, H)))
PY
echo "cfg4 rc=$?"; tail -3 gpurun_out/r02p5_cfg4.log | cut -c1-1500
