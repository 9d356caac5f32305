#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests/test_gpu_edge_cases.py tests/test_gpu_parity.py tests/test_gpu_baseline_configs.py -m gpu -q > gpurun_out/r02p14_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02p14_pytest.log | cut -c1-300
for f in 1 0; do RGP_IM2COL=$f timeout 300 python bench.py --mode train --batch 128 --no-sub --no-cpu-baseline --no-e2e > gpurun_out/r02p14_train_im2col$f.json 2> gpurun_out/r02p14_train$f.err; echo "train im2col=$f rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r02p14_train_im2col$f.json').read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["stage_ms"], d["roofline"].get("stage_ms_bwd"))
PY
done
