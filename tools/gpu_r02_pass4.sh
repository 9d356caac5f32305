#!/bin/bash
# round 2, pass 4: cascade parity, default bench line (with config5), launch list of the default bench
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_cascade.py tests/test_gpu_baseline_configs.py tests/test_gpu_optimizer.py -m gpu -q > gpurun_out/r02p4_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02p4_pytest.log
timeout 600 python bench.py > gpurun_out/r02p4_bench.json 2> gpurun_out/r02p4_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02p4_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02p4_bench.json').read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "train", d["train"]["value"], d["train"]["ms_per_step"])
print("config5", json.dumps(d.get("config5"))[:1800])
PY
