#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "persistent_recurrence" > gpurun_out/r02p11_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02p11_pytest.log | cut -c1-300
for nt in 2 1; do RGP_SEQ_NTILE=$nt timeout 300 python bench.py --no-sub --no-cpu-baseline --no-e2e > gpurun_out/r02p11_bench_nt$nt.json 2> gpurun_out/r02p11_bench$nt.err; echo "bench ntile=$nt rc=$?"; tail -2 gpurun_out/r02p11_bench$nt.err | cut -c1-300; python - <<PY
import json
d=json.loads(open('gpurun_out/r02p11_bench_nt$nt.json').read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["stage_ms"])
PY
done
timeout 600 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_cascade.py -m gpu -q > gpurun_out/r02p11_pytest2.log 2>&1; echo "pytest2 rc=$?"; tail -4 gpurun_out/r02p11_pytest2.log | cut -c1-300
