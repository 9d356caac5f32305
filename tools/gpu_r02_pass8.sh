#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k im2col > gpurun_out/r02p8_im2col.log 2>&1; echo "im2col rc=$?"; tail -15 gpurun_out/r02p8_im2col.log | cut -c1-300
for f in 1 0; do RGP_IM2COL=$f timeout 300 python bench.py --no-sub --no-cpu-baseline --no-e2e > gpurun_out/r02p8_bench_im2col$f.json 2> gpurun_out/r02p8_bench$f.err; echo "bench im2col=$f rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/r02p8_bench_im2col$f.json').read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["stage_ms"], d["roofline"]["frac"])
PY
done
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02p8_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02p8_pytest.log | cut -c1-300
