#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_cascade.py -m gpu -q -x > gpurun_out/r02p9_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02p9_pytest.log | cut -c1-300
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r02p9_bench.json 2> gpurun_out/r02p9_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02p9_bench.err | cut -c1-300; python - <<PY
import json
d=json.loads(open('gpurun_out/r02p9_bench.json').read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["stage_ms"], d["roofline"]["frac"])
print("train", d["train"]["value"], d["train"]["ms_per_step"], d["train"]["roofline"]["stage_ms"], d["train"]["roofline"].get("stage_ms_bwd"))
print("config5", d["config5"].get("value"), d["config5"].get("stage_ms"), d["config5"].get("error"))
print("config4", {k:(v.get("value"), v.get("stage_ms")) for k,v in d["config4"].items() if isinstance(v,dict)})
PY
