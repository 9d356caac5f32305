#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_cascade.py -m gpu -q -x > gpurun_out/r02p13_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02p13_pytest.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/r02p13_bench.json 2> gpurun_out/r02p13_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02p13_bench.err | cut -c1-300
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02p13_bench.json').read().strip().splitlines()[-1])
print("value", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "roof", d["roofline"]["frac"], d["roofline"]["stage_ms"])
print("train", d["train"]["value"], d["train"]["ms_per_step"])
print("refbatch", json.dumps(d.get("reference_batch"))[:900])
print("config5", d["config5"].get("value"), d["config5"].get("stage_ms"), d["config5"].get("error"))
print("config4", {k:v.get("value") for k,v in d["config4"].items() if isinstance(v,dict)})
print("cpu", d["cpu_baseline"])
PY
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02p13_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r02p13_smoke.log
