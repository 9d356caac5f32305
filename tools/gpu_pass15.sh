#!/bin/bash
# GPU pass 15: 256-bit epilogue accesses in the tap-GEMM engine — full GPU suite, inference + train bench lines.
mkdir -p gpurun_out
timeout 1400 python -m pytest tests -m gpu -q > gpurun_out/p15_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p15_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p15_bench_inf.json 2> gpurun_out/p15_bench_inf.err; echo "bench rc=$?"
timeout 600 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p15_bench_train.json 2> gpurun_out/p15_bench_train.err; echo "train rc=$?"
python - <<PY
import json
for f in ("p15_bench_inf","p15_bench_train"):
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
    print(f, round(d['value']), round(d['ms_per_step'],3), {k: round(x,3) for k,x in d['roofline']['stage_ms'].items()}, {k: round(x,3) for k,x in d['roofline'].get('stage_ms_bwd',{}).items()})
PY
