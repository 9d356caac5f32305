#!/bin/bash
# GPU pass 8: persistent recurrence with the resident state slab — parity (under a timeout: a pipeline bug would hang), then bench.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "persistent" > gpurun_out/p8_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -4 gpurun_out/p8_pytest_seq.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p8_bench_inf.json 2> gpurun_out/p8_bench_inf.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/p8_bench_inf.json').read().strip().splitlines()[-1])
print(round(d['value']), d['ms_per_step'], {k: round(x,3) for k,x in d['roofline']['stage_ms'].items()}, round(d['roofline']['frac'],3))
PY
timeout 300 python tools/seq_profile.py > gpurun_out/p8_seq_profile.txt 2>&1; tail -12 gpurun_out/p8_seq_profile.txt
