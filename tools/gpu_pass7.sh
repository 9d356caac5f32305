#!/bin/bash
# GPU pass 7: resident-A-slab tap-GEMM — parity, then A/B bench lines (RGP_SLAB=0 = previous per-tap A boxes).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_models.py -q -x > gpurun_out/p7_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/p7_pytest.log
for v in 0 1; do
  RGP_SLAB=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p7_bench_inf_slab$v.json 2> gpurun_out/p7_bench_inf_slab$v.err; echo "bench slab=$v rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/p7_bench_inf_slab$v.json').read().strip().splitlines()[-1])
print('slab=$v', round(d['value']), d['ms_per_step'], {k: round(x,3) for k,x in d['roofline']['stage_ms'].items()}, round(d['roofline']['frac'],3))
PY
done
for v in 0 1; do
  RGP_SLAB=$v timeout 600 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p7_bench_train_slab$v.json 2> gpurun_out/p7_bench_train_slab$v.err; echo "train slab=$v rc=$?"
  python -c "
import json
d=json.loads(open('gpurun_out/p7_bench_train_slab$v.json').read().strip().splitlines()[-1])
print('train slab=$v', round(d['value']), d['ms_per_step'])"
done
