#!/bin/bash
# GPU pass 10: A/B — projection with the L2 row prefetch (this build); training at 128 clips with the persistent recurrence forced.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "forward_parity or bptt" > gpurun_out/p10_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/p10_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p10_bench_inf.json 2> gpurun_out/p10_bench_inf.err; echo "bench rc=$?"
for v in 0 1; do
  RGP_SEQ=$v timeout 600 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p10_bench_train_seq$v.json 2> gpurun_out/p10_bench_train_seq$v.err; echo "train seq=$v rc=$?"
done
python - <<PY
import json
for f in ("p10_bench_inf","p10_bench_train_seq0","p10_bench_train_seq1"):
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
    print(f, round(d['value']), round(d['ms_per_step'],3), {k: round(x,3) for k,x in d['roofline']['stage_ms'].items()})
PY
