#!/bin/bash
# pass 22: small backward kernels (cb_dw3f, colsum, head_patches) rewritten: training parity + train bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_models.py -x -q -k "train or bptt or lstm or adam or grad or bwd or fc_head" > gpurun_out/p22_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p22_pytest.log
timeout 300 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/p22_bench_train.json 2> gpurun_out/p22_bench_train.err; echo "train rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/p22_bench_train.json')); print('frames/s %.0f' % d['value'], 'ms/step %.3f' % d['ms_per_step'], {k: round(v,3) for k,v in d['roofline']['stage_ms'].items()}, {k: round(v,3) for k,v in d['roofline'].get('stage_ms_bwd').items()})"
