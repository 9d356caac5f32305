#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_metrics.py -x -q > gpurun_out/metrics_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/metrics_pytest.log
timeout 600 python tools/bench_metrics.py > gpurun_out/bench_metrics.json 2> gpurun_out/bench_metrics.err; echo "rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_metrics.json')); print(d['value'], d['ms_total']); [print(k, {a: round(b,3) for a,b in v.items()}) for k,v in d['kernels'].items()]; print('fused', d['fused']['value'], d['fused']['ms_total']); [print(k, {a: round(b,3) for a,b in v.items()}) for k,v in d['fused']['kernels'].items()]; print(d['available_metrics_set'])"
