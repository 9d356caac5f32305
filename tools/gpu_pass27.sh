#!/bin/bash
# pass 27: pack_c3d grid order (channel group fastest) - forward parity tests + inference bench (stage_ms.embed)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "forward or pack or embed" > gpurun_out/p27_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p27_pytest.log
for i in 1 2; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p27_bench_inf$i.json 2> gpurun_out/p27_bench_inf$i.err; echo "inf rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/p27_bench_inf$i.json')); print('frames/s %.0f' % d['value'], 'ms/step %.3f' % d['ms_per_step'], d['clocks'], {k: round(v,3) for k,v in d['roofline']['stage_ms'].items()})"
done
