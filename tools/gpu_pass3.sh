#!/bin/bash
# GPU pass 3: validate HEAD — GPU tests, smoke, inference + train bench lines, reference arm.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p3_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p3_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/p3_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/p3_smoke.log
timeout 600 python bench.py > gpurun_out/p3_bench_inf.json 2> gpurun_out/p3_bench_inf.err; echo "bench inf rc=$?"
timeout 600 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/p3_bench_train.json 2> gpurun_out/p3_bench_train.err; echo "bench train rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/p3_bench_ref.json 2> gpurun_out/p3_bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/p3_bench_inf.json gpurun_out/p3_bench_train.json gpurun_out/p3_bench_ref.json
