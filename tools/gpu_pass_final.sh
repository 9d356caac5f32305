#!/bin/bash
# Final GPU pass of a round (was pass 9; artifacts ${PFX:-p16}_*): validate HEAD (tests, smoke, bench lines), then ncu evidence for the slab kernels and the fused metric kernels.
mkdir -p gpurun_out
timeout 1400 python -m pytest tests -m gpu -q > gpurun_out/${PFX:-p16}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${PFX:-p16}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${PFX:-p16}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${PFX:-p16}_smoke.log
timeout 600 python bench.py > gpurun_out/${PFX:-p16}_bench_inf.json 2> gpurun_out/${PFX:-p16}_bench_inf.err; echo "bench inf rc=$?"
timeout 600 python bench.py --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${PFX:-p16}_bench_train.json 2> gpurun_out/${PFX:-p16}_bench_train.err; echo "bench train rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${PFX:-p16}_bench_ref.json 2> gpurun_out/${PFX:-p16}_bench_ref.err; echo "bench ref rc=$?"
timeout 600 python tools/bench_metrics.py > gpurun_out/${PFX:-p16}_bench_metrics.json 2> gpurun_out/${PFX:-p16}_bench_metrics.err; echo "bench metrics rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${PFX:-p16}_launches_inf.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/${PFX:-p16}_ncu_inf.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gru_seq|tapgemm_tc" -c 4 -o gpurun_out/${PFX:-p16}_fwd -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/${PFX:-p16}_ncu_full.log 2>&1; echo "ncu full rc=$?"
N=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fused_kernel|auc_sampled_kernel" -c 8 -o gpurun_out/${PFX:-p16}_metrics -f \
  python tools/bench_metrics.py > gpurun_out/${PFX:-p16}_ncu_metrics.log 2>&1; echo "ncu metrics rc=$?"
cat gpurun_out/${PFX:-p16}_bench_inf.json gpurun_out/${PFX:-p16}_bench_train.json gpurun_out/${PFX:-p16}_bench_ref.json | cut -c1-600
