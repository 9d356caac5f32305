#!/bin/bash
# pass 28: ncu --set full of the training-only kernels (weight-gradient engine, per-step BPTT tap-GEMMs, batched small backward kernels)
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wgrad_tc_kernel" -s 4 -c 4 -o gpurun_out/p28_wgrad -f \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p28_ncu_wgrad.log 2>&1; echo "ncu wgrad rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tapgemm_tc_kernel<128" -s 40 -c 6 -o gpurun_out/p28_bptt -f \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p28_ncu_bptt.log 2>&1; echo "ncu bptt rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"bn_bwd|colsum|head_patches|cb_dw3f|clip_adam" -c 6 -o gpurun_out/p28_small -f \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p28_ncu_small.log 2>&1; echo "ncu small rc=$?"
ls -la gpurun_out/p28_*.ncu-rep
