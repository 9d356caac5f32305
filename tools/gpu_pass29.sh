#!/bin/bash
# pass 29: ncu --set full of the config-4 ConvLSTM element-wise kernels and of per-step BPTT tap-GEMM launches of the headline training step
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none -k regex:"lstm_gates|fc_head" -s 6 -c 6 -o gpurun_out/p29_lstm -f \
  python bench.py --model gaze_lstm --mode train --batch 32 --T 128 --grid 14 --hidden 128 --entry emb --head fc --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p29_ncu_lstm.log 2>&1; echo "ncu lstm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tapgemm_tc_kernel" -s 120 -c 6 -o gpurun_out/p29_bptt -f \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/p29_ncu_bptt.log 2>&1; echo "ncu bptt rc=$?"
ls -la gpurun_out/p29_*.ncu-rep
