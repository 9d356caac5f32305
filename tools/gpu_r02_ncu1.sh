#!/bin/bash
# round 2 ncu pass: launch list of the default inference step + --set full captures of the round's kernels
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
CMD="python bench.py --no-sub --no-cpu-baseline --no-e2e --steps 2 --warmup 1"
$CMD > gpurun_out/r02ncu_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_inference_B256.csv $CMD > gpurun_out/r02ncu_list.log 2>&1
echo "list rc=$?"
for c in fwd256 cfg4b8 cascade; do
  python tools/prof_cases.py $c 2 > gpurun_out/r02ncu_plain_$c.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tapgemm_tc_kernel|gru_seq_kernel|lstm_seq_kernel|topgru_seq_kernel|pack_c3d" -s 0 -c 14 -o gpurun_out/r02_prof_$c -f python tools/prof_cases.py $c 1 > gpurun_out/r02ncu_full_$c.log 2>&1
  echo "full $c rc=$?"
done
ls -la gpurun_out/*.ncu-rep
