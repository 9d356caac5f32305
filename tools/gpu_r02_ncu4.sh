#!/bin/bash
# round 2, fourth ncu pass: position-major i2h / head (launch list + --set full of the inference step's kernels), then the default bench
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
CMD="python bench.py --no-sub --no-cpu-baseline --no-e2e --steps 2 --warmup 1"
$CMD > gpurun_out/r02ncu4_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_inference_B256_pm.csv $CMD > gpurun_out/r02ncu4_list.log 2>&1
echo "list rc=$?"
python tools/prof_cases.py fwd256 2 > gpurun_out/r02ncu4_plain_fwd256.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tapgemm_tc_kernel|gru_seq_kernel|pack_c3d" -s 0 -c 7 -o gpurun_out/r02_prof4_fwd256 -f python tools/prof_cases.py fwd256 1 > gpurun_out/r02ncu4_full_fwd256.log 2>&1
echo "full rc=$?"
timeout 900 python bench.py > gpurun_out/r02_bench_n1_pass22.json 2> gpurun_out/r02_bench_n1_pass22.err; echo "bench rc=$?"; tail -1 gpurun_out/r02_bench_n1_pass22.json | cut -c1-600
