#!/bin/bash
# round 2, second ncu pass: the kernels added after the first one
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for c in refshape cascade2 cfg4b64; do
  python tools/prof_cases.py $c 2 > gpurun_out/r02ncu2_plain_$c.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gru_small|lstm_seq_multi|topgru_seq|tapgemm_tc_kernel<16|tapgemm_tc_kernel<256, true>|maxout" -s 0 -c 8 -o gpurun_out/r02_prof2_$c -f python tools/prof_cases.py $c 1 > gpurun_out/r02ncu2_full_$c.log 2>&1
  echo "full $c rc=$?"; tail -2 gpurun_out/r02ncu2_plain_$c.log
done
ls -la gpurun_out/r02_prof2_*.ncu-rep
