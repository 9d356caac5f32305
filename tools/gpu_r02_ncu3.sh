#!/bin/bash
# round 2, third ncu pass: weight-gradient kernel, dense-row vs position-major k-steps
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for pm in 0 1; do
  RGP_PMAJOR=$pm python tools/prof_cases.py train128 1 > gpurun_out/r02ncu3_plain_$pm.log 2>&1 && \
  RGP_PMAJOR=$pm timeout 600 ncu --set full --clock-control none --import-source on -k regex:"wgrad_tc" -s 0 -c 4 -o gpurun_out/r02_prof3_wgrad_pm$pm -f python tools/prof_cases.py train128 1 > gpurun_out/r02ncu3_full_$pm.log 2>&1
  echo "pm=$pm rc=$?"
done
ls -la gpurun_out/r02_prof3_*.ncu-rep
