#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sim_cc_nss_lean -c 2 -o gpurun_out/r02p26_lean -f python tools/ab_sim_fused.py gpurun_out/r02p26_lean.npy > gpurun_out/r02p26_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r02p26_ncu.log
ls -la gpurun_out/*.ncu-rep
