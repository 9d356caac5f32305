#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
RGP_SIM_FUSED_KERNEL=old timeout 300 python tools/ab_sim_fused.py gpurun_out/q_old.npy smooth 2>&1 | tail -1
for i in 1 2; do timeout 300 python tools/ab_sim_fused.py gpurun_out/q_lean.npy smooth 2>&1 | tail -1; done
python - <<Q
import numpy as np
a=np.load("gpurun_out/q_old.npy"); b=np.load("gpurun_out/q_lean.npy")
print("max|d| sim cc nss:", np.nanmax(np.abs(a[:,:3]-b[:,:3]),0), "min/max/nfix equal:", np.array_equal(a[:,3:6],b[:,3:6]))
Q
rm -f gpurun_out/q_*.npy
