#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
P=${1:-r02p32}
timeout 600 python -m pytest tests/test_gpu_metrics.py -m gpu -q -x > gpurun_out/${P}_metrics_tests.log 2>&1; echo "metrics tests rc=$?"; tail -8 gpurun_out/${P}_metrics_tests.log | cut -c1-400
RGP_JUDD_FUSED_KERNEL=old timeout 300 python tools/ab_judd_fused.py gpurun_out/${P}_old.npz 2>&1 | tail -1
timeout 300 python tools/ab_judd_fused.py gpurun_out/${P}_lean.npz 2>&1 | tail -1
python - <<Q
import numpy as np
a=np.load("gpurun_out/${P}_old.npz"); b=np.load("gpurun_out/${P}_lean.npz")
print({k: bool(np.array_equal(a[k], b[k], equal_nan=True)) for k in a.files})
Q
rm -f gpurun_out/${P}_*.npz
for mode in smooth noise; do
  RGP_SIM_FUSED_KERNEL=old timeout 300 python tools/ab_sim_fused.py gpurun_out/${P}_old_$mode.npy $mode 2>&1 | tail -1
  timeout 300 python tools/ab_sim_fused.py gpurun_out/${P}_lean_$mode.npy $mode 2>&1 | tail -1
  python - <<Q
import numpy as np
a=np.load("gpurun_out/${P}_old_$mode.npy"); b=np.load("gpurun_out/${P}_lean_$mode.npy")
print("$mode", "max|d| sim cc nss:", np.nanmax(np.abs(a[:,:3]-b[:,:3]),0), "min/max/nfix equal:", np.array_equal(a[:,3:6],b[:,3:6]), "sums rel:", np.max(np.abs(a[:,6:]-b[:,6:])/np.abs(a[:,6:])))
Q
done
rm -f gpurun_out/${P}_*.npy
timeout 600 ncu --set full --clock-control none --import-source on -k regex:auc_judd_lean -c 1 -o gpurun_out/${P}_judd -f python tools/ab_judd_fused.py /tmp/x.npz > gpurun_out/${P}_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sim_cc_nss_lean -c 1 -o gpurun_out/${P}_sim -f python tools/ab_sim_fused.py /tmp/x.npy > gpurun_out/${P}_ncu2.log 2>&1; echo "ncu rc=$?"
