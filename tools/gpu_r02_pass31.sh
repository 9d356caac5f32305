#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
P=${1:-r02p31}
timeout 600 python -m pytest tests/test_gpu_metrics.py -m gpu -q -x > gpurun_out/${P}_metrics_tests.log 2>&1; echo "metrics tests rc=$?"; tail -8 gpurun_out/${P}_metrics_tests.log | cut -c1-400
RGP_JUDD_FUSED_KERNEL=old timeout 300 python tools/ab_judd_fused.py gpurun_out/${P}_old.npz 2>&1 | tail -1
timeout 300 python tools/ab_judd_fused.py gpurun_out/${P}_lean.npz 2>&1 | tail -1
python - <<Q
import numpy as np
a=np.load("gpurun_out/${P}_old.npz"); b=np.load("gpurun_out/${P}_lean.npz")
for k in a.files: print(k, np.array_equal(a[k], b[k], equal_nan=True))
Q
rm -f gpurun_out/${P}_*.npz
timeout 600 ncu --set full --clock-control none --import-source on -k regex:auc_judd_lean -c 1 -o gpurun_out/${P}_judd -f python tools/ab_judd_fused.py /tmp/x.npz > gpurun_out/${P}_ncu.log 2>&1; echo "ncu rc=$?"
