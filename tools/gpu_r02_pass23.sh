#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_cascade.py -m gpu -q -x > gpurun_out/r02p23_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02p23_pytest.log | cut -c1-400
for f in 0 1; do
  echo "RGP_CASCADE_FUSE=$f"
  RGP_CASCADE_FUSE=$f timeout 300 python tools/cascade_times.py 2>&1 | tail -1 | cut -c1-600
done
