#!/bin/bash
# round 2, pass 3: cascade GPU parity first (new kernels), then the whole GPU suite
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_cascade.py -m gpu -q -x > gpurun_out/r02p3_cascade.log 2>&1; echo "cascade rc=$?"; tail -30 gpurun_out/r02p3_cascade.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_cascade.py > gpurun_out/r02p3_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02p3_pytest.log
