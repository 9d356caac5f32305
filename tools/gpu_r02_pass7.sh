#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_models.py -m gpu -q -x -k dense > gpurun_out/r02p7_dense.log 2>&1; echo "dense rc=$?"; tail -25 gpurun_out/r02p7_dense.log | cut -c1-300
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02p7_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02p7_pytest.log | cut -c1-300
