#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests/test_gpu_metrics.py -m gpu -q > gpurun_out/r02p36_metrics_tests.log 2>&1; echo "metrics tests rc=$?"; tail -30 gpurun_out/r02p36_metrics_tests.log | cut -c1-300
timeout 300 python tools/prof_eval_host.py 2>&1 | grep -A4 "rng ="
