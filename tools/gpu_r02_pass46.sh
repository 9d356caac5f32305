#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_cascade.py -m gpu -q -x 2>&1 | tail -3
RGP_TOPGRU_SPLIT=1 timeout 600 python -m pytest tests/test_gpu_cascade.py -m gpu -q -x 2>&1 | tail -2
echo "--- one CTA per clip"; RGP_TOPGRU_SPLIT=1 timeout 300 python tools/cascade_times.py 2>&1 | tail -4
echo "--- two CTAs per clip"; timeout 300 python tools/cascade_times.py 2>&1 | tail -4
