#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02p24_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r02p24_alone.log | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02p24_pytest_all.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/r02p24_pytest_all.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
