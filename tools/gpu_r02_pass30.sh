#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
P=${1:-r02p30}
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${P}_pytest_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/${P}_pytest_all.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/${P}_bench_n1.json 2> gpurun_out/${P}_bench_n1.err; echo "bench rc=$?"; tail -c 600 gpurun_out/${P}_bench_n1.json
