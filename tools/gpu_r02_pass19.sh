#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for n in 1 2 4; do
  echo "RGP_NACC=$n"
  RGP_NACC=$n timeout 300 python bench.py --steps 10 --warmup 3 --no-sub --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['stage_ms'])"
  RGP_NACC=$n timeout 300 python tools/cascade_times.py 2>&1 | tail -1 | cut -c1-600
done
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02p19_pytest_all.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/r02p19_pytest_all.log | cut -c1-300
