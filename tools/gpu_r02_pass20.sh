#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "position_major" > gpurun_out/r02p20_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02p20_pytest.log | cut -c1-300
for pm in 0 1; do
  echo "RGP_PMAJOR=$pm"
  RGP_PMAJOR=$pm timeout 300 python bench.py --steps 10 --warmup 3 --no-sub --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['stage_ms'])"
done
RGP_PMAJOR=1 timeout 300 python tools/stage_times.py gaze_grcn 28 42 128 2>&1 | tail -1 | cut -c1-900
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02p20_pytest_all.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/r02p20_pytest_all.log | cut -c1-300
