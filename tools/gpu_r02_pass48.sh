#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "persistent_recurrence" 2>&1 | tail -2
echo "--- cascade, nsplit auto (2)"; timeout 300 python tools/cascade_times.py 2>&1 | tail -1 | cut -c1-400
echo "--- cascade, nsplit 4"; RGP_SEQ_NSPLIT=4 timeout 300 python tools/cascade_times.py 2>&1 | tail -1 | cut -c1-400
for b in 32 64; do
echo "--- gaze_grcn inference $b clips, auto"; timeout 300 python bench.py --batch $b --steps 10 --warmup 3 --no-sub --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys;d=json.loads(sys.stdin.read());print(d['value'],d['ms_per_step'],d['roofline']['stage_ms'])"
echo "--- nsplit 4"; RGP_SEQ_NSPLIT=4 timeout 300 python bench.py --batch $b --steps 10 --warmup 3 --no-sub --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys;d=json.loads(sys.stdin.read());print(d['value'],d['ms_per_step'],d['roofline']['stage_ms'])"
done
