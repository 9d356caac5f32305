#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
run() { timeout 300 python bench.py --hidden 128 --batch $1 --steps 10 --warmup 3 --no-sub --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys;d=json.loads(sys.stdin.read());print(round(d['value']),round(d['ms_per_step'],3),'rec',round(d['roofline']['stage_ms']['recurrence'],3))"; }
for b in 48 96; do
  echo "--- hidden 128, $b clips: per-step"; RGP_SEQ=0 run $b
  echo "    seq nsplit 1"; RGP_SEQ=1 RGP_SEQ_NSPLIT=1 run $b
  echo "    seq nsplit 2"; RGP_SEQ=1 RGP_SEQ_NSPLIT=2 run $b
done
