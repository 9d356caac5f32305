#!/bin/bash
# persistent-recurrence iteration loop: parity tests of the kernel, then stage timings with it on / off
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_parity.py -x -q -k "persistent" > gpurun_out/seq_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/seq_pytest.log
B=256 timeout 60 python tools/seq_profile.py; B=128 TRAIN=1 timeout 60 python tools/seq_profile.py
for cfg in "inference 256" "train 128"; do
  set -- $cfg
  for ns in "1 1" "1 2" "1 4" "1 8" "2 1" "2 2" "2 4"; do
    set -- $cfg $ns; ns="$3 mcast=$4"
    RGP_SEQ_NSPLIT=$3 RGP_SEQ_MCAST=$4 timeout 90 python bench.py --mode $1 --batch $2 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1 B=$2 nsplit=$ns', 'ms/step %.3f' % d['ms_per_step'], {k: round(v,3) for k,v in d['roofline']['stage_ms'].items()})"
  done
done
