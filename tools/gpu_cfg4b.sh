#!/bin/bash
# config-4 family after the batched FC-head backward: parity of the new kernel, training parity tests, then the bench lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "fc_head or grid14 or train or bptt or lstm" > gpurun_out/cfg4b_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/cfg4b_pytest.log
for cfg in "gaze_lstm train 32" "gaze_grcn77 train 32" "gaze_lstm inference 64"; do
  set -- $cfg
  timeout 300 python bench.py --model $1 --mode $2 --batch $3 --T 128 --grid 14 --hidden 128 --entry emb --head fc --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/cfg4_$1_$2_B$3.json 2> gpurun_out/cfg4_$1_$2_B$3.err; echo "$cfg rc=$?"
  python -c "
import json,sys
d=json.load(open('gpurun_out/cfg4_$1_$2_B$3.json')); print('$cfg', 'frames/s %.0f' % d['value'], 'ms/step %.2f' % d['ms_per_step'], 'e2e %.0f' % d['e2e']['value'], {k: round(v,3) for k,v in d['roofline']['stage_ms'].items()}, d['roofline'].get('stage_ms_bwd'))" || tail -5 gpurun_out/cfg4_$1_$2_B$3.err
done
