#!/bin/bash
# pass 21: ConvLSTM gate-backward v4 parity + bench, then ncu launch lists of the two training steps (never captured before)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_models.py -x -q -k "fc_head or grid14 or train or bptt or lstm or adam" > gpurun_out/p21_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p21_pytest.log
timeout 300 python bench.py --model gaze_lstm --mode train --batch 32 --T 128 --grid 14 --hidden 128 --entry emb --head fc --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/cfg4_gaze_lstm_train_B32.json 2> gpurun_out/cfg4_gaze_lstm_train_B32.err; echo "cfg4 train rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/cfg4_gaze_lstm_train_B32.json')); print('frames/s %.0f' % d['value'], 'ms/step %.2f' % d['ms_per_step'], {k: round(v,3) for k,v in d['roofline']['stage_ms'].items()}, d['roofline'].get('stage_ms_bwd'))"
agg() {
python - "$1" <<'PY'
import csv, collections, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        k = r[ki][:100]; v = float(r[vi].replace(",", ""))
    except Exception:
        continue
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
    print("%-100s n=%5d  total %9.1f us  avg %7.2f us  %5.1f%%" % (k, a[0], a[1] / 1e3, a[1] / 1e3 / a[0], 100 * a[1] / tot))
PY
}
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/p21_launches_train_B128.csv \
  python bench.py --mode train --batch 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/p21_ncu_train.log 2>&1; echo "ncu train rc=$?"
agg gpurun_out/p21_launches_train_B128.csv
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/p21_launches_cfg4_lstm_train.csv \
  python bench.py --model gaze_lstm --mode train --batch 32 --T 128 --grid 14 --hidden 128 --entry emb --head fc --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/p21_ncu_cfg4_train.log 2>&1; echo "ncu cfg4 train rc=$?"
agg gpurun_out/p21_launches_cfg4_lstm_train.csv
