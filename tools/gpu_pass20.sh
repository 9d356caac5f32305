#!/bin/bash
# pass 20: clock sampler check on the two headline lines + launch list of the config-4 ConvLSTM inference step
mkdir -p gpurun_out
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/p20_bench_inf.json 2> gpurun_out/p20_bench_inf.err; echo "inf rc=$?"
timeout 300 python bench.py --mode train --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/p20_bench_train.json 2> gpurun_out/p20_bench_train.err; echo "train rc=$?"
python - <<'PY'
import json
for f in ("inf", "train"):
    d = json.load(open("gpurun_out/p20_bench_%s.json" % f))
    print(f, round(d["value"]), round(d["ms_per_step"], 3), d["clocks"], round(d["e2e"]["value"]))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/p20_launches_cfg4_lstm.csv \
  python bench.py --model gaze_lstm --mode inference --batch 64 --T 128 --grid 14 --hidden 128 --entry emb --head fc --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/p20_ncu_cfg4.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = list(csv.reader(l for l in open("gpurun_out/p20_launches_cfg4_lstm.csv") if l.startswith('"')))
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        k = r[ki][:90]; v = float(r[vi].replace(",", ""))
    except Exception:
        continue
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print("%-90s n=%5d  total %9.1f us  avg %7.2f us  %5.1f%%" % (k, a[0], a[1] / 1e3, a[1] / 1e3 / a[0], 100 * a[1] / tot))
PY
