#!/bin/bash
# two-GPU check of HEAD: reference arm, inference and training lines under torchrun exactly as the driver launches them
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/n8_ref.json 2> gpurun_out/n8_ref.err; echo "ref rc=$?"
timeout 600 $TR bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/n8_inf.json 2> gpurun_out/n8_inf.err; echo "inf rc=$?"
timeout 600 $TR bench.py --gpus 8 --mode train --batch 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/n8_train.json 2> gpurun_out/n8_train.err; echo "train rc=$?"
python - <<'PY'
import json
for f in ("ref", "inf", "train"):
    try:
        d = json.loads(open("gpurun_out/n8_%s.json" % f).read().strip().splitlines()[-1])
        print(f, d.get("n_gpus"), round(d["value"]), round(d["ms_per_step"], 3), d.get("clocks"), (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/n8_%s.err" % f).read()[-1500:])
PY
