#!/bin/bash
# round 2: the driver's N=8 launch
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 > gpurun_out/r02n8_bench.json 2> gpurun_out/r02n8_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r02n8_bench.err | cut -c1-300
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02n8_bench.json').read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "e2e fp32", d["e2e_fp32_features"]["value"])
t=d["train"]; print("train", t["value"], t["ms_per_step"], "allreduce_ms", t.get("allreduce_ms"), "e2e", t["e2e"]["value"])
print("numa", d.get("numa")); print("clocks", d.get("clocks"))
PY
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/r02n8_ref.json 2> gpurun_out/r02n8_ref.err; echo "ref rc=$?"; head -c 300 gpurun_out/r02n8_ref.json
