#!/bin/bash
# round 2: the driver's N=2 launch (torchrun, one rank per GPU): default bench line with the train sub-record (NCCL all-reduce), reference arm
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r02n2_bench.json 2> gpurun_out/r02n2_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r02n2_bench.err | cut -c1-300
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02n2_bench.json').read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value", d["value"], "e2e", d["e2e"]["value"])
t=d["train"]; print("train", t["value"], t["ms_per_step"], "allreduce_ms", t.get("allreduce_ms"), "e2e", t["e2e"]["value"])
print("config4", {k:(v.get("value"), v.get("persistent_recurrence")) for k,v in d["config4"].items() if isinstance(v,dict)})
print("config5", d["config5"].get("value"), d["config5"].get("error"))
PY
