#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02_final_bench_n8.json 2> gpurun_out/r02_final_bench_n8.err; echo "n8 rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r02_final_bench_n8.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['train']['value'], d['train'].get('allreduce_ms'))"
tail -3 gpurun_out/r02_final_bench_n8.err | cut -c1-300
