#!/bin/bash
# round 2: the driver's N=2 launch (torchrun, one rank per GPU): default bench line with the train sub-record (NCCL all-reduce), reference arm
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_optimizer.py tests/test_gpu_baseline_configs.py -m gpu -q > gpurun_out/r02n2_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02n2_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r02n2_bench.json 2> gpurun_out/r02n2_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r02n2_bench.err; head -c 1500 gpurun_out/r02n2_bench.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r02n2_ref.json 2> gpurun_out/r02n2_ref.err; echo "ref rc=$?"; head -c 400 gpurun_out/r02n2_ref.json
