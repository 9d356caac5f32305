#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 420 python -m pytest tests/test_gpu_baseline_configs.py -m gpu -q -x -k "persistent_convlstm or config4" > gpurun_out/r02p6_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02p6_pytest.log | cut -c1-300
for B in 8 2; do timeout 120 python tools/lstm_seq_timeline.py $B 14 128 > gpurun_out/r02p6_timeline_B$B.json 2> gpurun_out/r02p6_timeline_B$B.err; echo "timeline rc=$?"; cat gpurun_out/r02p6_timeline_B$B.json | cut -c1-1200; done
timeout 120 python tools/lstm_seq_timeline.py 28 7 35 > gpurun_out/r02p6_timeline_G7.json 2>&1; cat gpurun_out/r02p6_timeline_G7.json | cut -c1-1200
