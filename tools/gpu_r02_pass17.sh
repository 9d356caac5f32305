#!/bin/bash
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 600 python -m pytest tests/test_gpu_baseline_configs.py -m gpu -q -x -k "lstm_bptt" > gpurun_out/r02p17_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02p17_pytest.log | cut -c1-300
python tools/stage_times.py gaze_lstm 28 42 128 2>&1 | tail -1 | tee -a gpurun_out/r02_stage_times.log | cut -c1-900
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02p17_pytest_all.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/r02p17_pytest_all.log | cut -c1-300
