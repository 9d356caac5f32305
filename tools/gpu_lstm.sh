#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "bptt or training or learns or entry_points" > gpurun_out/lstm_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/lstm_pytest.log
