#!/bin/bash
# round 2, pass 2: full GPU test suite (no -x), small-case kernel checks (compute-sanitizer is closed on this pool)
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02p2_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02p2_pytest.log
timeout 300 python tools/sanitize_cases.py tapgemm256 tapgemm128 head48 wgrad seq_1_1 seq_2_1 seq_1_4 seq_2_4 bptt lstm > gpurun_out/r02p2_cases.log 2>&1; echo "cases rc=$?"; tail -12 gpurun_out/r02p2_cases.log
