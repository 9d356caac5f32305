#!/bin/bash
# round 2, pass 1: full GPU test suite, smoke, default bench line (with sub-records), reference arm, sanitizer runs
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02p1_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02p1_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02p1_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r02p1_smoke.log
timeout 900 python bench.py > gpurun_out/r02p1_bench.json 2> gpurun_out/r02p1_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r02p1_bench.err; head -c 3000 gpurun_out/r02p1_bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02p1_ref.json 2> gpurun_out/r02p1_ref.err; echo "ref rc=$?"; head -c 600 gpurun_out/r02p1_ref.json
for tool in memcheck synccheck racecheck; do
  timeout 240 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py tapgemm256 tapgemm128 head48 wgrad seq_1_1 seq_2_1 seq_1_4 seq_2_4 bptt lstm > gpurun_out/r02p1_san_$tool.log 2>&1
  echo "sanitizer $tool rc=$?"; grep -E "ERROR SUMMARY|case .* ok|RACECHECK SUMMARY" gpurun_out/r02p1_san_$tool.log | tail -14
done
