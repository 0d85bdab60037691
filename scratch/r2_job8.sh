#!/bin/bash
set -u
O=gpurun_out/r2h; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_api.py tests/test_gpu_stages.py tests/test_gpu_retrieval.py -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log | cut -c1-300
timeout 200 python tests/sanitizer_smoke.py > $O/smoke_plain.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke_plain.log | cut -c1-300
timeout 400 python bench.py --steps 10 --warmup 3 --stages --no-cpu-baseline > $O/bench_msmt17.log 2>&1; tail -12 $O/bench_msmt17.log | cut -c1-330
timeout 300 python bench.py --workload duke --steps 20 --warmup 3 --stages --no-cpu-baseline > $O/bench_duke.log 2>&1; tail -11 $O/bench_duke.log | cut -c1-330
for tool in memcheck synccheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 python tests/sanitizer_smoke.py > $O/sanitizer_$tool.log 2>&1; echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitizer smoke|Error|error" $O/sanitizer_$tool.log | head -6 | cut -c1-250
done
