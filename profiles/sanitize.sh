#!/bin/bash
# compute-sanitizer evidence (on the GPU box): memcheck + racecheck of smoke(), memcheck of the small-world / generic-tier
# tests, racecheck of the passes test.  Every leg is bounded; the summaries land in gpurun_out/${TAG}_sanitizer_*.log
TAG=${1:-rXX}
for tool in memcheck racecheck; do
  timeout 400 compute-sanitizer --tool $tool --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_sanitizer_${tool}_smoke.log 2>&1
  echo "$tool smoke: rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/${TAG}_sanitizer_${tool}_smoke.log | tail -1)"
done
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests -m gpu -q -x -k "${SANITIZE_K:-list_and_trees or passes or multi_light or mixed}" > gpurun_out/${TAG}_sanitizer_memcheck_tests.log 2>&1
echo "memcheck tests: rc=$? $(grep -E 'ERROR SUMMARY' gpurun_out/${TAG}_sanitizer_memcheck_tests.log | tail -1) $(grep -E 'passed|failed' gpurun_out/${TAG}_sanitizer_memcheck_tests.log | tail -1)"
