#!/bin/bash
# compute-sanitizer evidence (on the GPU box): memcheck + racecheck of smoke(), memcheck of the peer / multi-device tests
TAG=${1:-rXX}
for tool in memcheck racecheck; do
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_sanitizer_${tool}_smoke.log 2>&1
  echo "$tool smoke: rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/${TAG}_sanitizer_${tool}_smoke.log | tail -1)"
done
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests -m gpu -q -x -k "fused_peer or rs_multi or list_and_trees" > gpurun_out/${TAG}_sanitizer_memcheck_tests.log 2>&1
echo "memcheck tests: rc=$? $(grep -E 'ERROR SUMMARY' gpurun_out/${TAG}_sanitizer_memcheck_tests.log | tail -1) $(grep -E 'passed|failed' gpurun_out/${TAG}_sanitizer_memcheck_tests.log | tail -1)"
timeout 900 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests -m gpu -q -x -k "fused_peer" > gpurun_out/${TAG}_sanitizer_racecheck_tests.log 2>&1
echo "racecheck tests: rc=$? $(grep -E 'RACECHECK SUMMARY' gpurun_out/${TAG}_sanitizer_racecheck_tests.log | tail -1) $(grep -E 'passed|failed' gpurun_out/${TAG}_sanitizer_racecheck_tests.log | tail -1)"
