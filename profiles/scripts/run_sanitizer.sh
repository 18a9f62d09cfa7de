#!/bin/bash
cd $GRAFT_REPO_ROOT
for tool in memcheck racecheck; do
  AMCLB200_NO_GRAPH=1 timeout 1200 compute-sanitizer --tool $tool --print-limit 20 python profiles/scripts/sanitize_case.py > gpurun_out/r2_sanitizer_$tool.log 2>&1
  echo "rc $?" >> gpurun_out/r2_sanitizer_$tool.log
done
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python profiles/scripts/sanitize_case.py > gpurun_out/r2_sanitizer_memcheck_graph.log 2>&1
echo "rc $?" >> gpurun_out/r2_sanitizer_memcheck_graph.log
tail -5 gpurun_out/r2_sanitizer_*.log
