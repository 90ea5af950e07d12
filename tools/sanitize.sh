#!/bin/bash
# compute-sanitizer on smoke() (one small frame through the C ABI, checked against the oracle): memcheck, racecheck,
# initcheck and synccheck logs into gpurun_out/<tag>_sanitizer_*.log (copied to profiles/ when clean).
#   gpurun --timeout 900 -- 'bash tools/sanitize.sh r02'
tag=${1:-san}
o=gpurun_out
mkdir -p $o
for tool in memcheck racecheck initcheck synccheck; do
  timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 python __graft_entry__.py smoke > $o/${tag}_sanitizer_$tool.log 2>&1
  echo "sanitizer $tool rc $?"
  tail -3 $o/${tag}_sanitizer_$tool.log
done
