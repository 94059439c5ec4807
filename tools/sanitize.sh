#!/bin/bash
# compute-sanitizer over the small fixed workload (run on the GPU box: gpurun -- bash tools/sanitize.sh).
# memcheck: out-of-bounds / misaligned global, local and shared accesses and leaks of device allocations;
# racecheck: shared-memory hazards (the PLOC window of the device BVH build is the only shared-memory user);
# synccheck: the __syncwarp / __ballot_sync masks of the persistent traversal warps and the queue appends;
# initcheck: reads of device memory nobody wrote (queues, counters, path state).
# Logs land in gpurun_out/ (copy the summaries to profiles/).
set -u
mkdir -p gpurun_out
rc=0
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ "$tool" = memcheck ] && extra="--leak-check full"
  timeout 900 compute-sanitizer --tool $tool $extra --error-exitcode 3 --log-file gpurun_out/sanitize_$tool.log \
      python tools/sanitize_target.py > gpurun_out/sanitize_$tool.out 2>&1
  r=$?
  echo "$tool: exit $r; $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|LEAK SUMMARY' gpurun_out/sanitize_$tool.log | tr '\n' ' ')"
  [ $r -ne 0 ] && rc=$r
done
exit $rc
