#!/bin/bash
# Memory-safety run of the small fixed workload (tools/sanitize_target.py) on the GPU box:
#   gpurun -- bash tools/sanitize.sh
# 1. The library's own checked build (libakari_b200_debug.so = -DAKB_DEBUG_CHECKS -DAKB_DEBUG_TRAVERSAL, built by
#    __graft_entry__.build()): every node / leaf-triangle / queue / slot index is bounds-checked on the device and the
#    traversal stack traps on overflow; a violation is a CUDA error on the host, and the target also compares every
#    result with the unchecked expectations (wavefront = megakernel = sorted, any-hit = closest-hit).
# 2. compute-sanitizer (memcheck, racecheck, synccheck, initcheck) where the pool allows it. On this round's pool it
#    answers "compute-sanitizer is closed on this pool and stays closed" (exit 86) - recorded in profiles/r02_sanitize.md.
set -u
mkdir -p gpurun_out
rc=0
AKB_LIB=$PWD/akarirender-1_b200/libakari_b200_debug.so timeout 600 python tools/sanitize_target.py > gpurun_out/sanitize_debug_build.out 2>&1
r=$?; echo "checked build: exit $r; $(tail -1 gpurun_out/sanitize_debug_build.out)"; [ $r -ne 0 ] && rc=$r
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ "$tool" = memcheck ] && extra="--leak-check full"
  timeout 900 compute-sanitizer --tool $tool $extra --error-exitcode 3 --log-file gpurun_out/sanitize_$tool.log \
      python tools/sanitize_target.py > gpurun_out/sanitize_$tool.out 2>&1
  r=$?
  echo "$tool: exit $r; $(head -c 300 gpurun_out/sanitize_$tool.out | tr '\n' ' ') $(grep -sE 'ERROR SUMMARY|RACECHECK SUMMARY|LEAK SUMMARY' gpurun_out/sanitize_$tool.log | tr '\n' ' ')"
  [ $r -ne 0 ] && [ $r -ne 86 ] && rc=$r
done
exit $rc
