#!/bin/bash
# Memory-safety pass over every kernel of the library, on 64x64 frames of six scenes x every traversal x
# both wavefront formulations x both strategies (tools/sanitize_run.py).  Run on the GPU box from the repo root:
#   bash tools/sanitize.sh            -> gpurun_out/sanitize_*.log   (summaries are copied to profiles/)
# 1. compute-sanitizer memcheck / racecheck / initcheck / synccheck where the pool allows it (this pool answers
#    "compute-sanitizer is closed on this pool"; the attempt and its answer are logged);
# 2. the library built with -DPT_DEBUG_CHECKS (device-side asserts on queue slots, list positions, object ids, light picks and
#    traversal stacks: pt_math.cuh PT_CHECK) under the same workload AND under the whole `-m gpu` parity suite: an assert that
#    fires aborts the context, so a green run means none fired.
mkdir -p gpurun_out
rc=0
for tool in memcheck racecheck initcheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 3 python tools/sanitize_run.py > gpurun_out/sanitize_$tool.log 2>&1
  r=$?
  echo "== compute-sanitizer $tool exit $r: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|closed on this pool' gpurun_out/sanitize_$tool.log | tail -1 | cut -c1-120)"
done
DBG=$PWD/pathtrace_b200/lib/libpt_debug.so
if [ ! -f $DBG ]; then PT_LIB=$DBG PT_NVCC_EXTRA="-DPT_DEBUG_CHECKS" python pathtrace_b200/build.py --force > /dev/null; python pathtrace_b200/build.py --force > /dev/null; fi
PT_LIB=$DBG timeout 900 python tools/sanitize_run.py > gpurun_out/sanitize_debug_checks.log 2>&1; r=$?
echo "== PT_DEBUG_CHECKS workload exit $r: $(tail -1 gpurun_out/sanitize_debug_checks.log)"; [ $r -ne 0 ] && rc=$r
PT_LIB=$DBG timeout 1500 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/sanitize_debug_checks_pytest.log 2>&1; r=$?
echo "== PT_DEBUG_CHECKS parity suite exit $r: $(tail -1 gpurun_out/sanitize_debug_checks_pytest.log)"; [ $r -ne 0 ] && rc=$r
exit $rc
