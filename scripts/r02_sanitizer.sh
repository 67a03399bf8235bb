#!/bin/bash
# compute-sanitizer over the small-size GPU tests.  ONE tool per gpurun call (B200_PROFILING.md): $1 = memcheck | racecheck | synccheck
# memcheck: every small test incl. the pool kernel, the wavefront queues, the LBVH build, filt_im, the resolve; racecheck / synccheck: the kernels
# that communicate through shared memory and warp votes (pool kernel, wavefront push, LBVH refit).  Summaries -> profiles/r02_sanitizer_<tool>.txt
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
TOOL=${1:-memcheck}
O=gpurun_out/sanitizer; mkdir -p $O
if [ "$TOOL" = memcheck ]; then
  SEL='tests/test_gpu_edges.py tests/test_gpu_output.py'; K='not 1920 and not grid1m and not c1_'
  EXTRA="--leak-check no"
else
  SEL='tests/test_gpu_edges.py'; K='ragged or mirror or bounce or duplicate'
  EXTRA=""
  [ "$TOOL" = racecheck ] && EXTRA="--racecheck-report all"
fi
timeout 1500 compute-sanitizer --tool $TOOL $EXTRA --error-exitcode 9 --log-file $O/$TOOL.log python -m pytest $SEL -q -m gpu -x -p no:cacheprovider -k "$K" > $O/$TOOL.pytest 2>&1; echo "$TOOL rc=$?"
echo "== $TOOL log"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|Error|error" $O/$TOOL.log | sort | uniq -c | sort -rn | head -12
echo "== pytest"; tail -3 $O/$TOOL.pytest
