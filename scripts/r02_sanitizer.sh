#!/bin/bash
# compute-sanitizer over the small-size GPU tests: memcheck on everything small, racecheck + synccheck on the
# kernels that communicate through shared memory / ballots (pool kernel, wavefront queues, LBVH build); summaries -> profiles/
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/sanitizer; mkdir -p $O
T="tests/test_gpu_edges.py"
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 --log-file $O/memcheck_edges.log python -m pytest $T -q -m gpu -x -p no:cacheprovider > $O/memcheck_edges.pytest 2>&1; echo "memcheck edges rc=$?"
timeout 900 compute-sanitizer --tool racecheck --racecheck-report all --error-exitcode 9 --log-file $O/racecheck_edges.log python -m pytest $T -q -m gpu -x -p no:cacheprovider -k "ragged or mirror or bounce" > $O/racecheck_edges.pytest 2>&1; echo "racecheck edges rc=$?"
timeout 900 compute-sanitizer --tool synccheck --error-exitcode 9 --log-file $O/synccheck_edges.log python -m pytest $T -q -m gpu -x -p no:cacheprovider -k "ragged or mirror" > $O/synccheck_edges.pytest 2>&1; echo "synccheck edges rc=$?"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 --log-file $O/memcheck_out.log python -m pytest tests/test_gpu_output.py -q -m gpu -x -p no:cacheprovider -k "filt_im and not 1920 or texture or resolve or checkpoint" > $O/memcheck_out.pytest 2>&1; echo "memcheck output rc=$?"
for f in $O/*.log; do echo "== $f"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|error" $f | sort | uniq -c | head -8; done
for f in $O/*.pytest; do echo "== $f"; tail -2 $f; done
