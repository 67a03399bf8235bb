#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call6; mkdir -p $O
timeout 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x --timeout 100 --timeout-method thread > $O/noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"; tail -3 $O/noalloc_alone.log
timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method thread > $O/pytest.log 2>&1; echo "pytest all rc=$?"; tail -6 $O/pytest.log
run() { timeout 60 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
print(round(json.loads(l)['value']) if l.strip() else 'FAILED', end='  ')"; }
{
for r in 1 4 8 12 16 24; do echo -n "cornell regen_min=$r: "; PTB_REGEN_MIN=$r run cornell_1080p_1024spp_d4 256 auto; echo; done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
