#!/bin/bash
# pool kernel with the hot / cold state split: GPU tests, then pool vs wavefront on the LBVH workloads and on rank 0's share of an 8-way tile split
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call14; mkdir -p $O
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; rc=$?; echo "smoke rc=$rc"; tail -2 $O/smoke.log
[ $rc -ne 0 ] && exit 1
timeout 600 python -m pytest tests -m gpu -q -x ${PYTEST_SEL:-} --timeout 200 --timeout-method thread > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
if not l.strip(): print('FAILED', end='  ')
else:
    d=json.loads(l); print(round(d['value']), end='  ')"; }
{
echo "# pool kernel, hot / cold state chunks (scripts/r02_call14.sh): Mrays/s, spheres100k 1080p x64 | grid1m 4K x16 | grid1m rank 0 of 8 x128 | spheres rank 0 of 8 x256"
for pipe in pool wavefront; do
  echo -n "$pipe: "; run spheres100k_1080p_256spp_d4 64 $pipe; run grid1m_4k_64spp_d4 16 $pipe; run grid1m_4k_8192spp_d4 128 $pipe --emulate-shard 8; run spheres100k_1080p_256spp_d4 256 $pipe --emulate-shard 8; echo
done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
