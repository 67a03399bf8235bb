#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call7; mkdir -p $O
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; rc=$?; tail -2 $O/smoke.log
[ $rc -ne 0 ] && { tail -20 $O/smoke.log; exit 1; }
timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method thread > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest.log
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
if not l.strip(): print('FAILED', end='  ')
else:
    d=json.loads(l); r=d['roofline']; lb=d['detail']['lbvh'] or {}
    print(round(d['value']), 'B=%.1f T=%.1f depth=%s build=%.2fms' % (r['box_tests_per_ray'], r['tri_tests_per_ray'], lb.get('max_depth'), lb.get('build_ms', 0)), end='  ')"; }
{
for ploc in 1 0; do
  for pipe in wavefront pool; do
    echo -n "PLOC=$ploc $pipe spheres|grid: "; PTB_LBVH_PLOC=$ploc run spheres100k_1080p_256spp_d4 32 $pipe; PTB_LBVH_PLOC=$ploc run grid1m_4k_64spp_d4 8 $pipe; echo
  done
done
echo -n "PLOC=1 leaf 2|4 wavefront spheres: "; PTB_LBVH_LEAF=2 run spheres100k_1080p_256spp_d4 32 wavefront; PTB_LBVH_LEAF=4 run spheres100k_1080p_256spp_d4 32 wavefront; echo
echo -n "PLOC=1 shard 1/8 grid pool|wavefront: "; run grid1m_4k_8192spp_d4 64 pool --emulate-shard 8; run grid1m_4k_8192spp_d4 64 wavefront --emulate-shard 8; echo
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
