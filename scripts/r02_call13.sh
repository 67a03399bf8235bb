#!/bin/bash
# PLOC search radius 8 / 16 / 32 (PTB_PLOC_R) and registers per thread of k_wf_shade after the state split (WF_SHADE_MINB 3 / 4 / 5 / 6)
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call13; mkdir -p $O
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
if not l.strip(): print('FAILED', end='  ')
else:
    d=json.loads(l); r=d['roofline']; lb=d['detail']['lbvh'] or {}
    print(round(d['value']), 'B=%.1f T=%.1f depth=%s build=%.2fms' % (r['box_tests_per_ray'], r['tri_tests_per_ray'], lb.get('max_depth'), lb.get('build_ms', 0)), end='  ')"; }
{
echo "# PLOC search radius (PTB_PLOC_R) and WF_SHADE_MINB variants: Mrays/s, box / triangle tests per ray; spheres100k 1080p x32 spp | grid1m 4K x8 spp | grid1m rank 0 of 8 (pool) (scripts/r02_call13.sh)"
for r in 8 16 32; do
  echo -n "PLOC_R=$r wavefront spheres|grid|pool grid 1/8: "; PTB_PLOC_R=$r run spheres100k_1080p_256spp_d4 32 wavefront; PTB_PLOC_R=$r run grid1m_4k_64spp_d4 8 wavefront; PTB_PLOC_R=$r run grid1m_4k_8192spp_d4 64 pool --emulate-shard 8; echo
done
for v in 3 5 6; do
  echo -n "WF_SHADE_MINB=$v wavefront spheres|grid: "; PTB200_LIB=build_ab/libptb200_shade$v.so run spheres100k_1080p_256spp_d4 32 wavefront; PTB200_LIB=build_ab/libptb200_shade$v.so run grid1m_4k_64spp_d4 8 wavefront; echo
done
echo -n "WF_SHADE_MINB=4 (shipping) wavefront spheres|grid: "; run spheres100k_1080p_256spp_d4 32 wavefront; run grid1m_4k_64spp_d4 8 wavefront; echo
for s in 1 3 4; do
  echo -n "PTB_WF_SUBS=$s wavefront spheres|grid: "; PTB_WF_SUBS=$s run spheres100k_1080p_256spp_d4 32 wavefront; PTB_WF_SUBS=$s run grid1m_4k_64spp_d4 8 wavefront; echo
done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
