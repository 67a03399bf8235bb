#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call2; mkdir -p $O
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; rc=$?; tail -3 $O/smoke.log
[ $rc -ne 0 ] && exit 1
timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method thread --deselect tests/test_gpu_multi.py -k "rays or pool or output or grid1m or wavefront_parity or spheres" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
run() { timeout 120 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline pool --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), 'B=%.1f T=%.1f' % (r['box_tests_per_ray'], r['tri_tests_per_ray']), end='  ')"; }
two() { run spheres100k_1080p_256spp_d4 32; run grid1m_4k_64spp_d4 8; echo; }
{
echo -n "wavefront: "; timeout 120 python bench.py --workload spheres100k_1080p_256spp_d4 --workloads none --spp 32 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; print(round(json.loads(sys.stdin.readline())['value']))"
echo -n "pool, quantised nodes via tex (default): "; two
for v in qf0 qf1 ns64; do echo -n "variant $v: "; PTB200_LIB=$PWD/build_ab/libptb200_$v.so two; done
echo -n "pool, float nodes: "; PTB_QUANT_NODES=0 two
echo -n "pool, quantised, node_steps=3: "; PTB_POOL_NODE_STEPS=3 two
echo -n "pool, quantised, node_steps=3 shade_min=24: "; PTB_POOL_NODE_STEPS=3 PTB_POOL_SHADE_MIN=24 two
echo -n "pool, quantised, 6 blocks/SM: "; PTB_POOL_BLOCKS=6 two
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
