#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call8; mkdir -p $O
timeout 400 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method thread -k "rays or pool or edges or grid1m or spheres or quantised or wavefront" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
if not l.strip(): print('FAILED', end='  ')
else:
    d=json.loads(l); r=d['roofline']
    print(round(d['value']), 'B=%.1f T=%.1f' % (r['box_tests_per_ray'], r['tri_tests_per_ray']), end='  ')"; }
{
for q in 1 0; do for pipe in wavefront pool; do
echo -n "$pipe PTB_QUANT_NODES=$q spheres|grid|grid 1/8: "; PTB_QUANT_NODES=$q run spheres100k_1080p_256spp_d4 32 $pipe; PTB_QUANT_NODES=$q run grid1m_4k_64spp_d4 8 $pipe; PTB_QUANT_NODES=$q run grid1m_4k_8192spp_d4 64 $pipe --emulate-shard 8; echo
done; done
echo -n "auto pipeline, auto node format spheres|grid|grid 1/8: "; run spheres100k_1080p_256spp_d4 32 auto; run grid1m_4k_64spp_d4 8 auto; run grid1m_4k_8192spp_d4 64 auto --emulate-shard 8; echo
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
