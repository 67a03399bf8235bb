#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call11; mkdir -p $O
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
print(round(json.loads(l)['value']) if l.strip() else 'FAILED', end='  ')"; }
{
for w in 1 0; do echo -n "pool whole_rounds=$w grid 1/8 x128spp | grid 1/4 | spheres full x64 | spheres 1/8 x256: "; PTB_POOL_WHOLE_ROUNDS=$w run grid1m_4k_8192spp_d4 128 pool --emulate-shard 8; PTB_POOL_WHOLE_ROUNDS=$w run grid1m_4k_8192spp_d4 64 pool --emulate-shard 4; PTB_POOL_WHOLE_ROUNDS=$w run spheres100k_1080p_256spp_d4 64 pool; PTB_POOL_WHOLE_ROUNDS=$w run spheres100k_1080p_256spp_d4 256 pool --emulate-shard 8; echo; done
for ns in 1 3 4; do echo -n "pool node_steps=$ns grid 1/8 | spheres full: "; PTB_POOL_NODE_STEPS=$ns run grid1m_4k_8192spp_d4 128 pool --emulate-shard 8; PTB_POOL_NODE_STEPS=$ns run spheres100k_1080p_256spp_d4 64 pool; echo; done
for sm in 8 24 32; do echo -n "pool shade_min=$sm grid 1/8 | spheres full: "; PTB_POOL_SHADE_MIN=$sm run grid1m_4k_8192spp_d4 128 pool --emulate-shard 8; PTB_POOL_SHADE_MIN=$sm run spheres100k_1080p_256spp_d4 64 pool; echo; done
for lm in 8 24; do echo -n "pool leaf_min=$lm grid 1/8 | spheres full: "; PTB_POOL_LEAF_MIN=$lm run grid1m_4k_8192spp_d4 128 pool --emulate-shard 8; PTB_POOL_LEAF_MIN=$lm run spheres100k_1080p_256spp_d4 64 pool; echo; done
echo -n "wavefront grid 1/8 | spheres full: "; run grid1m_4k_8192spp_d4 128 wavefront --emulate-shard 8; run spheres100k_1080p_256spp_d4 64 wavefront; echo
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
