#!/bin/bash
# A/B in one run: pool kernel with the hot / cold state chunks (shipping build) against the previous three-chunk state (build_ab/libptb200_oldpool.so)
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call15; mkdir -p $O
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 3 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
if not l.strip(): print('FAILED', end='  ')
else:
    d=json.loads(l); print(round(d['value']), end='  ')"; }
{
echo "# pool kernel state chunks A/B (scripts/r02_call15.sh): Mrays/s, spheres100k 1080p x64 | grid1m 4K x16 | grid1m rank 0 of 8 x128 | spheres rank 0 of 8 x256"
for rep in 1 2; do
echo -n "new (2 hot + 1 cold, cold stored when dirty): "; run spheres100k_1080p_256spp_d4 64 pool; run grid1m_4k_64spp_d4 16 pool; run grid1m_4k_8192spp_d4 128 pool --emulate-shard 8; run spheres100k_1080p_256spp_d4 256 pool --emulate-shard 8; echo
echo -n "old (3 chunks loaded and stored every step):  "; export PTB200_LIB=build_ab/libptb200_oldpool.so; run spheres100k_1080p_256spp_d4 64 pool; run grid1m_4k_64spp_d4 16 pool; run grid1m_4k_8192spp_d4 128 pool --emulate-shard 8; run spheres100k_1080p_256spp_d4 256 pool --emulate-shard 8; echo; unset PTB200_LIB
done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
