#!/bin/bash
# which pipeline for which share of the 4K frame, after the wavefront's state split: rank 0's share of an N-way tile split (--emulate-shard N)
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call16; mkdir -p $O
run() { timeout 100 python bench.py --workload $1 --workloads none --spp $2 --steps 3 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
l=sys.stdin.readline()
if not l.strip(): print('FAILED', end='  ')
else:
    d=json.loads(l); print(round(d['value']), end='  ')"; }
{
echo "# grid1m 4K, rank 0's share of an N-way tile split, Mrays/s: pool | wavefront 2 sub-wavefronts | 1 | 3 (scripts/r02_call16.sh)"
for n in 2 4 8; do
  echo -n "1/$n: "; run grid1m_4k_8192spp_d4 $((16*n)) pool --emulate-shard $n; run grid1m_4k_8192spp_d4 $((16*n)) wavefront --emulate-shard $n; PTB_WF_SUBS=1 run grid1m_4k_8192spp_d4 $((16*n)) wavefront --emulate-shard $n; PTB_WF_SUBS=3 run grid1m_4k_8192spp_d4 $((16*n)) wavefront --emulate-shard $n; echo
done
echo "# spheres100k 1080p"
for n in 2 4; do
  echo -n "1/$n: "; run spheres100k_1080p_256spp_d4 $((32*n)) pool --emulate-shard $n; run spheres100k_1080p_256spp_d4 $((32*n)) wavefront --emulate-shard $n; PTB_WF_SUBS=1 run spheres100k_1080p_256spp_d4 $((32*n)) wavefront --emulate-shard $n; echo
done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
