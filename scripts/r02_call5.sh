#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call5; mkdir -p $O
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; rc=$?; tail -2 $O/smoke.log
[ $rc -ne 0 ] && exit 1
timeout 900 python -m pytest tests -m gpu -q --timeout 240 --timeout-method thread > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest.log
run() { timeout 150 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(round(d['value']), end='  ')"; }
{
for sh in 2 4 8; do
  echo -n "grid1m 4K shard 1/$sh pool chunked: "; run grid1m_4k_8192spp_d4 128 pool --emulate-shard $sh; echo -n " unchunked: "; PTB_POOL_ROUNDS=0 run grid1m_4k_8192spp_d4 128 pool --emulate-shard $sh; echo -n " wavefront: "; run grid1m_4k_8192spp_d4 128 wavefront --emulate-shard $sh; echo
done
echo -n "spheres 1080p full pool chunked/unchunked: "; run spheres100k_1080p_256spp_d4 64 pool; PTB_POOL_ROUNDS=0 run spheres100k_1080p_256spp_d4 64 pool; echo
echo -n "spheres 1080p shard 1/8 auto: "; run spheres100k_1080p_256spp_d4 256 auto --emulate-shard 8; echo
for r in 1 4 8 12 16 24; do echo -n "cornell regen_min=$r: "; PTB_REGEN_MIN=$r run cornell_1080p_1024spp_d4 256 auto; echo; done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
