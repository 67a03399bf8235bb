#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call3; mkdir -p $O
run() { timeout 150 python bench.py --workload $1 --workloads none --spp $2 --steps 2 --warmup 3 --pipeline $3 --no-cpu-baseline --no-e2e ${@:4} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(round(d['value']), end='  ')"; }
{
for sh in 1 2 4 8; do
  for pipe in wavefront pool; do echo -n "grid1m 4K shard 1/$sh $pipe: "; run grid1m_4k_8192spp_d4 $((8*sh)) $pipe --emulate-shard $sh; echo; done
done
for sh in 2 8; do
  for pipe in wavefront pool; do echo -n "spheres 1080p shard 1/$sh $pipe: "; run spheres100k_1080p_256spp_d4 $((16*sh)) $pipe --emulate-shard $sh; echo; done
done
echo -n "pool st4 variant (spheres, grid): "; PTB200_LIB=$PWD/build_ab/libptb200_st4.so run spheres100k_1080p_256spp_d4 32 pool; PTB200_LIB=$PWD/build_ab/libptb200_st4.so run grid1m_4k_64spp_d4 8 pool; echo
echo -n "pool float nodes shard 1/8 grid: "; PTB_QUANT_NODES=0 run grid1m_4k_8192spp_d4 64 pool --emulate-shard 8; echo
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
bash scripts/r02_sanitizer.sh memcheck
