#!/bin/bash
# ncu --set full of ONE k_pool launch on the 104 k / 1 M-triangle scenes (the same command first runs plain)
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/ncu; mkdir -p $O
LIB=${1:-}
[ -n "$LIB" ] && export PTB200_LIB=$PWD/$LIB
A="--workloads none --steps 1 --warmup 0 --pipeline pool --no-cpu-baseline --no-e2e --no-parity-check"
C1="python bench.py --workload spheres100k_1080p_256spp_d4 --spp 8 $A"
C2="python bench.py --workload grid1m_4k_64spp_d4 --spp 2 $A"
$C1 > $O/plain_spheres.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pool -c 1 -f -o $O/pool_spheres${2:-} $C1 > $O/ncu_spheres.log 2>&1; echo "spheres rc=$?"
$C2 > $O/plain_grid.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pool -c 1 -f -o $O/pool_grid${2:-} $C2 > $O/ncu_grid.log 2>&1; echo "grid rc=$?"
ls -la $O
