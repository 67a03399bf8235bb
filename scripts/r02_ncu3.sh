#!/bin/bash
# ncu --set full of k_wf_trace + k_wf_shade on the THIRD bounce of the first sample (incoherent rays; launches 12-13 of the two
# sub-wavefronts' trace / shade sequence), final round-2 code; each command first runs plain
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/ncu3; mkdir -p $O
A="--workloads none --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-parity-check"
cap() { # name, kernel regex, skip, count, command...
  local name=$1 k=$2 s=$3 c=$4; shift 4
  "$@" > $O/plain_$name.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c $c -f -o $O/$name "$@" > $O/ncu_$name.log 2>&1; echo "$name rc=$?"
}
cap wf_spheres_b3 'k_wf_(trace|shade)' 12 2 python bench.py --workload spheres100k_1080p_256spp_d4 --spp 2 --pipeline wavefront $A
cap wf_grid_b3 'k_wf_(trace|shade)' 12 2 python bench.py --workload grid1m_4k_64spp_d4 --spp 2 --pipeline wavefront $A
ls -la $O | head
