#!/bin/bash
# ncu --set full: one launch each of k_render (Cornell), k_wf_trace + k_wf_shade (spheres100k, grid1m), k_pool (spheres100k); each command first runs plain
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/ncu2; mkdir -p $O
A="--workloads none --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-parity-check"
cap() { # name, kernel regex, skip, count, command...
  local name=$1 k=$2 s=$3 c=$4; shift 4
  "$@" > $O/plain_$name.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c $c -f -o $O/$name "$@" > $O/ncu_$name.log 2>&1; echo "$name rc=$?"
}
cap render k_render 0 1 python bench.py --workload cornell_1080p_1024spp_d4 --spp 16 $A
cap wf_spheres 'k_wf_(trace|shade)' 2 2 python bench.py --workload spheres100k_1080p_256spp_d4 --spp 2 --pipeline wavefront $A
cap wf_grid 'k_wf_(trace|shade)' 2 2 python bench.py --workload grid1m_4k_64spp_d4 --spp 1 --pipeline wavefront $A
cap pool_spheres k_pool 0 1 python bench.py --workload spheres100k_1080p_256spp_d4 --spp 8 --pipeline pool $A
python bench.py --workload spheres100k_1080p_256spp_d4 --spp 16 --pipeline wavefront $A > $O/plain_launches.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_c3_wavefront.csv python bench.py --workload spheres100k_1080p_256spp_d4 --spp 16 --pipeline wavefront $A > $O/ncu_launches.log 2>&1; echo "launch list rc=$?"
ls -la $O | head -30
