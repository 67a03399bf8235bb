# experiment driver: vote thresholds of k_wf_trace and the 9-blocks/SM build (build_ab/libptb200_b9.so, -DWF_MINB=9)
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(round(d['value']), end=' ')"; }
both() { run spheres100k_1080p_256spp_d4 32; run grid1m_4k_64spp_d4 8; echo; }
echo -n "default: "; both
echo -n "b9: "; PTB200_LIB=$PWD/build_ab/libptb200_b9.so both
for f in 2 4 12 16; do echo -n "fetch_min=$f: "; PTB_WF_FETCH_MIN=$f both; done
for l in 1 2 6 8; do echo -n "leaf_min=$l: "; PTB_WF_LEAF_MIN=$l both; done
for s in 1 2 8; do echo -n "node_steps=$s: "; PTB_WF_NODE_STEPS=$s both; done
