# experiment driver: A/B of two builds of the library (PTB200_LIB) on the LBVH scenes; $1 = the alternative .so
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), 'B=%.1f' % r['box_tests_per_ray'], end='  ')"; }
three() { run spheres100k_1080p_256spp_d4 32; run grid1m_4k_64spp_d4 8; run grid1m_4k_64spp_d4 32 --emulate-shard 8; echo; }
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_edges.py tests/test_gpu_rays.py -x -q -m gpu -k "wavefront or lbvh or pipeline or edge or rays or closest" 2>&1 | tail -2
echo -n "default build: "; three
echo -n "alt build $1: "; PTB200_LIB=$PWD/$1 three
for l in 1 2 8; do echo -n "default build, leaf_min=$l: "; PTB_WF_LEAF_MIN=$l three; done
