# experiment driver: the two-ray-pool trace kernel (PTB_WF_POOL=1) against the default, parity tests first
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), end='  ')"; }
three() { run spheres100k_1080p_256spp_d4 32; run grid1m_4k_64spp_d4 8; echo; }
PTB_WF_POOL=1 PTB_WF_FETCH_MIN=16 PTB_WF_LEAF_MIN=16 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_edges.py tests/test_gpu_rays.py -x -q -m gpu -k "wavefront or lbvh or pipeline or edge or rays or closest" 2>&1 | tail -2
echo -n "default: "; three
for f in 12 20 28; do for l in 8 16 24; do echo -n "pool fetch_min=$f leaf_min=$l: "; PTB_WF_POOL=1 PTB_WF_FETCH_MIN=$f PTB_WF_LEAF_MIN=$l three; done; done
