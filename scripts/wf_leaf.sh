# experiment driver: triangles per LBVH leaf (PTB_LBVH_LEAF) on the two LBVH scenes
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 2 --pipeline wavefront --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), 'B=%.1f T=%.1f' % (r['box_tests_per_ray'], r['tri_tests_per_ray']), end='  ')"; }
for l in 2 3 4 6 8; do echo -n "leaf=$l: "; PTB_LBVH_LEAF=$l run spheres100k_1080p_256spp_d4 24; PTB_LBVH_LEAF=$l run grid1m_4k_64spp_d4 6; echo; done
