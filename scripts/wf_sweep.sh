# experiment driver: LBVH scenes (104k spheres at 1080p, 1M grid at 4K), Mrays/s and box/triangle tests per ray
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline ${3:-wavefront} --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), 'B=%.1f T=%.1f' % (r['box_tests_per_ray'], r['tri_tests_per_ray']))"; }
for steps in 4 2; do for leaf in 8 4; do
  echo -n "spheres100k node_steps=$steps leaf_min=$leaf : "; PTB_WF_NODE_STEPS=$steps PTB_WF_LEAF_MIN=$leaf run spheres100k_1080p_256spp_d4 8
  echo -n "grid1m node_steps=$steps leaf_min=$leaf : "; PTB_WF_NODE_STEPS=$steps PTB_WF_LEAF_MIN=$leaf run grid1m_4k_64spp_d4 8
done; done
