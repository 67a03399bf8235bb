# experiment driver: LBVH scenes (104k spheres at 1080p, 1M grid at 4K), Mrays/s and box/triangle tests per ray
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline ${3:-wavefront} --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), 'B=%.1f T=%.1f' % (r['box_tests_per_ray'], r['tri_tests_per_ray']))"; }
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu -k "wavefront or lbvh or pipeline" 2>&1 | tail -2
for sort in 0 1; do
  echo -n "spheres100k sort=$sort : "; PTB_WF_SORT=$sort run spheres100k_1080p_256spp_d4 16
  echo -n "grid1m sort=$sort : "; PTB_WF_SORT=$sort run grid1m_4k_64spp_d4 8
done
