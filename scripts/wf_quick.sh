# experiment driver: LBVH parity tests, then Mrays/s of the wavefront pipeline on the 104k / 1M-triangle scenes (full frames and one rank of an 8-way tile split)
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), '%.1f ms' % d['ms_per_step'], 'B=%.1f T=%.1f' % (r['box_tests_per_ray'], r['tri_tests_per_ray']))"; }
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_edges.py tests/test_gpu_rays.py -x -q -m gpu -k "wavefront or lbvh or pipeline or edge or rays or closest" 2>&1 | tail -2
echo -n "spheres100k 1080p : "; run spheres100k_1080p_256spp_d4 32
echo -n "grid1m 4K : "; run grid1m_4k_64spp_d4 8
echo -n "grid1m 4K, rank 0 of 8 : "; run grid1m_4k_64spp_d4 32 --emulate-shard 8
