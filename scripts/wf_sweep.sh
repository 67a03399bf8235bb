# experiment driver: wavefront pipeline variants on the 104k-triangle scene (and the 1M grid), Mrays/s
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(round(d['value']))"; }
for tex in 0 1; do for wide in 0 1; do
  echo -n "spheres100k tex=$tex wide=$wide : "; PTB_TEX_NODES=$tex PTB_WIDE_BVH=$wide run spheres100k_1080p_256spp_d4 8
  echo -n "grid1m tex=$tex wide=$wide : "; PTB_TEX_NODES=$tex PTB_WIDE_BVH=$wide run grid1m_4k_64spp_d4 8
done; done
