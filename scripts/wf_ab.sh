# experiment driver: A/B of builds of the library (PTB200_LIB) on the LBVH scenes; arguments = alternative .so files
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); r=d['roofline']; print(round(d['value']), 'B=%.1f' % r['box_tests_per_ray'], end='  ')"; }
three() { run spheres100k_1080p_256spp_d4 32; run grid1m_4k_64spp_d4 8; run grid1m_4k_64spp_d4 32 --emulate-shard 8; echo; }
echo -n "default build: "; three
for lib in "$@"; do echo -n "alt build $lib: "; PTB200_LIB=$PWD/$lib three; done
