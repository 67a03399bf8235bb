# experiment driver: persisting L2 window on the wide-node array (PTB_L2_PERSIST=2)
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(round(d['value']), end='  ')"; }
three() { run spheres100k_1080p_256spp_d4 32; run grid1m_4k_64spp_d4 8; run grid1m_4k_64spp_d4 32 --emulate-shard 8; echo; }
echo -n "default: "; three
echo -n "L2 persist: "; PTB_L2_PERSIST=2 PTB_VERBOSE=1 three 2>gpurun_out/l2.err; grep "L2 persisting" gpurun_out/l2.err | head -1
echo -n "default: "; three
echo -n "L2 persist: "; PTB_L2_PERSIST=2 three
