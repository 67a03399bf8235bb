# experiment driver: number of interleaved sub-wavefronts (PTB_WF_SUBS) on the LBVH scenes, full frames and a 1/8 frame (= one rank of the 8-GPU tile split)
run() { python bench.py --workload $1 --spp $2 --steps 2 --warmup 3 --pipeline wavefront --no-cpu-baseline --no-e2e ${@:3} 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(round(d['value']), '%.1f ms' % d['ms_per_step'])"; }
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_edges.py -x -q -m gpu -k "wavefront or lbvh or pipeline or edge" 2>&1 | tail -2
for subs in 1 2 3 4; do
  echo -n "spheres100k 1080p subs=$subs : "; PTB_WF_SUBS=$subs run spheres100k_1080p_256spp_d4 32
  echo -n "grid1m 4K subs=$subs : "; PTB_WF_SUBS=$subs run grid1m_4k_64spp_d4 8
  echo -n "grid1m 1/8 of 4K (3840x270) subs=$subs : "; PTB_WF_SUBS=$subs run grid1m_4k_64spp_d4 32 --height 270
done
