#!/bin/bash
# the reference's own main.cpp (oracle/_ref/ref_host_b200: its Scene class over the cl:: shim + libptb200.so, no Python in the render)
# on N GPUs through the one added call ptb_set_gpus: BASELINE configs[3]'s scene (1 M triangles, 3840x2160, depth 4)
cd "$(dirname "$0")/.." || exit 1
N=${1:-4}
O=gpurun_out/refhost$N; mkdir -p $O
CAM=$(python -c "
from opencl_path_tracer_b200 import scenes
d = scenes.describe('grid1m'); scenes.write_ptbs(d, '$O/grid1m.ptbs')
print('--eye', *[repr(float(v)) for v in d['eye']], '--fov', repr(float(d['fov'])), '--cam-yaw', repr(float(d['yaw'])), '--cam-pitch', repr(float(d['pitch'])))")
H=oracle/_ref/ref_host_b200
COMMON="--w 3840 --h 2160 --iterations 4 --tribin $O/grid1m.ptbs --no-bvh $CAM"
{
echo "# ref_host_b200 (reference main.cpp over the shim), grid1m 3840x2160 depth 4, $N GPUs in ONE process (ptb_set_gpus); scripts/r02_refhost_multi.sh"
echo "## bit-exact flavour, 4 samples through Scene::render (main.cpp:683-687: gen_ray + trace_ray per sample)"
for g in 1 $N; do
  timeout 300 $H $COMMON --spp 4 --math 0 --gpus $g --split tile --out $O/par_g$g.bin --rnds-out $O/par_r$g.bin | grep '^{' | tail -1 | cut -c1-400
done
cmp $O/par_g1.bin $O/par_g$N.bin && cmp $O/par_r1.bin $O/par_r$N.bin && echo "tile split on $N GPUs == 1 GPU: colours and RNG states bit-identical"
echo "## shipping flavour, 64 samples in one ptb_render (in-kernel sample loop)"
for cfg in "1 tile" "$N tile" "$N spp"; do
  set -- $cfg
  echo -n "gpus=$1 split=$2: "; timeout 300 $H $COMMON --spp 64 --math 1 --render-ext --gpus $1 --split $2 --out $O/fast_$1_$2.bin | grep '^{' | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.readline()); print('%.3f s, %.0f Mrays/s' % (d['seconds'], d['rays'] / d['seconds'] / 1e6), d.get('error', ''))"
done
python - <<PY
import numpy as np
a = np.fromfile('$O/fast_1_tile.bin', dtype=np.float32).reshape(-1, 4)[:, :3]; b = np.fromfile('$O/fast_${N}_tile.bin', dtype=np.float32).reshape(-1, 4)[:, :3]
c = np.fromfile('$O/fast_${N}_spp.bin', dtype=np.float32).reshape(-1, 4)[:, :3]
rel = np.abs(a - b).max(axis=1) / np.maximum(np.abs(a).max(axis=1), 1e-6)
print('fast flavour, tile split vs 1 GPU: %.4f %% of pixels bit-equal, max rel diff %.2e (pool kernel on the shares, wavefront on the whole frame)' % (100 * (a.view(np.uint32) == b.view(np.uint32)).all(axis=1).mean(), rel.max()))
print('fast flavour, spp split vs 1 GPU (other random numbers): mean ratio %.4f, RMSE / mean %.3f' % (c.mean() / a.mean(), np.sqrt(((c - a) ** 2).mean()) / a.mean()))
PY
} 2>&1 | tee $O/report.txt
rm -f $O/*.bin $O/grid1m.ptbs
