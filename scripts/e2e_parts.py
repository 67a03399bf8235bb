"""Where does an end-to-end step go?  Times the pieces of bench.py's e2e_step on one workload (host clock around a finish())."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from opencl_path_tracer_b200 import capi, host, multigpu

name = sys.argv[1] if len(sys.argv) > 1 else "grid1m_4k_64spp_d4"
scene_name, W, H, spp, depth = bench.WORKLOADS[name]
spp = int(os.environ.get('SPP', '4'))
ptb = capi.Ptb(0)
dev = torch.device("cuda", 0)
if os.environ.get("USE_STREAM", "1") == "1":       # what bench.py does: a real (non-default) stream shared by the kernels and the timing events
    stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream); ptb.set_stream(stream.cuda_stream)
ptb.set_math(capi.MATH_FAST)
scene = host.Scene(ptb)
scene.build_trees = False
bench.build_scene(scene_name, scene)
scene.init_Scene(W, H, seed_skip=0)
frame = multigpu.DeviceFrame(ptb, W, H, dev)
seeds_host = torch.from_numpy(host.minstd_rand0_seeds(W * H, 0)).pin_memory()
frame.rnds.copy_(seeds_host)
frame.attach(scene)
scene.upload_Triangles(); scene.upload_Materials()
scene.iterations = depth
scene.camera = scene._camera(); scene._set_trace_args()
ptb.build_accel(); ptb.finish()
tris_b, mats_b = np.ascontiguousarray(scene.tris), np.stack(scene.mats)
tris_pinned = torch.from_numpy(tris_b.view(np.uint8).reshape(-1)).pin_memory()
colors_host = torch.empty(W * H, 4, dtype=torch.float32).pin_memory()

def timed(label, fn):
    ptb.finish(); torch.cuda.synchronize()
    t0 = time.perf_counter(); fn(); ptb.finish(); torch.cuda.synchronize()
    print(f"{label:40s} {1e3 * (time.perf_counter() - t0):9.2f} ms")

for rep in range(int(os.environ.get("PASSES", "2"))):
    print("--- pass", rep)
    timed("write seeds (pinned)", lambda: ptb.write_ptr(frame.h_rnds, seeds_host.data_ptr(), seeds_host.numel() * 4))
    timed("write tris (pageable numpy)", lambda: ptb.write(scene.buffer_tris, tris_b))
    timed("build_accel after tris write", lambda: ptb.build_accel())
    timed("write tris (pinned)", lambda: ptb.write_ptr(scene.buffer_tris, tris_pinned.data_ptr(), tris_pinned.numel()))
    timed("build_accel after tris write", lambda: ptb.build_accel())
    timed("write mats", lambda: ptb.write(scene.buffer_mats, mats_b))
    def r():
        scene.current_sample = 0; scene.render_spp(spp)
    timed(f"render {spp} spp", r)
    timed(f"render {spp} spp again", r)
    timed("read colors (pinned)", lambda: ptb.read_ptr(frame.h_colors, colors_host.data_ptr(), colors_host.numel() * 4))
