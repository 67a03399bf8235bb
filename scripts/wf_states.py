"""throw-away: lane-state census of k_wf_trace (needs the -DWF_DEBUG_STATES build copied over libptb200.so)"""
import sys
import numpy as np
from opencl_path_tracer_b200 import capi, host, scenes
ptb = capi.Ptb(0)
ptb.set_accel(capi.ACCEL_LBVH); ptb.set_math(capi.MATH_FAST); ptb.set_pipeline(capi.PIPE_WAVEFRONT)
d = scenes.describe("spheres100k")
s = host.Scene(ptb); s.build_trees = False
s.init_Scene(1920, 1080)
scenes.populate(s, d)
s.upload_Triangles(); s.upload_Materials(); s.iterations = 4
cam = host.camera_at(1920, 1080, d["fov"], d.get("yaw", 0), d.get("pitch", 0), d["eye"])
s._camera = lambda: cam
ptb.set_count_work(1); ptb.reset_counters()
s.render_spp(2); ptb.finish()
c = ptb.counters()
it = c["pixel_samples"] - 1920 * 1080 * 2
print(c)
print("votes", it, "avg lanes NODE+POP %.1f (POP %.1f) LEAF %.1f FETCH %.1f EXIT/idle %.1f" % (c["tri_stage1"] / it, c["box_tests"] / it, c["tri_stage2"] / it, c["tri_stage3"] / it, 32 - (c["tri_stage1"] + c["tri_stage2"] + c["tri_stage3"]) / it))
