"""Cornell 1080p, parity flavour, a few samples: the command the ncu source-level capture of the bit-exact k_render runs"""
import sys, time
from opencl_path_tracer_b200 import capi, host
from tests import util
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ptb = capi.Ptb(0)
ptb.set_accel(capi.ACCEL_BRUTE); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
s = host.Scene(ptb)
s.init_Scene(1920, 1080)
s.add_Obj(util.CORNELL, (555, 0, 0), (1, 1, 1), 0, 0)
s.upload_Triangles(); s.upload_Materials(); s.iterations = 4
cam = host.camera_at(1920, 1080, 39.3, 0, 0, (277.5, 277.5, -800))
s._camera = lambda: cam
s.render_spp(1); ptb.finish()
t0 = time.time(); s.render_spp(spp); ptb.finish(); dt = time.time() - t0
print("parity %d spp: %.1f ms" % (spp, dt * 1e3))
