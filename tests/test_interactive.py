"""Interactive / progressive semantics (opencl_path_tracer_b200/interactive.py) against the rules of
main.cpp:1040-1241 — CPU part: the callback logic on a recording stand-in for the device; GPU part: a scripted camera
path rendered through the C-ABI, checked frame by frame against the oracle run with the same cameras, sample indices
and carried-over RNG state."""
import os

import numpy as np
import pytest

from opencl_path_tracer_b200 import host, interactive
from tests import util

f32 = np.float32


class _Recorder:
    """Just enough of capi.Ptb for host.Scene.render()/render_spp(): records launches, owns no device."""

    def __init__(self):
        self.launches, self.args = [], {}

    def set_arg_mem(self, k, i, h): self.args[(k, i)] = h
    def set_arg(self, k, i, v): self.args[(k, i)] = v
    def launch(self, k, gx, gy=1): self.launches.append((k, dict(self.args)))
    def render(self, spp): self.launches.append(("render", spp, dict(self.args)))
    def finish(self): pass


def _scene():
    s = host.Scene(ptb=_Recorder())
    s.screen_width, s.screen_height, s.rays_size = 8, 4, 32
    s.buffer_rays = s.buffer_rnds = s.cl_screen = s.buffer_colors = 1
    s.buffer_tris = s.buffer_mats = s.buffer_kd_tree = s.buffer_kd_tree_shift = 1
    s.tris_size = s.kd_tree_shift_size = 0
    return s


def _sample_indices(rec):
    from opencl_path_tracer_b200 import capi
    return [a[(capi.KERNEL_TRACE_RAY, 10)] for k, a in [(l[0], l[-1]) for l in rec.launches] if k == capi.KERNEL_TRACE_RAY]


def test_progressive_accumulation_and_reset_rules():
    s = _scene()
    ses = interactive.Session(s)
    for _ in range(3):
        ses.on_idle(0.01)
    assert s.current_sample == 3 and _sample_indices(s.ptb) == [0, 1, 2]
    ses.on_keyboard("w")                                  # key down alone resets nothing (main.cpp:1064-1099) ...
    assert s.current_sample == 3
    ses.on_idle(0.01)                                     # ... the next idle does, and renders sample 0 (main.cpp:1179-1183)
    assert _sample_indices(s.ptb)[-1] == 0 and s.current_sample == 1
    ses.on_idle(0.01)
    assert _sample_indices(s.ptb)[-1] == 0                # still held: every frame starts over
    ses.on_keyboard_up("W")                               # release resets (main.cpp:1101-1136), case-insensitive
    assert s.current_sample == 0 and not ses.moving()
    ses.on_idle(0.01); ses.on_idle(0.01)
    assert _sample_indices(s.ptb)[-2:] == [0, 1]
    ses.on_mouse(interactive.GLUT_LEFT_BUTTON, interactive.GLUT_DOWN, 10, 10)
    assert s.current_sample == 0 and ses.moving()
    ses.on_mouse(interactive.GLUT_LEFT_BUTTON, interactive.GLUT_UP, 10, 10)
    assert not ses.moving()


def test_depth_keys_clamp_and_reset():
    s = _scene()
    ses = interactive.Session(s)
    s.iterations, s.current_sample = 1, 5
    ses.on_keyboard("-")                                  # already 1: no change, NO reset (main.cpp:1042-1047)
    assert s.iterations == 1 and s.current_sample == 5
    ses.on_keyboard("+")
    assert s.iterations == 2 and s.current_sample == 0
    s.iterations, s.current_sample = interactive.MAX_ITERATIONS, 7
    ses.on_keyboard("+")
    assert s.iterations == 50 and s.current_sample == 7
    ses.on_keyboard("r")
    assert ses.real_time is False
    ses.on_keyboard("\x1b")
    assert ses.closed


def test_camera_motion_integrates_along_the_rotated_basis():
    s = _scene()
    s.global_fov, s.global_yaw, s.global_pitch = 60.0, 0.0, 0.0
    s.global_shift = np.zeros(3, dtype=np.float32)
    ses = interactive.Session(s)
    ses.on_keyboard("w")
    ses.on_idle(0.016)                                    # forward = 1000*dt along +z at yaw = pitch = 0
    assert np.allclose(s.global_shift, [0, 0, 16.0], atol=1e-4)
    ses.on_keyboard_up("w"); ses.on_keyboard("d"); ses.on_keyboard("q")
    ses.on_idle(0.01)
    assert np.allclose(s.global_shift, [10.0, 10.0, 16.0], atol=1e-4)
    ses.on_keyboard_up("d"); ses.on_keyboard_up("q")
    before = s.global_shift.copy()
    ses.on_idle(0.01)                                     # nothing held: the camera rests
    assert (s.global_shift == before).all()
    # yaw 90 degrees: "ahead" is +x (rotate_y, main.cpp:55-62)
    s.global_yaw = 90.0
    ses.on_keyboard("s"); ses.on_idle(0.01); ses.on_keyboard_up("s")
    assert np.allclose(s.global_shift - before, [-10.0, 0, 0], atol=1e-3)
    # the camera record handed to the kernels sees the moved eye (main.cpp:343)
    cam = np.frombuffer(s.camera.tobytes(), dtype=host.CAMERA)[0]
    assert np.allclose(cam["eye"][:3], np.array([500, 500, -1299.037842], dtype=np.float32) + s.global_shift, atol=1e-3)


def test_fov_keys_and_mouse_speed_tiers():
    s = _scene()
    ses = interactive.Session(s)
    s.global_fov = 75.0
    ses.on_keyboard("e"); ses.on_idle(0.1); ses.on_keyboard_up("e")
    assert abs(s.global_fov - 73.0) < 1e-4                # -20*dt above 10 degrees
    s.global_fov = 5.0
    ses.on_keyboard("e"); ses.on_idle(0.1); ses.on_keyboard_up("e")
    assert abs(s.global_fov - 4.8) < 1e-4                 # -2*dt below
    s.global_fov = 0.05
    ses.on_keyboard("e"); ses.on_idle(0.1); ses.on_keyboard_up("e")
    assert abs(s.global_fov - 0.1) < 1e-7                 # clamp
    s.global_fov = 95.0
    ses.on_keyboard("c"); ses.on_idle(0.1); ses.on_keyboard_up("c")
    assert s.global_fov == 90.0
    s.global_fov, s.global_yaw, s.global_pitch = 75.0, 0.0, 0.0
    ses.on_mouse(interactive.GLUT_LEFT_BUTTON, interactive.GLUT_DOWN, 100, 100)
    ses.on_mouse_motion(110, 95)
    assert abs(s.global_yaw - 2.0) < 1e-5 and abs(s.global_pitch + 1.0) < 1e-5      # 0.2 degrees per pixel
    s.global_fov = 5.0
    ses.on_mouse_motion(120, 95)
    assert abs(s.global_yaw - 2.5) < 1e-5                                           # 0.05 below 10 degrees
    s.global_fov = 1.0
    ses.on_mouse_motion(130, 95)
    assert abs(s.global_yaw - 2.6) < 1e-5                                           # 0.01 below 2 degrees


def test_idle_still_is_one_launch_and_refuses_while_moving():
    s = _scene()
    ses = interactive.Session(s)
    ses.idle_still(64)
    assert s.current_sample == 64 and s.ptb.launches[-1][:2] == ("render", 64)
    ses.on_keyboard("a")
    with pytest.raises(RuntimeError):
        ses.idle_still(4)


def test_scripted_run_orders_events_by_frame():
    s = _scene()
    ses = interactive.Session(s)
    ev = [(2, lambda q: q.on_keyboard("w")), (4, lambda q: q.on_keyboard_up("w")), (0, lambda q: q.on_keyboard("+"))]
    n = ses.run(7, 0.01, ev)
    assert s.iterations == 2
    assert _sample_indices(s.ptb) == [0, 1, 0, 0, 0, 1, 2] and n == 3


# ---------------------------------------------------------------- GPU: scripted path vs the oracle, frame by frame
@pytest.mark.gpu
def test_scripted_camera_path_matches_oracle(ptb):
    from opencl_path_tracer_b200 import capi
    from oracle import pt_oracle_py as po
    W, H = 64, 48
    ptb.set_accel(capi.ACCEL_AUTO); ptb.set_math(capi.MATH_PARITY); ptb.set_pipeline(capi.PIPE_MEGAKERNEL)
    s = host.Scene(ptb)
    s.init_Scene(W, H)
    s.add_Obj(util.CORNELL, (555, 0, 0), (1, 1, 1), 0, 0)
    s.upload_Triangles(); s.upload_Materials()
    s.iterations = 3
    s.global_fov, s.global_yaw, s.global_pitch = 39.3, 0.0, 0.0
    s.global_shift = np.array([277.5 - 500, 277.5 - 500, -800 + 1299.037842], dtype=np.float32)
    surface = {"W": W, "H": H, "tris": s.tris, "mats": np.stack(s.mats), "kd_tree": host.nodes_to_records(s.kd_tree),
               "kd_tree_shift": np.asarray(s.kd_tree_shift, dtype=np.int32), "iterations": s.iterations,
               "camera": host.Camera(W, H, s.global_fov, s.global_yaw, s.global_pitch, s.global_shift)}
    orc = po.OracleScene(surface, brute=True)
    ses = interactive.Session(s)
    events = {2: lambda q: q.on_keyboard("w"), 4: lambda q: (q.on_keyboard_up("w"), q.on_keyboard("+")),
              6: lambda q: (q.on_mouse(0, 0, 5, 5), q.on_mouse_motion(25, 0)), 7: lambda q: q.on_mouse(0, 1, 25, 0)}
    import ctypes as C
    for frame in range(10):
        if frame in events:
            events[frame](ses)
        ses.on_idle(0.02)
        # the oracle renders the same sample with the camera and sample index the session just used
        C.memmove(C.byref(orc.c.cam), s.camera.tobytes(), 80)
        orc.c.iterations = s.iterations
        orc.sample = s.current_sample - 1
        orc.render(1, threads=4)
    ptb.finish()
    got = s.download_colors()
    st = util.compare(got, orc.colors.reshape(H, W, 4))
    assert st["bit_equal_frac"] >= 1.0 - 1e-4, st
    assert (s.download_rnds() == orc.rnds).all()          # the RNG state carried over every reset
    assert s.current_sample == 3 and s.iterations == 4    # frames 7, 8, 9 accumulated after the last reset
    img = ses.refresh()
    assert img.shape == (H, W, 4) and img[..., 3].min() == 255
