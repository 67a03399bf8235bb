"""Headless mirror of the reference's interactive / progressive semantics (main.cpp:1040-1241).

The reference drives the path from GLUT callbacks: keys and the mouse move the camera, every move resets the
progressive accumulation (``current_sample = 0``) while the per-pixel RNG state simply carries on, ``+``/``-`` change
the path depth, and ``onIdle`` renders exactly one sample per call.  There is no window on this box, so the callbacks
become methods of :class:`Session` and wall-clock time becomes an explicit ``dt`` argument, which makes a camera path
a deterministic script.  The device work goes through the same ``Scene`` mirror (``host.py``) and therefore through
the same C-ABI calls the reference host makes.

Not reproduced: the window itself, full-screen toggle, the once-per-second console line, ``exit`` on ESC (it sets
``closed``)."""
import numpy as np

from . import host

f32 = np.float32
MOVE_KEYS = "wasdqyec"          # enum ControllKeys {W, A, S, D, Q, Y, E, C}, main.cpp:41
MAX_ITERATIONS = 50             # main.cpp:26
GLUT_LEFT_BUTTON, GLUT_DOWN, GLUT_UP = 0, 0, 1


class Session:
    def __init__(self, scene, real_time=True):
        self.scene = scene
        self.real_time = real_time                     # main.cpp:43
        self.keys_down = {k: False for k in MOVE_KEYS}
        self.mouse_down = False
        self.last_x = self.last_y = 0
        self.closed = False
        self.frames = 0
        self.resets = 0

    # ---- helpers
    def _reset(self):
        if self.scene.current_sample != 0:
            self.resets += 1
        self.scene.current_sample = 0

    def moving(self):
        return self.mouse_down or any(self.keys_down.values())

    # ---- onKeyboard, main.cpp:1041-1099
    def on_keyboard(self, key):
        s = self.scene
        if key == "-" and s.iterations > 1:
            s.iterations -= 1
            self._reset()
        if key == "+" and s.iterations < MAX_ITERATIONS:
            s.iterations += 1
            self._reset()
        if key == "\x1b":
            self.closed = True
        if key == "r":
            self.real_time = not self.real_time
        k = key.lower()
        if k in self.keys_down:
            self.keys_down[k] = True

    # ---- onKeyboardUp, main.cpp:1101-1136: releasing a movement key resets the accumulation
    def on_keyboard_up(self, key):
        k = key.lower()
        if k in self.keys_down:
            self.keys_down[k] = False
            self._reset()

    # ---- onMouse / onMouseMotion, main.cpp:1138-1162
    def on_mouse(self, button, state, x, y):
        self.last_x, self.last_y = x, y
        if button == GLUT_LEFT_BUTTON and state in (GLUT_DOWN, GLUT_UP):
            self.mouse_down = state == GLUT_DOWN
            self._reset()

    def on_mouse_motion(self, x, y):
        s = self.scene
        dx, dy = x - self.last_x, y - self.last_y
        speed = f32(0.2)
        if s.global_fov < 10:
            speed = f32(0.05)
        if s.global_fov < 2:
            speed = f32(0.01)
        s.global_yaw = float(f32(f32(s.global_yaw) + f32(dx) * speed))
        s.global_pitch = float(f32(f32(s.global_pitch) + f32(dy) * speed))
        self.last_x, self.last_y = x, y

    # ---- onIdle, main.cpp:1170-1241: one sample per call
    def on_idle(self, dt):
        s = self.scene
        kd = self.keys_down
        if self.moving():
            self._reset()
        dt = f32(dt)
        speed = f32(1000.0)
        s.global_forward = float(speed * dt) if kd["w"] else float(-speed * dt) if kd["s"] else 0.0
        s.global_rightward = float(-speed * dt) if kd["a"] else float(speed * dt) if kd["d"] else 0.0
        s.global_upward = float(speed * dt) if kd["q"] else float(-speed * dt) if kd["y"] else 0.0
        fov = f32(s.global_fov)
        if kd["e"]:
            if fov > 10:
                fov = f32(fov - f32(20) * dt)
            elif fov > f32(0.1):
                fov = f32(fov - f32(2) * dt)
            else:
                fov = f32(0.1)
        elif kd["c"]:
            if fov < 10:
                fov = f32(fov + f32(2) * dt)
            elif fov < 90:
                fov = f32(fov + f32(20) * dt)
            else:
                fov = f32(90.0)
        s.global_fov = float(fov)
        s.render()                                     # generate_rays (moves the camera) + trace_rays, sample++
        self.frames += 1

    # ---- extension: `n` idle calls with nothing held down are `n` samples of a resting camera — one launch
    def idle_still(self, n):
        if self.moving():
            raise RuntimeError("idle_still() while a key or the mouse button is held")
        s = self.scene
        s.global_forward = s.global_rightward = s.global_upward = 0.0
        s.render_spp(n)
        self.frames += n

    # ---- scripted camera path: events = [(frame, callable(session))...]; frames rendered with on_idle(dt)
    def run(self, n_frames, dt, events=()):
        todo = sorted(events, key=lambda e: e[0])
        i = 0
        for frame in range(n_frames):
            while i < len(todo) and todo[i][0] <= frame:
                todo[i][1](self)
                i += 1
            if self.closed:
                break
            self.on_idle(dt)
        return self.scene.current_sample

    # ---- progressive refresh: what the window would show now (tone-mapped RGBA8, top row first)
    def refresh(self, mode=0):
        s = self.scene
        return s.ptb.resolve_rgba8(s.buffer_colors, s.screen_width, s.screen_height, mode)
