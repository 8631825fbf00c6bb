"""Camera.update replay: pt_camera_state_new / pt_camera_update / pt_camera_pack follow
src/Camera.zig:25-102 of the reference.  Known answers are exact; a random input trace is compared with an
independent double-precision restatement written here."""
import math
import os
import subprocess

import numpy as np

import b200pt

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "vulkan-pathtracer_b200", "b200pt")
FOV = math.pi * 0.25


def vec(a):
    return np.array(list(a), np.float64)


def bits(x):
    return np.asarray(list(x), np.float32).view(np.uint32).tolist()


def test_state_new_is_camera_new():
    c = b200pt.Camera(FOV, 16.0 / 9.0, (1.0, 2.0, 3.0))
    a = c.pack(9)
    b = b200pt.camera_new(FOV, 16.0 / 9.0, (1.0, 2.0, 3.0), frame_count=9)
    for f in ("position", "horizontal", "vertical", "forward"):
        assert bits(getattr(a, f)) == bits(getattr(b, f)), f
    assert a.frame_count == 9 and c.state.yaw == 0.0 and c.state.pitch == 0.0                   # Camera.zig:32-34
    assert list(c.state.forward) == [0.0, 0.0, 1.0] and list(c.state.left) == [-1.0, 0.0, 0.0]  # Vec3.forward / left
    assert list(c.state.up) == [0.0, 1.0, 0.0]


def test_keys_move_at_five_or_ten_units_per_second():
    c = b200pt.Camera(FOV, 1.0)
    c.update(keys=("w",), delta_time=1.0)
    assert list(c.state.position) == [0.0, 0.0, 5.0]                                            # move_speed 5, Camera.zig:67
    c.update(keys=("w", "left_shift"), delta_time=0.5)
    assert list(c.state.position) == [0.0, 0.0, 10.0]                                           # shift: 10
    c.update(keys=("s",), delta_time=2.0)
    assert list(c.state.position) == [0.0, 0.0, 0.0]
    c.update(keys=("d",), delta_time=1.0)                                                       # d ADDS left = -x (:71)
    assert list(c.state.position) == [-5.0, 0.0, 0.0]
    c.update(keys=("a",), delta_time=1.0)
    assert list(c.state.position) == [0.0, 0.0, 0.0]
    c.update(keys=("space",), delta_time=0.25)
    assert list(c.state.position) == [0.0, 1.25, 0.0]
    c.update(keys=("left_ctrl", "left_shift"), delta_time=0.125)
    assert list(c.state.position) == [0.0, 0.0, 0.0]


def test_opposite_keys_cancel_without_nan():
    c = b200pt.Camera(FOV, 1.0, (1.0, 1.0, 1.0))
    c.update(keys=("w", "s", "a", "d"), delta_time=1.0)                                         # norm() of the zero vector
    assert list(c.state.position) == [1.0, 1.0, 1.0]
    c.update(keys=("w", "d"), delta_time=1.0)                                                   # diagonal is normalised
    p = vec(c.state.position) - 1.0
    assert abs(np.linalg.norm(p) - 5.0) < 1e-5 and abs(p[0] + p[2]) < 1e-6


def test_mouse_look_needs_the_left_button():
    c = b200pt.Camera(FOV, 1.0)
    c.update(mouse_left=False, cursor_delta=(40, 40), delta_time=0.1)
    assert c.state.yaw == 0.0 and c.state.pitch == 0.0                                          # Camera.zig:49
    c.update(mouse_left=True, cursor_delta=(4, 8), delta_time=0.1)
    assert c.state.yaw == 359.0 and c.state.pitch == 2.0                                        # -= dx/4 wrapped, += dy/4
    c.update(mouse_left=True, cursor_delta=(-8, 0), delta_time=0.1)
    assert c.state.yaw == 1.0                                                                   # >= 360 wraps down
    c.update(mouse_left=True, cursor_delta=(0, 4000), delta_time=0.1)
    assert c.state.pitch == 90.0                                                                # clamp, :60
    c.update(mouse_left=True, cursor_delta=(0, -4000), delta_time=0.1)
    assert c.state.pitch == -90.0


def test_aspect_ratio_update_rescales_horizontal_only():
    c = b200pt.Camera(FOV, 1.0)
    v0, h0 = list(c.state.vertical), vec(c.state.horizontal)
    c.update_aspect_ratio(2.0)                                                                  # Camera.zig:81-84
    assert list(c.state.vertical) == v0
    np.testing.assert_allclose(vec(c.state.horizontal), 2.0 * h0, rtol=1e-7)


class RefCamera:
    """Camera.zig restated in double precision (zalgebra: fromAxis in degrees, rotateVec, norm)."""

    def __init__(self, fov, aspect, pos):
        self.fov, self.aspect, self.pos, self.yaw, self.pitch = fov, aspect, vec(pos), 0.0, 0.0
        self.orient()

    @staticmethod
    def from_axis(deg, axis):
        r = math.radians(deg) * 0.5
        return np.array([math.cos(r), *(vec(axis) * math.sin(r))])

    @staticmethod
    def qmul(a, b):
        w1, x1, y1, z1 = a
        w2, x2, y2, z2 = b
        return np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                         w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])

    @staticmethod
    def rotate(q, v):
        q = q / np.linalg.norm(q)
        b = q[1:]
        return v * (q[0] * q[0] - b @ b) + b * (2.0 * (v @ b)) + np.cross(b, v) * (2.0 * q[0])

    def orient(self):
        q = self.qmul(self.from_axis(self.yaw, (0, 1, 0)), self.from_axis(self.pitch, (1, 0, 0)))
        self.forward, self.left = self.rotate(q, vec((0, 0, 1))), self.rotate(q, vec((-1, 0, 0)))
        self.up = np.cross(self.left, self.forward)
        vh = 2.0 * math.tan(self.fov * 0.5)
        self.horizontal, self.vertical = self.left * (vh * self.aspect), self.up * vh

    def update(self, keys, mouse, dx, dy, dt):
        f32 = np.float32
        if mouse and (dx or dy):
            self.yaw = float(f32(self.yaw) - f32(dx) * f32(0.25))
            self.pitch = float(f32(self.pitch) + f32(dy) * f32(0.25))
            if self.yaw < 0.0:
                self.yaw += 360.0
            if self.yaw >= 360.0:
                self.yaw -= 360.0
            self.pitch = min(max(self.pitch, -90.0), 90.0)
            self.orient()
        speed = 10.0 if "left_shift" in keys else 5.0
        d = np.zeros(3)
        if "w" in keys: d += self.forward
        if "s" in keys: d -= self.forward
        if "d" in keys: d += self.left
        if "a" in keys: d -= self.left
        n = np.linalg.norm(d)
        if n > 0:
            d /= n
        self.pos = self.pos + d * (speed * dt)
        if "space" in keys: self.pos[1] += speed * dt
        if "left_ctrl" in keys: self.pos[1] -= speed * dt


def test_random_input_trace_follows_the_restated_camera():
    rng = np.random.default_rng(5)
    names = list(b200pt.KEY_NAMES)
    c = b200pt.Camera(FOV, 16.0 / 9.0, (0.5, 1.0, -2.0))
    ref = RefCamera(FOV, 16.0 / 9.0, (0.5, 1.0, -2.0))
    for step in range(400):
        keys = [k for k in names if rng.random() < 0.3]
        mouse = bool(rng.random() < 0.6)
        dx, dy = (int(v) for v in rng.integers(-30, 31, 2))
        dt = float(np.float32(rng.uniform(0.001, 0.03)))
        c.update(keys=keys, mouse_left=mouse, cursor_delta=(dx, dy), delta_time=dt)
        ref.update(keys, mouse, dx, dy, dt)
        assert abs(c.state.yaw - ref.yaw) < 1e-3 and abs(c.state.pitch - ref.pitch) < 1e-3
        np.testing.assert_allclose(vec(c.state.forward), ref.forward, atol=2e-6)
        np.testing.assert_allclose(vec(c.state.left), ref.left, atol=2e-6)
    np.testing.assert_allclose(vec(c.state.position), ref.pos, atol=2e-3)
    cam = c.pack(400)
    np.testing.assert_allclose(vec(cam.horizontal)[:3], ref.horizontal, atol=1e-5)
    np.testing.assert_allclose(vec(cam.vertical)[:3], ref.vertical, atol=1e-5)
    assert cam.position[3] == 1.0 and cam.horizontal[3] == 0.0 and cam.frame_count == 400


def test_cli_rejects_a_malformed_camera_path(tmp_path):
    p = tmp_path / "path.txt"
    p.write_text("# dt keys mouse dx dy\n0.016 w+jump 0 0 0\n")
    r = subprocess.run([EXE, "-s", "none.glb", "--camera-path", str(p)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "unknown key 'jump'" in r.stderr
    p.write_text("0.016 w 0 0\n")
    r = subprocess.run([EXE, "-s", "none.glb", "--camera-path", str(p)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "expected" in r.stderr
