"""CPU tier: host Camera / uniform block vs the REFERENCE's own Camera.cpp (compiled in place by oracle/build_ref.sh;
its output is committed as tests/golden/camera_golden.json)."""
import json
import os

import numpy as np

from vkrt_b200 import camera

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "camera_golden.json")


def bits(a):
    return np.asarray(a, dtype=np.float32).view(np.uint32).tolist()


def _check(cam, ref):
    assert bits(cam.getPosition()) == ref["position"]
    assert bits(cam.getForward()) == ref["forward"]
    assert bits(cam.getLeft()) == ref["left"]
    assert bits(cam.getUp()) == ref["up"]
    assert bits(cam.getViewMatrix()) == ref["view"]
    assert bits(cam.getProjectionMatrix()) == ref["proj"]
    assert bits(cam.wvp(0.01)) == ref["wvp"]  # proj * view * scale(0.01) as glm multiplies it (src/Rasterizer.cpp:189-190)
    u = camera.uniforms_to_array(cam.uniforms())
    assert bits(u[0:16]) == ref["viewInverse"]
    assert bits(u[16:32]) == ref["projInverse"]
    assert np.array_equal(u[32:35], cam.getPosition()) and u[35] == 1.0
    assert np.array_equal(u[36:39], -cam.getLeft())
    assert np.array_equal(u[48:64].reshape(4, 4), camera.LIGHT_POSITIONS)


def test_camera_bit_exact_against_reference():
    g = {e["label"]: e for e in json.load(open(GOLDEN))}
    _check(camera.Camera(), g["default"])
    cam = camera.Camera()
    cam.setPosition((6.3, 4.5, -0.7))
    cam.setRotation((0.0, 1.57, 0.0))
    _check(cam, g["raytracer_default"])
    dt = np.float32(1.0) / np.float32(60.0)
    ta, ra = np.float32(5.0) * dt, np.float32(1.5) * dt
    for _ in range(150):
        cam.translate(cam.getForward() * ta)
    _check(cam, g["after_150_W"])
    for _ in range(150):
        cam.rotate((0, 1, 0), ra)
    _check(cam, g["after_150_W_150_Z"])
    cam = camera.Camera()
    cam.setPosition((-4.0, 2.0, -0.2))
    cam.setRotation((0.0, 1.51, 0.0))
    _check(cam, g["rasterizer_default"])
    cam = camera.Camera()
    cam.setPosition((-25.0, 45.0, -20.0))
    cam.setRotation((-0.6, -2.1, 0.1))
    _check(cam, g["pitch_yaw_roll"])
    cam = camera.Camera()
    cam.setPosition((0.25, -3.0, 12.5))
    cam.setRotation((0.3, 0.0, 0.0))
    cam.translate((1.0, 2.0, 3.0))
    cam.rotate((0, 1, 0), 0.7)
    _check(cam, g["translate_rotate"])


def test_fov_quirk_is_kept():
    """45.0 passed as radians: proj[0][0] = 1.344443, proj[1][1] = -1.792591 at 4:3."""
    p = camera.Camera().getProjectionMatrix()
    assert abs(p[0] - 1.344443) < 1e-6 and abs(p[5] + 1.792591) < 1e-6


def test_runtime_aspect():
    p43 = camera.Camera(4 / 3).getProjectionMatrix()
    p169 = camera.Camera(16 / 9).getProjectionMatrix()
    assert p43[5] == p169[5] and p169[0] < p43[0]
