"""GPU tier (`-m gpu`), last file on purpose: the C++ host mirror of the reference's Rasterizer class
(vkrt_b200/host/Rasterizer.cpp over the C ABI, `vkrt_headless --renderer raster`) renders the staged Sponza through its own
glTF + image loader; frame 0 must match the Python-side Rasterizer mirror of the same scene."""
import json
import os
import subprocess

import numpy as np
import pytest

from vkrt_b200 import binding as B
from vkrt_b200 import scenes
from vkrt_b200.raytracer import Rasterizer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def test_headless_host_rasterizer(tmp_path):
    exe = os.path.join(ROOT, "vkrt_b200", "host", "vkrt_headless")
    assets = scenes.find_asset_dir()
    if not os.path.exists(exe) or assets is None or not os.path.exists(os.path.join(assets, "Sponza.bin")):
        pytest.skip("headless executable or staged assets missing (run __graft_entry__.build())")
    W, Hh = 640, 360
    prefix = str(tmp_path / "raster_")
    r = subprocess.run([exe, "--renderer", "raster", "--width", str(W), "--height", str(Hh), "--frames", "2", "--models",
                        os.path.dirname(assets), "--out", prefix], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["renderer"] == "raster" and line["frames"] == 2 and line["triangles"] == 262267
    with open(prefix + "0000.ppm", "rb") as f:
        assert f.readline() == b"P6\n" and f.readline() == f"{W} {Hh}\n".encode() and f.readline() == b"255\n"
        rgb = np.frombuffer(f.read(), dtype=np.uint8).reshape(Hh, W, 3)
    rz = Rasterizer(scenes.sponza_standin(), W, Hh, device=0)
    rz.render()
    bgra = rz.readback(B.AOV_BGRA8)
    rz.close()
    diff = np.abs(rgb.astype(np.int16) - bgra[:, :, 2::-1].astype(np.int16))
    # stb-style JPEG decode (host loader) vs Pillow/libjpeg (Python side): texel LSBs differ, geometry does not
    assert (diff <= 4).mean() > 0.99 and diff.mean() < 1.0, (float((diff <= 4).mean()), float(diff.mean()))
