"""CPU tier: the C-ABI library loads without a GPU and exports every symbol include/vkrt.h declares; struct layouts
match the header; creating a context without a device fails loudly (there is no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from vkrt_b200 import binding as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_are_exported():
    header = open(os.path.join(ROOT, "include", "vkrt.h")).read()
    declared = sorted(set(re.findall(r"VKRT_API\s+[\w\s\*]+?\b(vkrt_\w+)\s*\(", header)))
    assert declared == sorted(B.EXPORTS), set(declared) ^ set(B.EXPORTS)
    lib = C.CDLL(B.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_layouts():
    assert C.sizeof(B.Instance) == 64       # VkAccelerationStructureInstanceKHR
    assert C.sizeof(B.Uniforms) == 256      # UniformBufferInfo, src/Raytracer.cpp:15-24
    assert C.sizeof(B.SubmeshInfo) == 16    # SubmeshInfo, src/Raytracer.cpp:33-39
    assert C.sizeof(B.GeometryRange) == 16
    assert B.NODE_DTYPE.itemsize == 80 and B.TRI_DTYPE.itemsize == 48
    assert B.Instance.accelerationStructureReference.offset == 56


def test_version_and_tile_owner_need_no_gpu():
    lib = B.load_library()
    assert lib.vkrt_version() >= 100
    tiles_x = (1920 + 63) // 64
    assert lib.vkrt_tile_owner(1920, 0, 0, 8) == 0
    assert lib.vkrt_tile_owner(1920, 64, 0, 8) == 1
    assert lib.vkrt_tile_owner(1920, 5, 64, 8) == tiles_x % 8
    assert lib.vkrt_tile_owner(1920, 700, 700, 1) == 0


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(B.VkrtError) as e:
        B.Context(0)
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)
