"""GPU tier (`-m gpu`): the mip chain built by k_mip_level (Raytracer::createMipmaps, src/Raytracer.cpp:572-640) and the
optional ray-cone texture level of detail of the shading kernels, through the C ABI, against the CPU oracle.  Bit equality
is the bar for both: the chain is byte work, and the level-of-detail arithmetic is fixed on both sides."""
import numpy as np
import pytest

import helpers as H
from vkrt_b200 import binding as B
from vkrt_b200 import scenes
from vkrt_b200.raytracer import Raytracer, flatten_model

pytestmark = pytest.mark.gpu

SIZES = [(64, 64), (32, 8), (5, 3), (1, 7), (17, 1), (1, 1), (6, 10), (1024, 512), (300, 200)]


def _random_images(seed, sizes):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8) for (w, h) in sizes]


def gpu_frame(rt):
    return {"rgb": rt.readback(B.AOV_RGB32F), "bgra": rt.readback(B.AOV_BGRA8), "tri": rt.readback(B.AOV_TRIANGLE_ID),
            "mask": rt.readback(B.AOV_SHADOW_MASK)}


def test_mip_chain_equals_oracle_and_api_errors():
    imgs = _random_images(11, SIZES)
    ctx = B.Context(0)
    with pytest.raises(B.VkrtError):  # no textures yet
        ctx.generate_mips()
    ctx.set_textures(imgs)
    assert ctx.texture_info(0) == (64, 64, 1)
    with pytest.raises(B.VkrtError):  # the ray-cone mode needs the chain
        ctx.set_texture_lod(B.TEXTURE_LOD_RAY_CONES)
    with pytest.raises(B.VkrtError):
        ctx.set_texture_lod(7)
    before = ctx.stats().kernelLaunches
    ctx.generate_mips()
    assert ctx.stats().kernelLaunches - before == 10  # one launch per level over all images: 1024 -> 11 levels
    model = scenes.analytic_scene()
    flat = dict(flatten_model(model))
    flat["images"] = imgs
    o = H.OracleScene(flat, scenes.identity_instance(1.0))
    o.generate_mips()
    for k, (w, h) in enumerate(SIZES):
        levels = o.level_count(k)
        assert levels == int(np.floor(np.log2(max(w, h)))) + 1
        for level in range(levels):
            assert ctx.texture_info(k, level) == (max(w >> level, 1), max(h >> level, 1), levels)
            assert np.array_equal(ctx.read_texture(k, level), o.texture_level(k, level)), (k, level)
        with pytest.raises(B.VkrtError):
            ctx.texture_info(k, levels)
    with pytest.raises(B.VkrtError):
        ctx.texture_info(len(SIZES))
    ctx.set_texture_lod(B.TEXTURE_LOD_RAY_CONES)
    ctx.set_texture_lod(B.TEXTURE_LOD_ZERO)
    # a new texture set discards the chain
    ctx.set_textures(imgs[:2])
    assert ctx.texture_info(0) == (64, 64, 1)
    with pytest.raises(B.VkrtError):
        ctx.set_texture_lod(B.TEXTURE_LOD_RAY_CONES)
    ctx.generate_mips()  # and it can be rebuilt
    assert np.array_equal(ctx.read_texture(1, 3), o.texture_level(1, 3))
    ctx.close()


def test_frame_parity_ray_cone_lod_analytic_instances():
    model = scenes.analytic_scene(metallic_floor=True)
    model["images"] = _random_images(6, [(16, 16), (64, 64), (32, 16), (8, 24)])
    inst = np.stack([scenes.identity_instance(1.0), scenes.identity_instance(0.5, (10, 0, -4)), scenes.identity_instance(1.0, (-18, 1, 2))])
    W, Hh = 200, 120
    rt = Raytracer(model, W, Hh, instances=inst)
    rt.camera.setPosition((0, 5, 16))
    rt.camera.setRotation((-0.25, 0.1, 0))
    rt.render()
    base = gpu_frame(rt)
    o = H.OracleScene(rt.flat, inst)
    c0 = H.compare_frames(o.render(rt.camera.uniforms(), W, Hh), base)
    assert c0["tri_agree"] == 1.0 and c0["rgb_bit_equal"], c0  # generating the chain does not touch the default mode
    rt.set_texture_lod(B.TEXTURE_LOD_RAY_CONES)
    rt.render()
    g = gpu_frame(rt)
    o.generate_mips()
    o.set_texture_lod(1)
    a = o.render(rt.camera.uniforms(), W, Hh)
    c = H.compare_frames(a, g)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"] and c["bgra_max_diff"] == 0, c
    assert np.array_equal(g["tri"], base["tri"])
    assert (np.abs(g["rgb"] - base["rgb"]) > 1e-3).mean() > 0.2  # the filter is live
    rt.set_texture_lod(B.TEXTURE_LOD_ZERO)
    rt.render()
    assert np.array_equal(gpu_frame(rt)["rgb"], base["rgb"])
    rt.close()


def test_frame_parity_ray_cone_lod_sponza_spp(sponza_flat):
    """Sponza stand-in, 640x360: 1 spp full frame and 8 spp (pixel-major waves, reflection bounce) on a pixel subset."""
    W, Hh = 640, 360
    rt = Raytracer(scenes.sponza_standin(), W, Hh)
    rt.set_texture_lod(B.TEXTURE_LOD_RAY_CONES)
    o = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    o.generate_mips()
    o.set_texture_lod(1)
    rt.render()
    c = H.compare_frames(o.render(rt.camera.uniforms(), W, Hh), gpu_frame(rt))
    print("ray-cone LOD parity, 1 spp:", c)
    assert c["tri_agree"] >= 0.999 and c["mask_agree"] >= 0.999 and c["bgra_max_diff"] <= 1, c
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"], c
    rt.render(spp=8)
    sel = np.zeros((Hh, W), dtype=bool)
    sel[::3, ::3] = True
    c = H.compare_frames(o.render(rt.camera.uniforms(), W, Hh, spp=8, stride=(3, 3)), gpu_frame(rt), sel)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"], c
    rt.close()
