"""Mip chain (Raytracer::createMipmaps, src/Raytracer.cpp:572-640) and the optional ray-cone texture level of detail:
the oracle against independent numpy restatements, and the kernels' host-compiled logic (tests/emu) against the oracle."""
import numpy as np
import pytest

import helpers as H
from vkrt_b200 import camera, scenes
from vkrt_b200.raytracer import flatten_model


def _scene_with_images(images):
    model = scenes.analytic_scene(metallic_floor=True)
    flat = flatten_model(model)
    flat = dict(flat)
    n = len(flat["images"])
    flat["images"] = list(images) + list(flat["images"])[len(images):] if len(images) < n else list(images)
    # keep the material rows inside the image table
    assert np.asarray(flat["rows"])[:, :3].max() < len(flat["images"])
    return flat


def _random_images(seed, sizes):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8) for (w, h) in sizes]


def _blit_float64(src, dw, dh):
    """Vulkan linear blit in float64: the unrounded value of every destination texel."""
    sh, sw = src.shape[:2]
    i = (np.arange(dw) + 0.5) * (sw / dw) - 0.5
    j = (np.arange(dh) + 0.5) * (sh / dh) - 0.5
    x0f, y0f = np.floor(i), np.floor(j)
    a, b = (i - x0f)[None, :, None], (j - y0f)[:, None, None]
    x0 = np.clip(x0f.astype(int), 0, sw - 1)
    x1 = np.clip(x0f.astype(int) + 1, 0, sw - 1)
    y0 = np.clip(y0f.astype(int), 0, sh - 1)
    y1 = np.clip(y0f.astype(int) + 1, 0, sh - 1)
    s = src.astype(np.float64)
    return (s[y0][:, x0] * (1 - a) * (1 - b) + s[y0][:, x1] * a * (1 - b) + s[y1][:, x0] * (1 - a) * b + s[y1][:, x1] * a * b)


SIZES = [(64, 64), (32, 8), (5, 3), (1, 7), (17, 1), (1, 1), (6, 10)]


def test_oracle_mip_chain_matches_vulkan_blit_formula():
    imgs = _random_images(1, SIZES)
    o = H.OracleScene(_scene_with_images(imgs), scenes.identity_instance(1.0))
    assert o.level_count(0) == 1  # nothing before createMipmaps
    o.generate_mips()
    for k, (w, h) in enumerate(SIZES):
        levels = int(np.floor(np.log2(max(w, h)))) + 1  # src/Raytracer.cpp:467
        assert o.level_count(k) == levels
        prev = o.texture_level(k, 0)
        assert np.array_equal(prev, imgs[k])
        for level in range(1, levels):
            cur = o.texture_level(k, level)
            assert cur.shape == (max(h >> level, 1), max(w >> level, 1), 4)
            exact = _blit_float64(prev, cur.shape[1], cur.shape[0])
            # round to nearest; a value within float32 noise of .5 may fall either way
            assert np.abs(cur.astype(np.float64) - exact).max() <= 0.5 + 1e-3
            prev = cur
        assert prev.shape[:2] == (1, 1)


def test_oracle_mip_of_even_image_is_the_block_mean_and_constant_stays_constant():
    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, size=(16, 32, 4), dtype=np.uint8)
    const = np.full((8, 8, 4), 77, dtype=np.uint8)
    o = H.OracleScene(_scene_with_images([img, const]), scenes.identity_instance(1.0))
    o.generate_mips()
    mean = img.astype(np.float64).reshape(8, 2, 16, 2, 4).mean(axis=(1, 3))
    assert np.abs(o.texture_level(0, 1).astype(np.float64) - mean).max() <= 0.5 + 1e-3
    for level in range(4):
        assert (o.texture_level(1, level) == 77).all()


def test_emu_mip_chain_equals_oracle_bit_for_bit(sponza_flat):
    imgs = _random_images(3, SIZES)
    flat = _scene_with_images(imgs)
    o, e = H.OracleScene(flat, scenes.identity_instance(1.0)), H.EmuScene(flat, scenes.identity_instance(1.0))
    o.generate_mips(), e.generate_mips()
    for k in range(len(SIZES)):
        for level in range(o.level_count(k)):
            assert np.array_equal(o.texture_level(k, level), e.texture_level(k, level)), (k, level)
    # real textures of the stand-in scene
    o, e = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE), H.EmuScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    o.generate_mips(), e.generate_mips()
    for k in (0, 3, len(sponza_flat["images"]) - 1):
        for level in range(o.level_count(k)):
            assert np.array_equal(o.texture_level(k, level), e.texture_level(k, level)), (k, level)


def test_lod_sampling_oracle_properties_and_emu_bit_equality():
    imgs = _random_images(4, [(64, 32), (5, 3), (1, 1)])
    flat = _scene_with_images(imgs)
    o, e = H.OracleScene(flat, scenes.identity_instance(1.0)), H.EmuScene(flat, scenes.identity_instance(1.0))
    o.generate_mips(), e.generate_mips()
    rng = np.random.default_rng(5)
    # footprints of at most one texel (and NaN, negative, zero) sample level 0
    for fp in (0.0, -1.0, 1.0 / (64 * 32), 1e-9, float("nan")):
        u, v = rng.uniform(-2, 2, 2)
        assert np.array_equal(o.sample_texture_lod(0, u, v, fp), o.sample_texture(0, u, v))
    # a footprint far larger than the image samples the 1x1 level
    top = o.texture_level(0, o.level_count(0) - 1).reshape(4).astype(np.float32) / np.float32(255.0)
    assert np.array_equal(o.sample_texture_lod(0, 0.3, 0.7, 1e6), top)
    assert np.array_equal(o.sample_texture_lod(0, 0.3, 0.7, float("inf")), o.sample_texture(0, 0.3, 0.7))
    # an exact level: footprint2 * W * H = 4^k  ->  lambda = k, a single bilinear tap of level k
    lvl2 = o.texture_level(0, 2)
    single = H.OracleScene(_scene_with_images([lvl2]), scenes.identity_instance(1.0))
    assert np.array_equal(o.sample_texture_lod(0, 0.41, 0.13, 16.0 / (64 * 32)), single.sample_texture(0, 0.41, 0.13))
    # lambda is continuous and monotone in the footprint: the blend moves from level k to k+1
    a = o.sample_texture_lod(0, 0.41, 0.13, 16.0 / (64 * 32))
    b = o.sample_texture_lod(0, 0.41, 0.13, 64.0 / (64 * 32))
    mid = o.sample_texture_lod(0, 0.41, 0.13, 32.0 / (64 * 32))  # lambda = 2.5
    assert np.allclose(mid, 0.5 * (a + b), atol=2e-4)
    # kernels' logic == oracle, bit for bit, over random footprints
    for image in range(3):
        for _ in range(300):
            u, v = rng.uniform(-3, 3, 2)
            fp = float(10.0 ** rng.uniform(-7, 3))
            assert np.array_equal(o.sample_texture_lod(image, u, v, fp), e.sample_texture_lod(image, u, v, fp)), (image, u, v, fp)
    for fp in (0.0, float("nan"), float("inf"), -3.0, 1e30, 1e-30):
        assert np.array_equal(o.sample_texture_lod(0, 0.2, 0.9, fp), e.sample_texture_lod(0, 0.2, 0.9, fp), equal_nan=True)


def test_lod_log2_contract_is_accurate():
    # lambda = 0.5 * lodLog2(x): recover it from the blend weight between two constant levels
    lo = np.zeros((2, 2, 4), dtype=np.uint8)
    lo[...] = 0
    img = np.zeros((2, 2, 4), dtype=np.uint8)
    img[0, 0] = img[1, 1] = 255  # level 1 (1x1) = round(127.5) = 128, level 0 at a texel centre = 255 or 0
    o = H.OracleScene(_scene_with_images([img]), scenes.identity_instance(1.0))
    o.generate_mips()
    l1 = float(o.texture_level(0, 1)[0, 0, 0]) / 255.0
    for x in np.geomspace(1.001, 3.999, 40):  # texels2 in (1, 4) -> lambda in (0, 1)
        got = float(o.sample_texture_lod(0, 0.25, 0.25, x / 4.0)[0])  # texel centre of (0,0): level 0 = 1.0
        lam = (1.0 - got) / (1.0 - l1)
        assert abs(lam - 0.5 * np.log2(x)) < 2e-4


def test_frame_parity_lod_mode_analytic_and_instances():
    # noise textures of different, partly non-square sizes (the analytic scene's own are solid colours)
    flat = _scene_with_images(_random_images(6, [(16, 16), (64, 64), (32, 16), (8, 24)]))
    inst = np.stack([scenes.identity_instance(1.0), scenes.identity_instance(0.5, (10, 0, -4)), scenes.identity_instance(1.0, (-18, 1, 2))])
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    W, Hh = 128, 72
    u = camera.make_uniforms(W, Hh, position=(0, 5, 16), rotation=(-0.25, 0.1, 0))
    base = o.render(u, W, Hh)
    for s in (o, e):
        s.generate_mips()
        s.set_texture_lod(1)
    a, b = o.render(u, W, Hh), e.render(u, W, Hh)
    c = H.compare_frames(a, b)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"] and c["bgra_max_diff"] == 0
    assert np.array_equal(a["tri"], base["tri"])  # geometry is untouched by the texture mode
    assert (np.abs(a["rgb"] - base["rgb"]) > 1e-3).mean() > 0.2  # noise textures under minification: the filter is live
    o.set_texture_lod(0)
    again = o.render(u, W, Hh)
    assert np.array_equal(again["rgb"], base["rgb"])  # and the mode switches back to the reference's sampling


def test_frame_parity_lod_mode_sponza_subset(sponza_flat):
    o, e = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE), H.EmuScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    W, Hh = 640, 360
    u = camera.make_uniforms(W, Hh)
    base = o.render(u, W, Hh, stride=(5, 5))
    for s in (o, e):
        s.generate_mips()
        s.set_texture_lod(1)
    a, b = o.render(u, W, Hh, stride=(5, 5)), e.render(u, W, Hh, stride=(5, 5))
    sel = np.zeros((Hh, W), dtype=bool)
    sel[::5, ::5] = True
    c = H.compare_frames(a, b, sel)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] >= 0.999 and c["rgb_bit_equal"]
    assert np.array_equal(a["tri"][sel], base["tri"][sel])
    # the mode is live: minified textures are filtered, so colours move (but stay the same picture)
    diff = np.abs(a["rgb"][sel] - base["rgb"][sel])
    assert (diff > 1e-4).mean() > 0.05
    assert np.nanmean(diff) < 0.2 * np.nanmean(base["rgb"][sel])


def test_frame_parity_lod_mode_random_soup():
    """Random small triangles with random uv (degenerate uv areas, grazing angles, reflections off random normals) and noise
    textures: the ray-cone footprint takes every branch (magnification, clamped top level, NaN footprints)."""
    flat = flatten_model(scenes.random_soup(4000, seed=11, geometries=5, images=4))
    inst = scenes.identity_instance(1.0)
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    for s in (o, e):
        s.generate_mips()
        s.set_texture_lod(1)
    W, Hh = 96, 64
    u = camera.make_uniforms(W, Hh, position=(0, 2, 14), rotation=(-0.1, 0.0, 0))
    a, b = o.render(u, W, Hh, spp=2), e.render(u, W, Hh, spp=2)
    c = H.compare_frames(a, b)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"]
    assert (a["tri"] != 0xFFFFFFFF).mean() > 0.3
