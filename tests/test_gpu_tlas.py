"""GPU tier: the top-level acceleration structure (replaces Raytracer::createTLAS, src/Raytracer.cpp:1159-1283)
and the argument checks the C ABI makes before it launches anything."""
import numpy as np
import pytest

import helpers as H
from bvh_check import check_tlas, instance_world_boxes
from vkrt_b200 import binding as B
from vkrt_b200 import scenes
from vkrt_b200.raytracer import Raytracer, flatten_model

pytestmark = pytest.mark.gpu


def test_top_level_4096_instances_invariants_and_rays():
    """A 64 x 64 grid plus randomly rotated/scaled copies: 4,096 instances of one BLAS (the old flat list stopped at 256).
    Hierarchy invariants, then closest and any hit against the oracle, which loops over every instance."""
    flat = flatten_model(scenes.random_soup(300, seed=12))
    grid = np.stack([scenes.identity_instance(0.2, (6.0 * i, 0.0, 6.0 * j)) for j in range(48) for i in range(64)])
    rnd = scenes.random_instance_transforms(4096 - len(grid), seed=4, extent=(380.0, 40.0, 290.0), scale=(0.1, 0.6))
    rnd[:, 3] += 190.0
    rnd[:, 11] += 145.0
    inst = np.concatenate([grid, rnd]).astype(np.float32)
    assert inst.shape == (4096, 12)
    ctx = B.Context(0)
    blas = B.Blas(ctx, flat["vertices"], flat["indices"], flat["ranges"])
    tlas = B.Tlas(ctx, [blas], inst)
    info, binfo = tlas.info(), blas.info()
    assert info.instanceCount == 4096 and info.nodeCount > 4096 // 8 and 3 <= info.levels <= 16 and info.buildMilliseconds > 0
    nodes, leaves = tlas.read()
    lo, hi = instance_world_boxes(np.array(binfo.boundsMin), np.array(binfo.boundsMax), inst)
    assert check_tlas(nodes, leaves, lo, hi) == info.levels
    o = H.OracleScene(flat, inst)
    rays = H.random_rays(20000, 3, (-10, -15, -10), (390, 25, 300))
    rays["tmax"][::4] = 40.0
    # axis-parallel rays whose origin lies exactly in a box plane of the grid (0 * inf in the slab test)
    rays["origin"][:64] = [(6.0 * i, 0.1, -5.0) for i in range(64)]
    rays["direction"][:64] = (0.0, 0.0, 1.0)
    ho, _ = o.trace_rays(rays)
    hg = tlas.trace_rays(rays)
    assert np.array_equal(ho, hg)
    assert len(np.unique(hg["instance"][hg["instance"] != 0xFFFFFFFF])) > 1000
    ao, _ = o.trace_rays(rays, any_hit=True)
    ag = tlas.trace_rays(rays, any_hit=True)
    assert np.array_equal(ao["instance"] != 0xFFFFFFFF, ag["instance"] != 0xFFFFFFFF)
    # instance masks: every third instance invisible (mask 0), the rest with assorted non-zero masks (cullMask is 0xFF)
    masks = np.where(np.arange(4096) % 3 == 0, 0, 1 << (np.arange(4096) % 8)).astype(np.uint8)
    tlas_m = B.Tlas(ctx, [blas], inst, masks=masks)
    o.set_instance_masks(masks)
    hm = tlas_m.trace_rays(rays)
    hom, _ = o.trace_rays(rays)
    assert np.array_equal(hom, hm)
    hit = hm["instance"] != 0xFFFFFFFF
    assert hit.any() and (hm["instance"][hit] % 3 != 0).all() and not np.array_equal(hm, hg)
    tlas_m.close(), tlas.close(), blas.close(), ctx.close()


def test_two_blases_in_one_top_level_and_stale_detection():
    """Instances of two different BLASes; rebuilding one of them makes the TLAS stale (VKRT_ERROR_NOT_READY) until it is
    built again; destroying one does too."""
    fa, fb = flatten_model(scenes.random_soup(2000, seed=1)), flatten_model(scenes.random_soup(700, seed=2))
    ctx = B.Context(0)
    a = B.Blas(ctx, fa["vertices"], fa["indices"], fa["ranges"])
    b = B.Blas(ctx, fb["vertices"], fb["indices"], fb["ranges"], flags=B.BUILD_PREFER_FAST_BUILD)
    inst = scenes.random_instance_transforms(40, seed=7, extent=(60.0, 20.0, 60.0), scale=(0.5, 1.2))
    blases = [a if i % 2 == 0 else b for i in range(40)]
    tlas = B.Tlas(ctx, blases, inst)
    rays = H.random_rays(30000, 8, (-40, -15, -40), (40, 15, 40))
    hits = tlas.trace_rays(rays)
    # against the two single-BLAS oracles: the closer of the two answers, ties to the lower instance index
    oa, ob = H.OracleScene(fa, inst[0::2]), H.OracleScene(fb, inst[1::2])
    ha, _ = oa.trace_rays(rays)
    hb, _ = ob.trace_rays(rays)
    ia = np.where(ha["instance"] != 0xFFFFFFFF, ha["instance"] * 2, 0xFFFFFFFF).astype(np.uint32)
    ib = np.where(hb["instance"] != 0xFFFFFFFF, hb["instance"] * 2 + 1, 0xFFFFFFFF).astype(np.uint32)
    ta = np.where(ia != 0xFFFFFFFF, ha["t"], np.inf)
    tb = np.where(ib != 0xFFFFFFFF, hb["t"], np.inf)
    use_a = (ta < tb) | ((ta == tb) & (ia < ib))
    want_inst = np.where(use_a, ia, ib)
    assert np.array_equal(hits["instance"], want_inst)
    sel = want_inst != 0xFFFFFFFF
    assert np.array_equal(hits["t"][sel], np.where(use_a, ha["t"], hb["t"])[sel])
    assert np.array_equal(hits["primitive"][sel], np.where(use_a, ha["primitive"], hb["primitive"])[sel])
    moved = fb["vertices"].copy()
    moved[:, :3] += np.float32(0.25)
    b.rebuild(moved)
    with pytest.raises(B.VkrtError) as e:
        tlas.trace_rays(rays[:10])
    assert e.value.code == 6 and "rebuilt" in str(e.value)
    tlas2 = B.Tlas(ctx, blases, inst)
    assert not np.array_equal(tlas2.trace_rays(rays), hits)
    b.close()
    with pytest.raises(B.VkrtError) as e:
        tlas2.trace_rays(rays[:10])
    assert e.value.code == 6 and "destroyed" in str(e.value)
    with pytest.raises(B.VkrtError):
        B.Tlas(ctx, blases, inst)  # references a destroyed BLAS
    tlas.close(), tlas2.close(), a.close(), ctx.close()


def test_vertex_stride_contract():
    """a 16-byte-stride BLAS builds and answers ray queries; shading and rasterising it is refused
    (they read the 64-byte Model::Vertex with 16-byte loads) instead of faulting."""
    flat = flatten_model(scenes.random_soup(500, seed=3))
    v16 = np.ascontiguousarray(flat["vertices"][:, :4])
    ctx = B.Context(0)
    blas16 = B.Blas(ctx, v16, flat["indices"], flat["ranges"], stride=16)
    blas64 = B.Blas(ctx, flat["vertices"], flat["indices"], flat["ranges"])
    inst = scenes.identity_instance(1.0)
    t16, t64 = B.Tlas(ctx, [blas16], inst), B.Tlas(ctx, [blas64], inst)
    rays = H.random_rays(20000, 2, (-12, -8, -10), (12, 8, 10))
    assert np.array_equal(t16.trace_rays(rays), t64.trace_rays(rays))
    ctx.set_materials(flat["rows"])
    ctx.set_textures(flat["images"])
    ctx.generate_mips()
    frame = B.Frame(ctx, 64, 64)
    from vkrt_b200 import camera

    u = camera.make_uniforms(64, 64)
    frame.trace(t64, u)
    with pytest.raises(B.VkrtError) as e:
        frame.trace(t16, u)
    assert e.value.code == 1 and "vertexStride" in str(e.value)
    with pytest.raises(B.VkrtError) as e:
        frame.rasterize(blas16, np.eye(4, dtype=np.float32))
    assert e.value.code == 1
    ctx.synchronize()
    frame.close(), t16.close(), t64.close(), blas16.close(), blas64.close(), ctx.close()


def test_material_table_contract():
    """base colour index -1 (legal output of the reference's Model: only the other two indices are clamped,
    src/Raytracer.cpp:1424-1426) and indices past the texture array sample image 0, like the oracle; a table shorter
    than the BLAS's geometry count is refused."""
    model = scenes.analytic_scene(metallic_floor=True)
    W, Hh = 160, 90
    rt = Raytracer(model, W, Hh)
    rows = rt.flat["rows"].copy()
    rows[0, 0] = -1
    rows[-1, 1] = 1000
    rt.ctx.set_materials(rows)
    rt.camera.setPosition((0.3, 1.4, 4.2)), rt.camera.setRotation((0.05, 0.2, 0.0))
    rt.render()
    g = {"rgb": rt.readback(B.AOV_RGB32F), "bgra": rt.readback(B.AOV_BGRA8), "tri": rt.readback(B.AOV_TRIANGLE_ID),
         "mask": rt.readback(B.AOV_SHADOW_MASK)}
    flat = dict(rt.flat)
    flat["rows"] = rows
    a = H.OracleScene(flat, rt.instances).render(rt.camera.uniforms(), W, Hh)
    c = H.compare_frames(a, g)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"], c
    if len(rows) > 1:
        rt.ctx.set_materials(rows[:-1])
        with pytest.raises(B.VkrtError) as e:
            rt.render()
        assert e.value.code == 1 and "material table" in str(e.value)
    rt.close()
