"""The reference's rasteriser path (src/Rasterizer.cpp, shaders/shader.vert, shaders/shader.frag) on the CPU tier: the
oracle's z-buffer restatement against hand-checked scenes, and the kernels' host-compiled logic (tests/emu: 64-bit
visibility keys, reverse draw order) against the oracle, bit for bit."""
import numpy as np
import pytest

import helpers as H
from vkrt_b200 import camera, scenes
from vkrt_b200.raytracer import flatten_model

IDENTITY = np.eye(4, dtype=np.float32).T.reshape(16)  # column-major identity: positions are clip coordinates
CLEAR = np.array([51, 0, 0, 255], dtype=np.uint8)     # (0,0,0.2,1) as B,G,R,A


def _tri_model(tris, images, materials=None):
    """tris: list of (3x3 positions, 3x2 uv, material).  One submesh per triangle, drawn in list order."""
    subs = []
    for pos, uv, mat in tris:
        pos = np.asarray(pos, dtype=np.float32)
        n = np.tile(np.array([[0, 0, 1]], dtype=np.float32), (3, 1))
        tan = np.tile(np.array([[1, 0, 0, 1]], dtype=np.float32), (3, 1))
        subs.append({"vertices": scenes.make_vertex_array(pos, n, np.asarray(uv, dtype=np.float32), tan),
                     "indices": np.arange(3, dtype=np.uint32), "material": mat})
    mats = materials or [(k, 0, 0) for k in range(len(images))]
    return {"submeshes": subs, "materials": mats, "images": images}


def _solid(rgba, size=4):
    return np.tile(np.array(rgba, dtype=np.uint8), (size, size, 1))


UV = [(0, 0), (1, 0), (0, 1)]
# counter-clockwise in Vulkan's y-down framebuffer (x right, y down): (-,-) -> (-,+) -> (+,-) is clockwise on screen
# when y points up, i.e. counter-clockwise in framebuffer space
FRONT = [(-0.5, -0.5, 0.5), (-0.5, 0.5, 0.5), (0.5, -0.5, 0.5)]
BACK = [FRONT[0], FRONT[2], FRONT[1]]


def _both(model, wvp, W, Hh):
    flat = flatten_model(model)
    inst = scenes.identity_instance(1.0)
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    o.generate_mips(), e.generate_mips()
    a, b = o.rasterize(wvp, W, Hh), e.rasterize(wvp, W, Hh)
    assert np.array_equal(a["tri"], b["tri"]) and np.array_equal(a["bgra"], b["bgra"])
    return a


def test_clear_cull_and_coverage():
    W = Hh = 32
    red = _solid((255, 0, 0, 255))
    a = _both(_tri_model([(FRONT, UV, 0)], [red]), IDENTITY, W, Hh)
    inside = a["tri"] == 0
    # the triangle covers half of the central 16x16 square: x_fb in [8,24], y_fb in [8,24], x + y < 32
    assert 100 < inside.sum() < 140
    assert (a["bgra"][inside & np.roll(inside, 1, 0) & np.roll(inside, -1, 0) & np.roll(inside, 1, 1) & np.roll(inside, -1, 1)] == [0, 0, 255, 255]).all()
    assert (a["bgra"][0, 0] == CLEAR).all() and (a["tri"][0, 0] == 0xFFFFFFFF)
    # antialiased edge: some pixels on the diagonal blend triangle and clear colour
    edge = (a["bgra"][:, :, 2] > 0) & (a["bgra"][:, :, 2] < 255)
    assert edge.sum() >= 10
    # back faces are culled
    b = _both(_tri_model([(BACK, UV, 0)], [red]), IDENTITY, W, Hh)
    assert (b["tri"] == 0xFFFFFFFF).all() and (b["bgra"] == CLEAR).all()


def test_depth_less_draw_order_and_clip():
    W = Hh = 32
    red, green = _solid((255, 0, 0, 255)), _solid((0, 255, 0, 255))
    near = [(x, y, 0.25) for x, y, _ in FRONT]
    # nearer triangle wins regardless of draw order
    for order in ([(FRONT, UV, 0), (near, UV, 1)], [(near, UV, 1), (FRONT, UV, 0)]):
        a = _both(_tri_model(order, [red, green]), IDENTITY, W, Hh)
        assert (a["bgra"][12, 12] == [0, 255, 0, 255]).all()
    # equal depth: the first triangle drawn keeps the sample (LESS)
    a = _both(_tri_model([(FRONT, UV, 0), (FRONT, UV, 1)], [red, green]), IDENTITY, W, Hh)
    assert (a["tri"][12, 12] == 0) and (a["bgra"][12, 12] == [0, 0, 255, 255]).all()
    # beyond the far plane or in front of the near plane: clipped; z = 1 exactly fails LESS against the cleared depth
    for z in (1.5, -0.5, 1.0):
        a = _both(_tri_model([([(x, y, z) for x, y, _ in FRONT], UV, 0)], [red]), IDENTITY, W, Hh)
        assert (a["tri"] == 0xFFFFFFFF).all()


def test_alpha_test_discards_without_depth_write():
    W = Hh = 32
    holes = _solid((255, 255, 255, 10))      # alpha 10/255 < 0.1: every fragment discarded
    thin = _solid((9, 8, 7, 26))             # alpha 26/255 = 0.102 >= 0.1: kept
    red = _solid((255, 0, 0, 255))
    near = [(x, y, 0.25) for x, y, _ in FRONT]
    a = _both(_tri_model([(near, UV, 0), (FRONT, UV, 1)], [holes, red]), IDENTITY, W, Hh)
    assert (a["tri"][12, 12] == 1) and (a["bgra"][12, 12] == [0, 0, 255, 255]).all()  # the discarded one wrote no depth
    a = _both(_tri_model([(near, UV, 0)], [thin]), IDENTITY, W, Hh)
    assert (a["bgra"][12, 12] == [7, 8, 9, 26]).all()  # outColor = color, alpha included


def test_shared_edges_are_watertight_and_not_double_covered():
    # a fan of front-facing triangles around an interior vertex at an awkward position: every sample of the covered area
    # belongs to exactly one triangle, so additive coverage would be exactly one everywhere; here: no clear colour leaks
    W = Hh = 64
    rng = np.random.default_rng(3)
    c = (0.013, -0.027, 0.5)
    ring = [(0.8 * np.cos(t), 0.8 * np.sin(t), 0.5) for t in np.linspace(0, 2 * np.pi, 12, endpoint=False) + rng.uniform(-0.1, 0.1, 12)]
    tris = []
    for k in range(12):
        p, q = ring[k], ring[(k + 1) % 12]
        tris.append(([c, q, p], UV, 0))  # framebuffer-counter-clockwise
    white = _solid((255, 255, 255, 255))
    a = _both(_tri_model(tris, [white]), IDENTITY, W, Hh)
    assert (a["tri"] != 0xFFFFFFFF).sum() > 1500
    inner = np.zeros((Hh, W), dtype=bool)
    yy, xx = np.mgrid[0:Hh, 0:W]
    inner[((xx + 0.5 - 32) ** 2 + (yy + 0.5 - 32) ** 2) < (0.8 * 32 * 0.93) ** 2] = True
    assert (a["bgra"][inner] == 255).all()  # no sample of an interior pixel shows the clear colour


def test_near_plane_crossing_perspective_and_mips():
    # a big floor seen from a camera standing on it: triangles cross the near plane (no clipping in the product)
    model = scenes.analytic_scene(metallic_floor=False)
    # the analytic floor is wound with its geometric normal pointing down (the ray tracer does not cull); wind it upward
    fv, fi = scenes.quad((-8, 0, -8), (0, 0, 16), (16, 0, 0), (0, 1, 0), uv_scale=4.0)
    model["submeshes"][0] = {"vertices": fv, "indices": fi, "material": 0}
    rng = np.random.default_rng(8)
    model["images"] = [rng.integers(0, 256, size=(s, s, 4), dtype=np.uint8) for s in (8, 64, 16, 16)]
    for im in model["images"]:
        im[..., 3] = np.where(im[..., 3] < 40, 0, 255)  # some holes for the alpha test
    W, Hh = 160, 90
    cam = camera.Camera(np.float32(W) / np.float32(Hh))
    cam.setPosition((0.0, 1.0, 6.0))
    cam.setRotation((-0.1, 0.0, 0.0))
    wvp = cam.wvp(1.0)
    a = _both(model, wvp, W, Hh)
    seen = np.unique(a["tri"])
    assert {0, 1} <= set(seen.tolist()) and (a["tri"] != 0xFFFFFFFF).mean() > 0.6  # the floor runs under the camera
    assert (a["tri"][-1] <= 1).mean() > 0.7  # bottom row: floor triangles that cross the near plane (holes show the clear colour)
    # reference camera of the rasteriser over the Sponza-scale instance matrix: just has to agree and not be empty
    cam.setPosition((0.0, 3.0, 0.0))
    cam.setRotation((-1.2, 0.3, 0.0))
    a = _both(model, cam.wvp(1.0), W, Hh)
    assert (a["tri"] != 0xFFFFFFFF).mean() > 0.3


def test_sponza_subset_parity(sponza_flat):
    # the stand-in Sponza from the rasteriser's default camera (src/Rasterizer.cpp:207-211), small frame
    W, Hh = 192, 108
    cam = camera.Camera(np.float32(W) / np.float32(Hh))
    cam.setPosition((-4.0, 2.0, -0.2))
    cam.setRotation((0.0, 1.51, 0.0))
    wvp = cam.wvp(0.01)
    o, e = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE), H.EmuScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    o.generate_mips(), e.generate_mips()
    a, b = o.rasterize(wvp, W, Hh), e.rasterize(wvp, W, Hh)
    assert np.array_equal(a["tri"], b["tri"]) and np.array_equal(a["bgra"], b["bgra"])
    assert (a["tri"] != 0xFFFFFFFF).mean() > 0.9  # inside the atrium: geometry everywhere
    assert len(np.unique(a["tri"])) > 200


def test_raster_cross_checks_ray_traced_visibility(sponza_flat):
    """the rasteriser as a primary-visibility cross-check of the ray tracer (both oracles here; the GPU tier
    repeats it with the CUDA paths).  Same camera: the triangle owning sample 0 of a pixel and the triangle hit by the
    primary ray through the pixel centre agree except at silhouettes (sample 0 sits at (9/16, 5/16), not at the centre),
    on back faces (the ray tracer does not cull) and where the alpha test opens a surface."""
    W, Hh = 320, 180
    o = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    o.generate_mips()
    cam = camera.Camera(np.float32(W) / np.float32(Hh))
    cam.setPosition((6.3, 4.5, -0.7))
    cam.setRotation((0.0, 1.57, 0.0))
    traced = o.render(cam.uniforms(), W, Hh)["tri"]
    drawn = o.rasterize(cam.wvp(0.01), W, Hh)["tri"]
    both = (traced != 0xFFFFFFFF) & (drawn != 0xFFFFFFFF)
    assert both.mean() > 0.9 and (traced[both] == drawn[both]).mean() > 0.85


def test_standin_winding_is_counter_clockwise(sponza_flat):
    # glTF: front faces wind counter-clockwise seen from the side the normal points to (the rasteriser culls the others)
    v = np.asarray(sponza_flat["vertices"]).reshape(-1, 16)
    i = np.asarray(sponza_flat["indices"]).reshape(-1, 3)
    g = np.cross(v[i[:, 1], :3] - v[i[:, 0], :3], v[i[:, 2], :3] - v[i[:, 0], :3])
    assert ((g * v[i[:, 0], 4:7]).sum(1) >= 0).mean() > 0.999


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_soup_key_min_equals_ordered_z_buffer(seed):
    """Random triangles in clip space (many crossing the near and far planes and w = 0, duplicated triangles for equal-depth
    ties, textures with transparent texels): the product's order-independent (depth, triangle) keys, filled in REVERSE draw
    order by the emulation, must give what the oracle's in-order z-buffer with depth test LESS gives."""
    rng = np.random.default_rng(seed)
    images = [rng.integers(0, 256, size=(s, s, 4), dtype=np.uint8) for s in (16, 8, 32)]
    for im in images:
        im[..., 3] = np.where(rng.random(im.shape[:2]) < 0.3, rng.integers(0, 40, im.shape[:2]), 255)
    tris = []
    for k in range(300):
        c = rng.uniform(-1.2, 1.2, 3) * np.array([1, 1, 0.6]) + np.array([0, 0, 0.5])
        pos = c + rng.normal(0, 0.35, (3, 3))
        uv = rng.uniform(-2, 3, (3, 2))
        tris.append((pos, uv, int(rng.integers(0, 3))))
        if k % 10 == 0:
            tris.append((pos.copy(), uv.copy(), int(rng.integers(0, 3))))  # exact duplicate drawn later: loses every tie
    model = _tri_model(tris, images)
    # a perspective-like matrix so that w varies and changes sign across the soup
    wvp = np.array([[1.1, 0, 0, 0], [0, -1.3, 0, 0], [0, 0, 0.7, 0.9], [0.05, 0.02, 0.2, 0.6]], dtype=np.float32).reshape(16)
    a = _both(model, wvp, 96, 64)
    drawn = a["tri"] != 0xFFFFFFFF
    assert 0.2 < drawn.mean() < 1.0 and len(np.unique(a["tri"])) > 40
