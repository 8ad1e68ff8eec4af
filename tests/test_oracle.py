"""CPU tier: pins the oracle (brute force == BVH, analytic pixels, texture sampling, Halton, affine inverse)."""
import numpy as np
import pytest

import helpers as H
from vkrt_b200 import camera, scenes
from vkrt_b200.raytracer import flatten_model


def _analytic(metallic=False):
    flat = flatten_model(scenes.analytic_scene(metallic_floor=metallic))
    return flat, H.OracleScene(flat, scenes.identity_instance(1.0))


def test_brute_force_equals_bvh_analytic():
    flat, o = _analytic(True)
    u = camera.make_uniforms(96, 64, position=(0, 3, 9), rotation=(-0.2, 0, 0))
    a = o.render(u, 96, 64, mode=H.MODE_BRUTE)
    b = o.render(u, 96, 64, mode=H.MODE_BVH)
    c = H.compare_frames(a, b)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"]
    assert (a["tri"] != 0xFFFFFFFF).mean() > 0.5


@pytest.mark.parametrize("seed", [1, 2])
def test_brute_force_equals_bvh_soup_rays(seed):
    flat = flatten_model(scenes.random_soup(4000, seed=seed))
    o = H.OracleScene(flat, scenes.identity_instance(1.0))
    rays = H.random_rays(20000, seed + 10, (-12, -8, -10), (12, 8, 10))
    h1, _ = o.trace_rays(rays, mode=H.MODE_BRUTE)
    h2, _ = o.trace_rays(rays, mode=H.MODE_BVH)
    assert np.array_equal(h1, h2)
    a1, _ = o.trace_rays(rays, any_hit=True, mode=H.MODE_BRUTE)
    a2, _ = o.trace_rays(rays, any_hit=True, mode=H.MODE_BVH)
    assert np.array_equal(a1["instance"] != 0xFFFFFFFF, a2["instance"] != 0xFFFFFFFF)
    assert 0.2 < (h1["instance"] != 0xFFFFFFFF).mean() < 0.9


def test_brute_force_equals_bvh_sponza_subset(sponza_flat):
    o = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    W, Hh = 1280, 720
    u = camera.make_uniforms(W, Hh)
    a = o.render(u, W, Hh, mode=H.MODE_BRUTE, stride=(97, 61))  # 14 x 12 pixels, ~1.5e8 ray-triangle tests
    b = o.render(u, W, Hh, mode=H.MODE_BVH, stride=(97, 61))
    sel = np.zeros((Hh, W), dtype=bool)
    sel[::61, ::97] = True
    c = H.compare_frames(a, b, sel)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"]


def test_analytic_pixel_matches_float64_shading():
    """One pixel of the analytic scene recomputed independently in float64 straight from the GLSL formulas."""
    flat, o = _analytic(False)
    W, Hh = 64, 48
    u = camera.make_uniforms(W, Hh, position=(0, 3, 9), rotation=(-0.2, 0, 0))
    r = o.render(u, W, Hh, mode=H.MODE_BRUTE)
    x, y = 40, 30
    ua = camera.uniforms_to_array(u).astype(np.float64)
    vi, pi = ua[:16].reshape(4, 4).T, ua[16:32].reshape(4, 4).T  # column-major -> matrices
    uv = (np.array([x + 0.5, y + 0.5]) / np.array([W, Hh])) * 2 - 1
    tgt = pi @ np.array([uv[0], uv[1], 1, 1])
    d = (vi @ np.append(tgt[:3] / np.linalg.norm(tgt[:3]), 0))[:3]
    org = (vi @ np.array([0, 0, 0, 1.0]))[:3]
    assert d[1] < 0
    t = -org[1] / d[1]
    pos = org + t * d  # on the floor y = 0
    assert abs(pos[0]) < 8 and abs(pos[2]) < 8
    n_map = np.array([128, 128, 255]) / 255.0 * 2 - 1
    n_map /= np.linalg.norm(n_map)
    T, N = np.array([1.0, 0, 0]), np.array([0, 1.0, 0])
    Bv = np.cross(T, N)
    n = T * n_map[0] + Bv * n_map[1] + N * n_map[2]
    n /= np.linalg.norm(n)
    total = 0.0
    occl_lo, occl_hi = np.array([-1, -1]), np.array([1, 1])
    for L in camera.LIGHT_POSITIONS[:, :3].astype(np.float64):
        lv = L - pos
        dist = np.linalg.norm(lv)
        ld = lv / dist
        ndl = float(n @ ld)
        mult = 1.0
        if ndl > 0:
            # the only possible occluder is the quad at y = 2, x,z in [-1,1] (the wall is behind the floor's far edge)
            s = (2 - pos[1]) / ld[1]
            hit = pos + s * ld
            if 0.001 < s < dist and (hit[[0, 2]] >= occl_lo).all() and (hit[[0, 2]] <= occl_hi).all():
                mult = 0.3
        total += min(max(ndl, 0), 1) * (10.0 / dist ** 2) * mult
    expected = np.ones(3) * total + np.ones(3) * 0.1
    assert np.allclose(r["rgb"][y, x], expected, rtol=2e-5, atol=2e-6)
    bgra = r["bgra"][y, x]
    assert bgra[3] == 255 and bgra[2] == int(min(max(expected[0], 0), 1) * 255 + 0.5)


def test_miss_is_sky_and_reflection_adds_sky():
    flat, o = _analytic(True)
    W, Hh = 32, 24
    u = camera.make_uniforms(W, Hh, position=(0, 3, 9), rotation=(0.9, 0, 0))  # looking up: all sky
    r = o.render(u, W, Hh)
    assert (r["tri"] == 0xFFFFFFFF).all()
    assert np.allclose(r["rgb"], np.array([0.8, 0.8, 1.0], dtype=np.float32))
    assert (r["bgra"][..., 0] == 255).all() and (r["bgra"][..., 1] == 204).all() and (r["bgra"][..., 2] == 204).all()
    # metallic floor: a pixel looking at the floor reflects into the sky -> final = shaded*(1-att) + sky
    u = camera.make_uniforms(W, Hh, position=(0, 3, 9), rotation=(-0.6, 0, 0))
    r = o.render(u, W, Hh)
    assert r["stats"].raysReflection > 0
    floor = r["tri"] < 2
    assert floor.any() and (r["rgb"][floor][:, 2] > 1.0).any()


def test_texture_sampling_bilinear_repeat():
    flat = flatten_model(scenes.analytic_scene())
    img = np.zeros((2, 2, 4), dtype=np.uint8)
    img[0, 0] = (0, 0, 0, 255)
    img[0, 1] = (255, 0, 0, 255)
    img[1, 0] = (0, 255, 0, 255)
    img[1, 1] = (0, 0, 255, 255)
    flat["images"] = [img]
    flat["rows"][:, :3] = 0
    o = H.OracleScene(flat, scenes.identity_instance(1.0))
    assert np.allclose(o.sample_texture(0, 0.25, 0.25), [0, 0, 0, 1])          # texel centre (0,0)
    assert np.allclose(o.sample_texture(0, 0.75, 0.25), [1, 0, 0, 1])          # texel centre (1,0)
    assert np.allclose(o.sample_texture(0, 0.5, 0.25), [0.5, 0, 0, 1])         # halfway
    assert np.allclose(o.sample_texture(0, 0.0, 0.25), [0.5, 0, 0, 1])         # REPEAT: wraps to texel 1
    assert np.allclose(o.sample_texture(0, 1.25, -0.75), o.sample_texture(0, 0.25, 0.25))
    assert np.allclose(o.sample_texture(0, 0.5, 0.5), [0.25, 0.25, 0.25, 1])


def test_halton_and_affine_inverse():
    lib = H.oracle_lib()
    out = np.zeros(2, dtype=np.float32)
    exp = [(0.5, 1 / 3), (0.25, 2 / 3), (0.75, 1 / 9), (0.125, 4 / 9)]
    for s, e in enumerate(exp):
        lib.orc_halton23(s, out.ctypes.data)
        assert np.allclose(out, e, atol=1e-7)
    m = np.array([0.01, 0, 0, 40, 0, 0.01, 0, 0, 0, 0, 0.01, 25], dtype=np.float32)
    inv = np.zeros(12, dtype=np.float32)
    lib.orc_affine_inverse(m.ctypes.data, inv.ctypes.data)
    assert inv[0] == np.float32(100) and np.allclose(inv[[3, 7, 11]], [-4000, 0, -2500], rtol=1e-6)


def test_nan_tangent_path_is_defined(sponza_flat):
    """Primitive 50 has zero tangents -> NaN perturbed normal: no shadow rays, ambient only, NaN reflection ray misses."""
    o = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    first = int(sponza_flat["rows"][50, 3])
    v = sponza_flat["vertices"]
    idx = sponza_flat["indices"][3 * first:3 * first + 3]
    c = v[idx, :3].mean(axis=0) * 0.01
    n = v[idx[0], 4:7]
    rays = np.zeros(1, dtype=H.B.RAY_DTYPE)
    rays["origin"] = c + n * 0.5
    rays["direction"] = -n
    rays["tmin"], rays["tmax"] = 0.001, 1000
    h, _ = o.trace_rays(rays)
    assert h["geometry"][0] == 50


# ---------------------------------------------------------------------------------------------------------------
# Closed-form check of whole frames (textured quad, two-quad occluder, mirror quad).  The analytic scene is
# three axis-aligned quads, so every ray/quad intersection, every shadow test and the reflection bounce have closed
# forms; _Quads64 evaluates shader.rgen / shader.rchit / shader.rmiss straight from the GLSL in float64, sharing no
# code with the oracle (no BVH, no watertight test, no texture sampler: the textures are constant images).
# ---------------------------------------------------------------------------------------------------------------
class _Quads64:
    # (origin, edge u, edge v, normal, tangent, metallic blue channel) of floor, occluder, wall - scenes.analytic_scene
    def __init__(self, metallic_floor):
        f8 = np.float64
        self.quads = [
            (np.array([-8, 0, -8], f8), np.array([16, 0, 0], f8), np.array([0, 0, 16], f8), np.array([0, 1, 0], f8), np.array([1, 0, 0], f8), 1.0 if metallic_floor else 0.0),
            (np.array([-1, 2, -1], f8), np.array([2, 0, 0], f8), np.array([0, 0, 2], f8), np.array([0, 1, 0], f8), np.array([1, 0, 0], f8), 0.0),
            (np.array([-8, 0, -8], f8), np.array([16, 0, 0], f8), np.array([0, 8, 0], f8), np.array([0, 0, 1], f8), np.array([1, 0, 0], f8), 0.0),
        ]
        self.lights = camera.LIGHT_POSITIONS[:, :3].astype(f8)

    def hit(self, o, d, tmin, tmax):
        """(t, quad, margin): nearest quad with tmin < t < tmax; margin = distance (in quad-parameter units) of the hit from
        the nearest quad edge, over all candidate quads - tiny margins mean float32 may legitimately decide otherwise."""
        best, margin = (np.inf, -1), np.inf
        for k, (p0, pu, pv, n, _, _) in enumerate(self.quads):
            dn = d @ n
            if dn == 0:
                continue
            t = ((p0 - o) @ n) / dn
            if not (tmin < t < tmax):
                continue
            q = o + t * d - p0
            a, b = (q @ pu) / (pu @ pu), (q @ pv) / (pv @ pv)
            margin = min(margin, abs(a), abs(1 - a), abs(b), abs(1 - b))
            if 0 <= a <= 1 and 0 <= b <= 1 and t < best[0]:
                best = (t, k)
        return best[0], best[1], margin

    def shade(self, o, d, attenuation):
        """shader.rchit:86-171 / shader.rmiss:17 -> (hitValue, done, next origin, next direction, next attenuation, mask, margin)."""
        t, k, margin = self.hit(o, d, 0.001, 1000.0)
        if k < 0:
            return np.array([0.8, 0.8, 1.0]), 1, None, None, attenuation, 0, margin
        p0, pu, pv, n, tan, metallic = self.quads[k]
        pos = o + t * d
        n_map = np.array([128, 128, 255]) / 255.0 * 2 - 1
        n_map /= np.linalg.norm(n_map)
        bit = np.cross(tan, n)  # getTBN, rchit:78-84
        nn = tan * n_map[0] + bit * n_map[1] + n * n_map[2]
        nn /= np.linalg.norm(nn)
        total, mask = 0.0, 0
        for i, L in enumerate(self.lights):
            lv = L - pos
            dist = np.linalg.norm(lv)
            ld = lv / dist
            ndl = float(nn @ ld)
            mult = 1.0
            if ndl > 0:
                mask |= 16 << i
                ts, ks, ms = self.hit(pos, ld, 0.001, dist)
                margin = min(margin, ms)
                if ks >= 0:
                    mult = 0.3
                    mask |= 1 << i
            total += min(max(ndl, 0.0), 1.0) * (10.0 / dist ** 2) * mult
        base = np.ones(3)
        value = base * total * attenuation + base * 0.1
        if metallic > 0.1:
            attenuation = attenuation * 0.5 * metallic
            value = value * (1 - attenuation)
            return value, 0, pos, d - 2 * (nn @ d) * nn, attenuation, mask, margin
        return value, 1, None, None, attenuation, mask, margin

    def pixel(self, u, W, Hh, x, y):
        ua = camera.uniforms_to_array(u).astype(np.float64)
        vi, pi = ua[:16].reshape(4, 4).T, ua[16:32].reshape(4, 4).T
        uv = (np.array([x + 0.5, y + 0.5]) / np.array([W, Hh])) * 2 - 1
        tgt = pi @ np.array([uv[0], uv[1], 1, 1])
        d = (vi @ np.append(tgt[:3] / np.linalg.norm(tgt[:3]), 0))[:3]
        o = (vi @ np.array([0, 0, 0, 1.0]))[:3]
        final, att, margin, mask0 = np.zeros(3), 1.0, np.inf, 0
        for depth in range(2):  # rgen:49-73
            value, done, o2, d2, att, mask, m = self.shade(o, d, att)
            final += value
            margin = min(margin, m)
            if depth == 0:
                mask0 = mask
                first = self.hit(o, d, 0.001, 1000.0)[1]
            if done == 1:
                break
            o, d = o2, d2
        return final, mask0, first, margin


@pytest.mark.parametrize("metallic,pose", [(False, ((0, 3, 9), (-0.2, 0, 0))), (False, ((2.5, 6, 3), (-1.1, 0.3, 0))),
                                           (True, ((0, 3, 9), (-0.35, 0, 0))), (True, ((-3, 2, 6), (-0.3, -0.4, 0)))])
def test_frames_match_closed_form_float64(metallic, pose):
    """Two-quad occluder (floor + quad above it + wall: hard shadows of the four lights) and mirror quad (metallic floor:
    one reflection bounce into wall, occluder underside or sky): primary quad, shadow mask and RGB of EVERY pixel of a
    48 x 36 frame against the closed-form float64 evaluation.  Pixels whose rays pass within 1e-4 of a quad edge are
    skipped (there float32 and float64 may disagree legitimately); at least 95 % of the frame must remain."""
    flat, o = _analytic(metallic)
    W, Hh = 48, 36
    u = camera.make_uniforms(W, Hh, position=pose[0], rotation=pose[1])
    r = o.render(u, W, Hh, mode=H.MODE_BRUTE)
    ref = _Quads64(metallic)
    checked = shadowed = reflected = 0
    for y in range(Hh):
        for x in range(W):
            final, mask, first, margin = ref.pixel(u, W, Hh, x, y)
            if margin < 1e-4:
                continue
            checked += 1
            tri = int(r["tri"][y, x])
            assert (tri == 0xFFFFFFFF and first < 0) or (tri != 0xFFFFFFFF and tri // 2 == first), (x, y, tri, first)
            assert int(r["mask"][y, x]) == mask, (x, y, int(r["mask"][y, x]), mask)
            assert np.allclose(r["rgb"][y, x], final, rtol=5e-5, atol=5e-6), (x, y, r["rgb"][y, x], final)
            want = np.floor(np.clip(final, 0, 1) * 255 + 0.5)
            assert (np.abs(r["bgra"][y, x][2::-1].astype(np.float64) - want) <= 1).all()
            shadowed += (mask & 15) != 0
            reflected += metallic and first == 0
    assert checked > 0.95 * W * Hh and shadowed > 20
    assert reflected > 100 or not metallic
