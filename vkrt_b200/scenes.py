"""Synthetic workloads (numpy only): analytic test scenes, random triangle soups, and the Sponza STAND-IN.

`models/sponza/Sponza.bin` (all geometry) is absent from the reference checkout, so the Sponza
configurations run on a deterministic stand-in that honours the real glTF's structure: 103 primitives with the
real per-primitive vertex and index counts (262,267 triangles, 192,496 vertices), u16-range local indices, the
real POSITION bounds of every primitive, the real material table and (when present) the real textures.  Every
number measured on it is labelled "stand-in geometry".

A "model" here is the Python mirror of the reference's `Model` (src/Model.hpp:8-54):
    {"submeshes": [{"vertices": (V,16) f32, "indices": (I,) u32 local, "material": int}],
     "materials": [(baseColor, metallicRoughness, normal)], "images": [(H,W,4) u8]}
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LAYOUT_PATH = os.path.join(_HERE, "data", "sponza_layout.json")
ASSET_DIRS = [
    os.path.join(_HERE, "..", "assets", "_ref", "sponza"),  # staged by __graft_entry__.build() (git-ignored, travels to the GPU box)
    "/root/reference/models/sponza",
]


# ---------------------------------------------------------------------------------------------------------------
# small helpers
# ---------------------------------------------------------------------------------------------------------------
def make_vertex_array(position, normal=None, uv=None, tangent=None) -> np.ndarray:
    """(V,16) float32 laid out as Model::Vertex {vec4 position, normal, uv, tangent} (src/Model.hpp:12-18)."""
    position = np.asarray(position, dtype=np.float32)
    n = position.shape[0]
    v = np.zeros((n, 16), dtype=np.float32)
    v[:, 0:3] = position
    if normal is not None:
        v[:, 4:7] = np.asarray(normal, dtype=np.float32)
    if uv is not None:
        v[:, 8:10] = np.asarray(uv, dtype=np.float32)
    if tangent is not None:
        t = np.asarray(tangent, dtype=np.float32)
        v[:, 12:12 + t.shape[1]] = t
    return v


def solid_image(rgba, size=4) -> np.ndarray:
    im = np.zeros((size, size, 4), dtype=np.uint8)
    im[:, :] = np.asarray(rgba, dtype=np.uint8)
    return im


def noise_image(size: int, seed: int, base=(128, 128, 128), amp=100, alpha=255) -> np.ndarray:
    """Deterministic smooth-ish RGBA8 noise (sum of a few octaves of bilinear value noise)."""
    rng = np.random.default_rng(seed)
    acc = np.zeros((size, size, 3), dtype=np.float64)
    for octave, cells in enumerate([4, 16, 64]):
        cells = min(cells, size)
        g = rng.random((cells, cells, 3))
        idx = (np.arange(size) * cells // size) % cells
        acc += g[np.ix_(idx, idx)] * (0.5 ** octave)
    acc = acc / acc.max()
    im = np.zeros((size, size, 4), dtype=np.uint8)
    im[..., :3] = np.clip(np.asarray(base)[None, None, :] + (acc - 0.5) * 2 * amp, 0, 255).astype(np.uint8)
    im[..., 3] = alpha
    return im


def identity_instance(scale: float = 1.0, translation=(0.0, 0.0, 0.0)) -> np.ndarray:
    m = np.zeros(12, dtype=np.float32)
    m[0] = m[5] = m[10] = scale
    m[3], m[7], m[11] = translation
    return m


REFERENCE_INSTANCE = identity_instance(0.01)  # src/Raytracer.cpp:1165-1169


# ---------------------------------------------------------------------------------------------------------------
# analytic scenes (hand-checkable)
# ---------------------------------------------------------------------------------------------------------------
def quad(p0, pu, pv, normal, uv_scale=1.0, tangent=None):
    """Two triangles p0, p0+pu, p0+pu+pv, p0+pv."""
    p0, pu, pv = (np.asarray(a, dtype=np.float32) for a in (p0, pu, pv))
    pos = np.stack([p0, p0 + pu, p0 + pu + pv, p0 + pv])
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=np.float32) * uv_scale
    nrm = np.tile(np.asarray(normal, dtype=np.float32), (4, 1))
    if tangent is None:
        tangent = pu / np.linalg.norm(pu)
    tan = np.tile(np.concatenate([np.asarray(tangent, dtype=np.float32), [1.0]]), (4, 1))
    return make_vertex_array(pos, nrm, uv, tan), np.array([0, 1, 2, 0, 2, 3], dtype=np.uint32)


def analytic_scene(metallic_floor: bool = False) -> Dict:
    """A floor quad (y=0), an occluder quad above part of it, and a back wall; instance scale 1.
    Textures: 0 = flat normal map (128,128,255), 1 = white base colour, 2 = non-metallic (blue=0), 3 = metallic."""
    floor_v, floor_i = quad((-8, 0, -8), (16, 0, 0), (0, 0, 16), (0, 1, 0), uv_scale=4.0)
    occ_v, occ_i = quad((-1, 2, -1), (2, 0, 0), (0, 0, 2), (0, 1, 0))
    wall_v, wall_i = quad((-8, 0, -8), (16, 0, 0), (0, 8, 0), (0, 0, 1))
    images = [solid_image((128, 128, 255, 255)), solid_image((255, 255, 255, 255)), solid_image((0, 255, 0, 255)),
              solid_image((0, 128, 255, 255))]
    materials = [(1, 3 if metallic_floor else 2, 0), (1, 2, 0), (1, 2, 0)]
    submeshes = [
        {"vertices": floor_v, "indices": floor_i, "material": 0},
        {"vertices": occ_v, "indices": occ_i, "material": 1},
        {"vertices": wall_v, "indices": wall_i, "material": 2},
    ]
    return {"submeshes": submeshes, "materials": materials, "images": images}


def random_soup(triangles: int, seed: int = 1, extent=(10.0, 6.0, 8.0), size: float = 0.6, geometries: int = 3, images: int = 4) -> Dict:
    """Random small triangles in a box, split over several geometries, noise textures, random metallic."""
    rng = np.random.default_rng(seed)
    counts = np.full(geometries, triangles // geometries)
    counts[: triangles % geometries] += 1
    imgs = [solid_image((128, 128, 255, 255), 4)]
    for k in range(1, images):
        imgs.append(noise_image(32, seed * 131 + k, base=(150, 140, 130), amp=90))
    materials = [(1 + (g % max(1, images - 1)), (g * 7 + 2) % images, 0 if g % 2 == 0 else (g % images)) for g in range(geometries)]
    submeshes = []
    for g in range(geometries):
        t = int(counts[g])
        c = (rng.random((t, 1, 3)) - 0.5) * np.asarray(extent) * 2
        p = c + rng.normal(0, size, (t, 3, 3))
        pos = p.reshape(-1, 3).astype(np.float32)
        e1, e2 = p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]
        n = np.cross(e1, e2)
        n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-12)
        tng = e1 / np.maximum(np.linalg.norm(e1, axis=1, keepdims=True), 1e-12)
        nrm = np.repeat(n, 3, axis=0)
        tan = np.concatenate([np.repeat(tng, 3, axis=0), np.ones((3 * t, 1))], axis=1)
        uv = rng.random((3 * t, 2)) * 3 - 1
        submeshes.append({"vertices": make_vertex_array(pos, nrm, uv, tan), "indices": np.arange(3 * t, dtype=np.uint32), "material": g})
    return {"submeshes": submeshes, "materials": materials, "images": imgs}


def procedural_soup(triangles: int, seed: int = 0x5EED, frame: int = 0) -> Dict:
    """BASELINE.json config 5: one geometry, centroids uniform in [-20,20]x[0,15]x[-12,12], edge vectors
    ~N(0, 0.02^2) per axis, white texture, material row {0,0,0,0}; frame f moves centroid.y by 0.05*sin(0.1 f + hash)."""
    rng = np.random.default_rng(seed)
    c = rng.random((triangles, 3), dtype=np.float32) * np.array([40, 15, 24], dtype=np.float32) + np.array([-20, 0, -12], dtype=np.float32)
    e = rng.normal(0, 0.02, (triangles, 2, 3)).astype(np.float32)
    phase = (rng.random(triangles, dtype=np.float32) * np.float32(2 * np.pi))
    c[:, 1] += np.float32(0.05) * np.sin(np.float32(0.1 * frame) + phase)
    pos = np.empty((triangles, 3, 3), dtype=np.float32)
    pos[:, 0] = c
    pos[:, 1] = c + e[:, 0]
    pos[:, 2] = c + e[:, 1]
    n = np.cross(e[:, 0], e[:, 1])
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    v = np.zeros((triangles * 3, 16), dtype=np.float32)
    v[:, 0:3] = pos.reshape(-1, 3)
    v[:, 4:7] = np.repeat(n, 3, axis=0)
    t = e[:, 0] / np.maximum(np.linalg.norm(e[:, 0], axis=1, keepdims=True), 1e-20)
    v[:, 12:15] = np.repeat(t, 3, axis=0)
    v[:, 15] = 1
    sub = {"vertices": v, "indices": np.arange(3 * triangles, dtype=np.uint32), "material": 0}
    return {"submeshes": [sub], "materials": [(0, 0, 0)], "images": [solid_image((255, 255, 255, 255), 4)]}


# ---------------------------------------------------------------------------------------------------------------
# Sponza stand-in
# ---------------------------------------------------------------------------------------------------------------
def _spow(x, e):
    return np.sign(x) * np.abs(x) ** e


def _element(center, half, tris, vert_budget, inward, open_top, uv_tiles=4.0, dims=None):
    """A superellipsoid ("rounded box") shell with exactly `tris` triangles and at most `vert_budget` vertices.
    Returns (vertices (n,16), indices (3*tris,), (nu, nv)).  `dims` forces the grid (primitives that share one
    index accessor in the glTF must produce identical indices)."""
    tris = int(tris)
    center, half = np.asarray(center, dtype=np.float64), np.maximum(np.asarray(half, dtype=np.float64), 1e-3)
    if dims is None:
        perimeter = 4.0 * (half[0] + half[2])
        height = 2.0 * half[1] + (half[0] + half[2])
        quads = max(1, (tris + 1) // 2)
        ratio = float(np.clip(perimeter / max(height, 1e-6), 0.25, 4.0))
        nu_min = 3 if vert_budget >= 8 else 2
        nu = max(nu_min, int(round(np.sqrt(quads * ratio))))
        nv = max(1, int(np.ceil(quads / nu)))
        while (nu + 1) * (nv + 1) > vert_budget and (nu > nu_min or nv > 1):
            if nu > nu_min and (nu >= nv or nv == 1):
                nu -= 1
            else:
                nv -= 1
    else:
        nu, nv = dims
    e = 0.35  # squareness of the superellipsoid
    u = np.linspace(-np.pi, np.pi, nu + 1) + (np.pi / nu if nu < 8 and nu % 2 else 0.0)  # odd coarse rings: rotate so both axes get extent
    v_cap = 0.5 * np.pi - 0.05 if nv >= 4 else (np.pi / 3 if nv >= 2 else np.pi / 4)  # coarse grids stay off the poles
    v_hi = 0.12 * np.pi if open_top else v_cap
    v = np.linspace(-(0.5 * np.pi - 0.05) if open_top else -v_cap, v_hi, nv + 1)
    uu, vv = np.meshgrid(u, v)  # (nv+1, nu+1)
    cu, su, cv, sv = np.cos(uu), np.sin(uu), np.cos(vv), np.sin(vv)
    px = half[0] * _spow(cv, e) * _spow(cu, e)
    py = half[1] * _spow(sv, e)
    pz = half[2] * _spow(cv, e) * _spow(su, e)
    if open_top:
        py = py / max(np.max(py), 1e-9) * half[1]  # stretch the cut-off shell back to the full height
    nx = _spow(cv, 2 - e) * _spow(cu, 2 - e) / half[0]
    ny = _spow(sv, 2 - e) / half[1]
    nz = _spow(cv, 2 - e) * _spow(su, 2 - e) / half[2]
    nrm = np.stack([nx, ny, nz], axis=-1)
    nrm /= np.maximum(np.linalg.norm(nrm, axis=-1, keepdims=True), 1e-12)
    tan = np.stack([-su, np.zeros_like(su), cu], axis=-1)  # d/du direction, made orthogonal to the normal below
    tan = tan - nrm * np.sum(tan * nrm, axis=-1, keepdims=True)
    tan /= np.maximum(np.linalg.norm(tan, axis=-1, keepdims=True), 1e-12)
    if inward:
        nrm = -nrm
    pos = np.stack([px, py, pz], axis=-1) + center
    uv = np.stack([(uu + np.pi) / (2 * np.pi) * uv_tiles, (vv + 0.5 * np.pi) / np.pi * uv_tiles], axis=-1)
    verts = make_vertex_array(pos.reshape(-1, 3), nrm.reshape(-1, 3), uv.reshape(-1, 2),
                              np.concatenate([tan.reshape(-1, 3), np.ones(((nu + 1) * (nv + 1), 1))], axis=1))
    r, c = np.meshgrid(np.arange(nv), np.arange(nu), indexing="ij")
    a = (r * (nu + 1) + c).reshape(-1)
    b, d, cidx = a + 1, a + (nu + 1), a + (nu + 1) + 1
    if inward:
        quads_idx = np.stack([a, d, b, b, d, cidx], axis=1)
    else:
        quads_idx = np.stack([a, b, d, b, cidx, d], axis=1)
    tri_idx = quads_idx.reshape(-1, 3)
    if tri_idx.shape[0] >= tris:
        tri_idx = tri_idx[:tris]
    else:  # vertex budget too small for a full grid: repeat triangles with rotated vertex order (exact duplicates)
        reps = int(np.ceil(tris / tri_idx.shape[0]))
        tri_idx = np.concatenate([np.roll(tri_idx, k, axis=1) for k in range(reps)], axis=0)[:tris]
    return verts, tri_idx.reshape(-1).astype(np.uint32), (nu, nv)


# Free space kept clear of stand-in geometry (object units = world / 0.01): the nave of the atrium, which holds the
# default camera (630,450,-70), the config-2 camera path along -x and the four lights.
NAVE_LO = np.array([-1250.0, 140.0, -150.0])
NAVE_HI = np.array([1150.0, 1.0e9, 150.0])  # open to the sky


def _clip_out_of_nave(centre, half):
    """Returns the parts of the element box that lie outside the nave (split along z): 0, 1 or 2 boxes."""
    lo, hi = centre - half, centre + half
    if np.any(hi <= NAVE_LO) or np.any(lo >= NAVE_HI):
        return [(centre, half)]
    parts = []
    for zlo, zhi in ((lo[2], min(hi[2], NAVE_LO[2])), (max(lo[2], NAVE_HI[2]), hi[2])):
        if zhi - zlo >= 6.0:
            l, h = lo.copy(), hi.copy()
            l[2], h[2] = zlo, zhi
            parts.append((0.5 * (l + h), 0.5 * (h - l)))
    return parts


def load_layout() -> Dict:
    return json.load(open(LAYOUT_PATH))


def sponza_standin_submeshes(layout: Optional[Dict] = None) -> List[Dict]:
    layout = layout or load_layout()
    prims = layout["primitives"]
    lo_all = np.min([p["min"] for p in prims], axis=0)
    hi_all = np.max([p["max"] for p in prims], axis=0)
    scene_vol = float(np.prod(hi_all - lo_all))
    out = []
    topo_cache = {}  # index accessor -> [(nu, nv, tris) per element]: shared accessors must give identical indices
    for pi, p in enumerate(prims):
        V, T = int(p["vertexCount"]), int(p["indexCount"]) // 3
        lo, hi = np.asarray(p["min"], dtype=np.float64), np.asarray(p["max"], dtype=np.float64)
        ext = np.maximum(hi - lo, 1e-3)
        shell = float(np.prod(ext)) / scene_vol > 0.2
        # candidate element boxes: one box for object-sized primitives, a lattice over x/z for large ones
        if shell or max(ext[0], ext[2]) < 420.0:
            gx = gz = 1
        else:
            gx, gz = int(np.ceil(ext[0] / 220.0)), int(np.ceil(ext[2] / 220.0))
        boxes = []
        while True:
            boxes = []
            cell = ext / np.array([gx, 1, gz])
            for k in range(gx * gz):
                ix, iz = k % gx, k // gx
                fill = 1.0 if ext[1] < 0.15 * max(ext[0], ext[2]) else 0.62  # slab-like primitives tile without gaps
                half = 0.5 * cell * np.array([1.0 if gx == 1 else fill, 1.0, 1.0 if gz == 1 else fill])
                centre = lo + cell * (np.array([ix, 0, iz]) + 0.5)
                if gx * gz > 1:  # outermost elements touch the primitive's bounds so min/max are attained
                    if ix == 0:
                        centre[0] = lo[0] + half[0]
                    if ix == gx - 1:
                        centre[0] = hi[0] - half[0]
                    if iz == 0:
                        centre[2] = lo[2] + half[2]
                    if iz == gz - 1:
                        centre[2] = hi[2] - half[2]
                if shell or (gx * gz == 1 and max(ext[0], ext[2]) < 420.0):
                    boxes.append((centre, half))  # shells enclose the nave; object-sized primitives sit at their real place
                else:
                    boxes.extend(_clip_out_of_nave(centre, half))
            if not boxes:  # everything fell into the nave: keep one thin box on the primitive's low-z face
                boxes = [(np.array([0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), lo[2] + 3.0]), np.array([0.5 * ext[0], 0.5 * ext[1], 3.0]))]
            K = len(boxes)
            fits = T // K >= 16 and (T // 2 + K * (2 * int(np.sqrt(T / K)) + 4)) <= V
            if K == 1 or fits or (gx == 1 and gz == 1):
                break
            if gx >= gz and gx > 1:
                gx -= 1
            else:
                gz -= 1
        cached = topo_cache.get(p.get("indexAccessor", -1 - pi))
        if cached is not None and len(cached) != len(boxes):
            # same index data, different element count: reuse the first primitive's topology on the first boxes
            boxes = (boxes * len(cached))[: len(cached)]
        K = len(boxes)
        dims_list, verts_list, idx_list = [], [], []
        used_v, tri_left, vert_left = 0, T, V
        for k, (centre, half) in enumerate(boxes):
            if cached is not None:
                nu, nv, t_k = cached[k]
                dims = (nu, nv)
            else:
                t_k, dims = tri_left // (K - k), None
            v_k = vert_left // (K - k)
            if t_k == 0:
                dims_list.append((2, 1, 0))
                continue
            ev, ei, dims = _element(centre, half, t_k, max(v_k, 3), inward=shell, open_top=shell, dims=dims)
            dims_list.append((dims[0], dims[1], t_k))
            idx_list.append(ei + used_v)
            verts_list.append(ev)
            used_v += ev.shape[0]
            tri_left -= t_k
            vert_left -= ev.shape[0]
        if cached is None:
            topo_cache[p.get("indexAccessor", -1 - pi)] = dims_list
        verts = np.concatenate(verts_list, axis=0)
        idx = np.concatenate(idx_list, axis=0)
        assert idx.size == 3 * T, (pi, idx.size, T)
        # glTF front faces are counter-clockwise seen from the side the normal points to, and so are the real Sponza's;
        # _element emits the opposite order.  The ray tracer does not care (no culling, src/Raytracer.cpp:1176), the
        # rasteriser culls back faces (src/Rasterizer.cpp:803-804) and would show the insides of the elements.
        idx = idx.reshape(-1, 3)[:, [0, 2, 1]].reshape(-1)
        assert verts.shape[0] <= V, (pi, verts.shape[0], V)
        if verts.shape[0] < V:  # pad with unreferenced copies of vertex 0 to the accessor's count
            verts = np.concatenate([verts, np.repeat(verts[:1], V - verts.shape[0], axis=0)], axis=0)
        # clamp into the accessor bounds and snap the extremes onto them (POSITION min/max are honoured exactly)
        lo32, hi32 = np.asarray(p["min"], dtype=np.float32), np.asarray(p["max"], dtype=np.float32)
        verts[:, 0:3] = np.clip(verts[:, 0:3], lo32, hi32)
        for ax in range(3):
            cur_lo, cur_hi = float(verts[:, ax].min()), float(verts[:, ax].max())
            want = float(hi32[ax]) - float(lo32[ax])
            if cur_hi > cur_lo and want > 0 and want / (cur_hi - cur_lo) <= 1.6:
                # coarse elements do not reach the box faces: stretch the primitive onto its accessor bounds
                verts[:, ax] = (float(lo32[ax]) + (verts[:, ax].astype(np.float64) - cur_lo) * (want / (cur_hi - cur_lo))).astype(np.float32)
                verts[:, ax] = np.clip(verts[:, ax], lo32[ax], hi32[ax])
                verts[np.argmin(verts[:, ax]), ax] = lo32[ax]
                verts[np.argmax(verts[:, ax]), ax] = hi32[ax]
            elif want == 0:
                verts[:, ax] = lo32[ax]
        if not p["hasTangent"]:
            verts[:, 12:16] = 0.0  # primitive 50: no TANGENT accessor -> the loader leaves zeros (src/Model.hpp:17)
        assert int(idx.max()) < 65536
        out.append({"vertices": verts.astype(np.float32), "indices": idx.astype(np.uint32), "material": int(p["material"])})
    # primitives that share an attribute accessor in the glTF (instanced copies share NORMAL / TANGENT / TEXCOORD_0 and
    # index data, each with its own POSITION) must carry identical data for it: the first user's data wins
    columns = {"POSITION": (0, 3), "NORMAL": (4, 7), "TEXCOORD_0": (8, 10), "TANGENT": (12, 16)}
    first_user = {}
    for s, p in zip(out, prims):
        for name, acc_id in p.get("attributes", {}).items():
            if name not in columns:
                continue
            c0, c1 = columns[name]
            if acc_id in first_user:
                s["vertices"][:, c0:c1] = first_user[acc_id]
            else:
                first_user[acc_id] = s["vertices"][:, c0:c1].copy()
    return out


def find_asset_dir() -> Optional[str]:
    for d in ASSET_DIRS:
        if os.path.isfile(os.path.join(d, "white.png")) or os.path.isfile(os.path.join(d, "Sponza.gltf")):
            return os.path.abspath(d)
    return None


def load_sponza_images(layout: Optional[Dict] = None, max_size: Optional[int] = None):
    """Returns (images, provenance).  Real textures (decoded by the host C++ decoder, forced to RGBA8 as tinygltf does,
    submodules/tinygltf/tiny_gltf.h:2363) when the files are staged; deterministic procedural textures otherwise."""
    layout = layout or load_layout()
    d = find_asset_dir()
    images = []
    if d is not None and all(os.path.isfile(os.path.join(d, im["uri"])) for im in layout["images"]):
        from .model import decode_image  # host C++ decoder: bit-identical to the reference's stb_image on these files

        def one(im):
            with open(os.path.join(d, im["uri"]), "rb") as f:
                rgba = decode_image(f.read())
            if max_size and max(rgba.shape[:2]) > max_size:
                step = max(rgba.shape[:2]) // max_size
                rgba = np.ascontiguousarray(rgba[::step, ::step])
            return rgba

        # the reference decodes the 69 images one after the other (stb_image inside tinygltf, src/Model.cpp:138-151); the
        # decoder is called through ctypes, which releases the GIL, so a thread pool decodes them on all host cores
        from concurrent.futures import ThreadPoolExecutor

        with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as pool:
            images = list(pool.map(one, layout["images"]))
        return images, "reference textures (host decoder, == stb_image)" + (f", subsampled to {max_size}" if max_size else "")
    # procedural fall-back: role-dependent content so normal maps stay normal-like and metallic maps binary-ish
    roles = {}
    for m in layout["materials"]:
        roles.setdefault(m["baseColor"], "base")
        roles.setdefault(m["normal"], "normal")
        roles.setdefault(m["metallicRoughness"], "mr")
    for i, im in enumerate(layout["images"]):
        size = min(im["width"], max_size or im["width"])
        role = roles.get(i, "base")
        if im["width"] <= 4:
            images.append(solid_image((255, 255, 255, 255), im["width"]))
        elif role == "normal":
            images.append(noise_image(size, 7000 + i, base=(128, 128, 235), amp=20))
        elif role == "mr":
            n = noise_image(size, 9000 + i, base=(0, 160, 40), amp=60)
            if i % 3 == 0:
                n[..., 2] = np.where(n[..., 2] > 40, 200, 0)
            images.append(n)
        else:
            images.append(noise_image(size, 5000 + i, base=(150, 130, 110), amp=80))
    return images, "procedural textures (reference texture files not staged)"


def sponza_standin(max_texture_size: Optional[int] = None, with_images: bool = True) -> Dict:
    """with_images=False: geometry and materials only (a rank that receives the textures from another rank,
    Raytracer(textures_from_root=...))."""
    layout = load_layout()
    if not with_images:
        materials = [(m["baseColor"], m["metallicRoughness"], m["normal"]) for m in layout["materials"]]
        return {"submeshes": sponza_standin_submeshes(layout), "materials": materials, "images": None,
                "provenance": "stand-in geometry (Sponza.bin absent from the reference checkout) + textures received from the root rank"}
    images, provenance = load_sponza_images(layout, max_texture_size)
    materials = [(m["baseColor"], m["metallicRoughness"], m["normal"]) for m in layout["materials"]]
    return {"submeshes": sponza_standin_submeshes(layout), "materials": materials, "images": images,
            "provenance": "stand-in geometry (Sponza.bin absent from the reference checkout) + " + provenance}


def write_sponza_bin(path: str, gltf_path: str):
    """Writes a glTF-conformant stand-in Sponza.bin: every attribute at its bufferView's byteOffset with its
    byteStride, indices as uint16 (so the reference's own Model.cpp loads it, src/Model.cpp:48-120)."""
    g = json.load(open(gltf_path))
    subs = sponza_standin_submeshes()
    buf = bytearray(g["buffers"][0]["byteLength"])
    acc, bvs = g["accessors"], g["bufferViews"]
    for p, s in zip(g["meshes"][0]["primitives"], subs):
        ia = acc[p["indices"]]
        bv = bvs[ia["bufferView"]]
        off = bv.get("byteOffset", 0) + ia.get("byteOffset", 0)
        data = s["indices"].astype("<u2").tobytes()
        buf[off:off + len(data)] = data
        for name, col, ncomp in (("POSITION", 0, 3), ("NORMAL", 4, 3), ("TEXCOORD_0", 8, 2), ("TANGENT", 12, 4)):
            if name not in p["attributes"]:
                continue
            a = acc[p["attributes"][name]]
            bv = bvs[a["bufferView"]]
            off = bv.get("byteOffset", 0) + a.get("byteOffset", 0)
            stride = bv.get("byteStride", 4 * ncomp)
            arr = s["vertices"][: a["count"], col:col + ncomp].astype("<f4")
            view = np.frombuffer(buf, dtype=np.uint8, count=stride * a["count"], offset=off).reshape(a["count"], stride)
            view[:, : 4 * ncomp] = arr.view(np.uint8).reshape(a["count"], 4 * ncomp)
    with open(path, "wb") as f:
        f.write(bytes(buf))


def random_instance_transforms(n: int, seed: int = 1, extent=(60.0, 20.0, 60.0), scale=(0.2, 1.0)) -> np.ndarray:
    """n row-major 3x4 object->world matrices: random rotation, uniform scale in `scale`, translation inside +-extent/2
    (top-level tests: overlapping, rotated instance boxes)."""
    rng = np.random.default_rng(seed)
    q = rng.normal(0, 1, (n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    rot = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                    2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                    2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1).reshape(n, 3, 3)
    sc = rng.uniform(scale[0], scale[1], n)
    tr = (rng.random((n, 3)) - 0.5) * np.asarray(extent)
    m = np.zeros((n, 3, 4), dtype=np.float32)
    m[:, :, :3] = rot * sc[:, None, None]
    m[:, :, 3] = tr
    return m.reshape(n, 12)


def instanced_grid_transforms(nx: int = 4, nz: int = 4) -> np.ndarray:
    """BASELINE.json config 4: scale 0.01 + translation (40 i, 0, 25 j)."""
    out = []
    for j in range(nz):
        for i in range(nx):
            out.append(identity_instance(0.01, (40.0 * i, 0.0, 25.0 * j)))
    return np.stack(out)
