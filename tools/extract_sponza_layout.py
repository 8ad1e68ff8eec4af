#!/usr/bin/env python
"""Extracts the facts of models/sponza/Sponza.gltf that the stand-in scene generator needs:
per-primitive vertex/index counts, POSITION bounds, material; per-material image indices; per-image uri and size.
Run once in the build container (needs /root/reference); the result is committed as vkrt_b200/data/sponza_layout.json
so the generator also works where the reference checkout is absent (the GPU box)."""
import json
import os
import sys

from PIL import Image

REF = os.environ.get("REF", "/root/reference")
src = os.path.join(REF, "models", "sponza", "Sponza.gltf")
g = json.load(open(src))
acc = g["accessors"]
tex = g["textures"]


def source(t):
    return -1 if t is None else tex[t["index"]]["source"]


prims = []
for p in g["meshes"][0]["primitives"]:
    a = p["attributes"]
    pos = acc[a["POSITION"]]
    prims.append({
        "vertexCount": pos["count"],
        "indexCount": acc[p["indices"]]["count"],
        "indexAccessor": p["indices"],
        "min": pos["min"],
        "max": pos["max"],
        "material": p["material"],
        "hasTangent": "TANGENT" in a,
        "attributes": dict(a),
    })
mats = []
for m in g["materials"]:
    pbr = m.get("pbrMetallicRoughness", {})
    mats.append({
        "baseColor": source(pbr.get("baseColorTexture")),
        "metallicRoughness": source(pbr.get("metallicRoughnessTexture")),
        "normal": source(m.get("normalTexture")),
    })
images = []
for im in g["images"]:
    with Image.open(os.path.join(REF, "models", "sponza", im["uri"])) as f:
        images.append({"uri": im["uri"], "width": f.width, "height": f.height})
out = {"source": "models/sponza/Sponza.gltf (glTF-Transform v1.2.3)", "primitives": prims, "materials": mats, "images": images,
       "bufferByteLength": g["buffers"][0]["byteLength"]}
dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "vkrt_b200", "data", "sponza_layout.json")
json.dump(out, open(dst, "w"), indent=0)
print("wrote", dst, len(prims), "primitives", len(mats), "materials", len(images), "images",
      sum(p["indexCount"] for p in prims) // 3, "triangles", sum(p["vertexCount"] for p in prims), "vertices")
