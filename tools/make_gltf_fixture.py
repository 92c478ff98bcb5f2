#!/usr/bin/env python
"""Writes tests/golden/aux/textured.gltf: a self-contained glTF 2.0 file (buffers and the texture as data URIs) that exercises
what the reference's loader drops (src/gltf_loader.h:65, :312; src/triangle.h:35-39): TEXCOORD_0, per-primitive materials
(base-colour texture, metallic, emissive, transmission), u8 / u16 / u32 indices and a node transform."""
import base64
import io
import json
import os

import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "aux", "textured.gltf")


def main():
    rng = np.random.RandomState(5)
    blobs, views, accessors = [], [], []

    def add(arr, target=None):
        raw = np.ascontiguousarray(arr).tobytes()
        off = sum(len(b) for b in blobs)
        pad = (-len(raw)) % 4
        blobs.append(raw + b"\0" * pad)
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(raw)})
        comp = {np.dtype("float32"): 5126, np.dtype("uint16"): 5123, np.dtype("uint8"): 5121, np.dtype("uint32"): 5125}[arr.dtype]
        typ = {1: "SCALAR", 2: "VEC2", 3: "VEC3"}[1 if arr.ndim == 1 else arr.shape[1]]
        acc = {"bufferView": len(views) - 1, "componentType": comp, "count": len(arr), "type": typ}
        if typ == "VEC3":
            acc["min"], acc["max"] = arr.min(axis=0).tolist(), arr.max(axis=0).tolist()
        accessors.append(acc)
        return len(accessors) - 1

    def grid(n, z, x0):
        """An n x n grid of quads in the plane z, 2 x 2 units starting at x0, with uv over [0, 1]^2."""
        k = np.linspace(0, 1, n + 1, dtype=np.float32)
        gx, gy = np.meshgrid(k, k)
        pos = np.stack([x0 + 2 * gx.ravel(), 2 * gy.ravel(), np.full(gx.size, z, np.float32)], axis=1).astype(np.float32)
        uv = np.stack([gx.ravel(), 1 - gy.ravel()], axis=1).astype(np.float32)
        idx = []
        for j in range(n):
            for i in range(n):
                a = j * (n + 1) + i
                idx += [a, a + 1, a + n + 2, a, a + n + 2, a + n + 1]
        return pos, uv, np.array(idx)

    prims = []
    for m, (n, z, x0, idt) in enumerate([(4, 0.0, -3.0, np.uint16), (2, 0.2, -0.8, np.uint8), (1, 0.4, 1.4, np.uint32), (3, -0.5, -1.0, np.uint16)]):
        pos, uv, idx = grid(n, z, x0)
        prims.append({"attributes": {"POSITION": add(pos), "TEXCOORD_0": add(uv)}, "indices": add(idx.astype(idt)), "material": m})
    tex = (rng.rand(8, 8, 3) * 255).astype(np.uint8)
    tex[0, 0] = [255, 0, 0]
    tex[7, 7] = [0, 0, 255]
    png = io.BytesIO()
    Image.fromarray(tex, "RGB").save(png, format="PNG")
    j = {
        "asset": {"version": "2.0", "generator": "tools/make_gltf_fixture.py"},
        "scene": 0,
        "scenes": [{"nodes": [0]}],
        "nodes": [{"mesh": 0, "translation": [0.5, 1.0, -2.0], "scale": [1.5, 1.5, 1.5]}],
        "meshes": [{"primitives": prims}],
        "materials": [
            {"name": "textured", "pbrMetallicRoughness": {"baseColorTexture": {"index": 0}, "metallicFactor": 0.0}},
            {"name": "metal", "pbrMetallicRoughness": {"baseColorFactor": [0.9, 0.7, 0.3, 1.0], "metallicFactor": 1.0, "roughnessFactor": 0.25}},
            {"name": "lamp", "emissiveFactor": [1.0, 0.9, 0.8], "extensions": {"KHR_materials_emissive_strength": {"emissiveStrength": 6.0}}},
            {"name": "glass", "pbrMetallicRoughness": {"metallicFactor": 0.0},
             "extensions": {"KHR_materials_transmission": {"transmissionFactor": 1.0}, "KHR_materials_ior": {"ior": 1.45}}},
        ],
        "textures": [{"source": 0}],
        "images": [{"uri": "data:image/png;base64," + base64.b64encode(png.getvalue()).decode()}],
        "accessors": accessors,
        "bufferViews": views,
    }
    blob = b"".join(blobs)
    j["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        json.dump(j, f, indent=1)
    print(OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
