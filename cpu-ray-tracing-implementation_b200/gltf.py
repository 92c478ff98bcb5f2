"""glTF 2.0 -> triangle rows for b200rt: the host-side data path in front of the C ABI.

Replaces, on the host side of the boundary, `gltf::GltfLoader` (src/gltf_loader.h) plus the copy loops of the scene functions
(`glass_fox`, src/main.cc:346-393; `sponza`, src/main.cc:445-485): JSON + buffers -> float32 positions -> (n, 9) triangle rows,
ready for `scene.triangle_scene`.

mode="reference" reproduces what the reference does, so that the same file gives the same triangles in the same order:
  every primitive of every mesh in file order, POSITION as float32 VEC3, indices only if they are UNSIGNED_SHORT
  (src/main.cc:367-375 reads nothing else; a primitive with other index types yields no triangles there), NO node transforms
  (src/gltf_loader.h:295-303 never applies them: Sponza renders at its raw, un-scaled coordinates).
mode="full" removes those gaps: u8 / u16 / u32 indices, the default scene's node hierarchy with matrix / TRS transforms,
  a mesh instanced by several nodes emitted once per node.
"""
import base64
import json
import os
import struct

import numpy as np

from . import _abi as abi

_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_WIDTH = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


class Gltf:
    def __init__(self, path):
        self.dir = os.path.dirname(os.path.abspath(path))
        with open(path, "rb") as f:
            head = f.read(12)
            f.seek(0)
            raw = f.read()
        self.bin_chunk = None
        if head[:4] == b"glTF":  # binary container: 12-byte header, then JSON chunk, then optional BIN chunk
            off = 12
            chunks = []
            while off < len(raw):
                n, kind = struct.unpack_from("<II", raw, off)
                chunks.append((kind, raw[off + 8: off + 8 + n]))
                off += 8 + n
            self.j = json.loads(chunks[0][1].decode("utf-8"))
            self.bin_chunk = next((c for k, c in chunks if k == 0x004E4942), None)
        else:
            self.j = json.loads(raw.decode("utf-8"))
        self._buffers = {}

    def buffer(self, i):
        if i not in self._buffers:
            b = self.j["buffers"][i]
            uri = b.get("uri")
            if uri is None:
                data = self.bin_chunk
            elif uri.startswith("data:"):
                data = base64.b64decode(uri.split(",", 1)[1])
            else:
                with open(os.path.join(self.dir, uri), "rb") as f:
                    data = f.read()
            if data is None or len(data) < b["byteLength"]:
                raise ValueError(f"glTF buffer {i}: {0 if data is None else len(data)} bytes, {b['byteLength']} declared")
            self._buffers[i] = data
        return self._buffers[i]

    def accessor(self, i):
        a = self.j["accessors"][i]
        if "sparse" in a:
            raise ValueError("sparse accessors are not supported")
        dt, width, count = np.dtype(_COMPONENT[a["componentType"]]), _WIDTH[a["type"]], a["count"]
        if "bufferView" not in a:
            return np.zeros((count, width), dt)
        v = self.j["bufferViews"][a["bufferView"]]
        data = self.buffer(v["buffer"])
        start = v.get("byteOffset", 0) + a.get("byteOffset", 0)
        stride = v.get("byteStride") or dt.itemsize * width
        out = np.ndarray((count, width), dt, data, start, (stride, dt.itemsize))
        return np.ascontiguousarray(out)

    # ---- node transforms (mode="full")
    @staticmethod
    def _local(node):
        if "matrix" in node:
            return np.array(node["matrix"], np.float64).reshape(4, 4).T  # glTF stores column-major
        m = np.eye(4)
        if "scale" in node:
            m = np.diag(list(node["scale"]) + [1.0]) @ m
        if "rotation" in node:
            x, y, z, w = node["rotation"]
            r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                          [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                          [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
            rr = np.eye(4)
            rr[:3, :3] = r
            m = rr @ m
        if "translation" in node:
            t = np.eye(4)
            t[:3, 3] = node["translation"]
            m = t @ m
        return m

    def mesh_instances(self):
        """(mesh index, world matrix) for every node of the default scene that carries a mesh, depth first."""
        j = self.j
        scene = j.get("scenes", [{}])[j.get("scene", 0)] if j.get("scenes") else {"nodes": list(range(len(j.get("nodes", []))))}
        out, stack = [], [(n, np.eye(4)) for n in reversed(scene.get("nodes", []))]
        while stack:
            n, parent = stack.pop()
            node = j["nodes"][n]
            world = parent @ self._local(node)
            if "mesh" in node:
                out.append((node["mesh"], world))
            stack.extend((c, world) for c in reversed(node.get("children", [])))
        return out


def _primitive_triangles(g, prim, mode):
    if prim.get("mode", 4) != 4 or "POSITION" not in prim.get("attributes", {}):
        return np.zeros((0, 3, 3), np.float32)
    pos = g.accessor(prim["attributes"]["POSITION"])
    if pos.dtype != np.float32 or pos.shape[1] != 3:
        return np.zeros((0, 3, 3), np.float32)  # src/main.cc:358 reads DataType::kFloat only
    if "indices" in prim:
        idx = g.accessor(prim["indices"]).reshape(-1)
        if mode == "reference" and idx.dtype != np.uint16:
            return np.zeros((0, 3, 3), np.float32)  # src/main.cc:368: only kUnsignedShort indices are read
        idx = idx[: len(idx) // 3 * 3].astype(np.int64)
        return pos[idx].reshape(-1, 3, 3)
    return pos[: len(pos) // 3 * 3].reshape(-1, 3, 3)


def load_triangles(path, mode="reference"):
    """-> float32 array (n, 9): p0 p1 p2 of every triangle, in the order the reference's loops would push them."""
    if mode not in ("reference", "full"):
        raise ValueError("mode must be 'reference' or 'full'")
    g = Gltf(path)
    parts = []
    if mode == "reference":
        for mesh in g.j.get("meshes", []):
            for prim in mesh.get("primitives", []):
                parts.append(_primitive_triangles(g, prim, mode))
    else:
        for mesh_i, world in g.mesh_instances():
            for prim in g.j["meshes"][mesh_i].get("primitives", []):
                t = _primitive_triangles(g, prim, mode).astype(np.float64)
                t = t @ world[:3, :3].T + world[:3, 3]
                parts.append(t.astype(np.float32))
    if not parts:
        return np.zeros((0, 9), np.float32)
    return np.ascontiguousarray(np.concatenate(parts).reshape(-1, 9), dtype=np.float32)


# ------------------------------------------------------------------------------------------------ materials, textures, uv
# (second half: "per-primitive materials/textures, uv on triangles - all currently missing":
#  src/gltf_loader.h:65 `// Material material;`, :312 `// loadMaterials();`, src/triangle.h:35-39 never writes u, v.)
def _primitive_uv(g, prim, n_tris):
    acc = prim.get("attributes", {}).get("TEXCOORD_0")
    if acc is None:
        return None
    uv = g.accessor(acc)
    if uv.dtype == np.uint8:
        uv = uv.astype(np.float64) / 255.0
    elif uv.dtype == np.uint16:
        uv = uv.astype(np.float64) / 65535.0
    else:
        uv = uv.astype(np.float64)
    if "indices" in prim:
        idx = g.accessor(prim["indices"]).reshape(-1)
        idx = idx[: len(idx) // 3 * 3].astype(np.int64)
        uv = uv[idx]
    uv = uv[: n_tris * 3].reshape(n_tris, 3, 2).copy()
    uv[..., 1] = 1.0 - uv[..., 1]  # glTF's v runs down the image, picture_texture's j = H * (1 - v) runs up (src/texture.h:69)
    return uv.reshape(n_tris, 6)


def _decode_image(g, image):
    """-> uint8 (H, W, 3) LINEAR bytes, the way the reference's image class would hold a file of this kind: stb_image's
    stbi_loadf linearises LDR data with pow(x / 255, 2.2) (src/stb_image.h:1576, 1872) and image::convert_to_bytes quantises
    with float_to_byte (src/image.h:99-117)."""
    import io

    from PIL import Image

    if "uri" in image:
        uri = image["uri"]
        raw = base64.b64decode(uri.split(",", 1)[1]) if uri.startswith("data:") else open(os.path.join(g.dir, uri), "rb").read()
    else:
        v = g.j["bufferViews"][image["bufferView"]]
        data = g.buffer(v["buffer"])
        raw = bytes(data[v.get("byteOffset", 0): v.get("byteOffset", 0) + v["byteLength"]])
    rgb = np.asarray(Image.open(io.BytesIO(raw)).convert("RGB"), dtype=np.uint8)
    lin = np.power(rgb.astype(np.float32) / np.float32(255.0), np.float32(2.2)).astype(np.float32)
    out = np.where(lin <= 0, 0, np.where(lin >= 1, 255, (256.0 * lin.astype(np.float64)).astype(np.int64)))
    return out.astype(np.uint8)


def load_mesh(path, mode="full"):
    """-> dict(triangles float32 (n, 9), uv float64 (n, 6) | None, material int32 (n,), materials [glTF material dicts],
    images {glTF image index: uint8 (H, W, 3) linear bytes}): everything `load_triangles` gives plus what the reference drops."""
    if mode not in ("reference", "full"):
        raise ValueError("mode must be 'reference' or 'full'")
    g = Gltf(path)
    tris, uvs, mats = [], [], []
    have_uv = False
    inst = [(i, np.eye(4)) for i in range(len(g.j.get("meshes", [])))] if mode == "reference" else g.mesh_instances()
    for mesh_i, world in inst:
        for prim in g.j["meshes"][mesh_i].get("primitives", []):
            t = _primitive_triangles(g, prim, mode).astype(np.float64)
            if mode == "full":
                t = t @ world[:3, :3].T + world[:3, 3]
            n = len(t)
            tris.append(t.astype(np.float32))
            uv = _primitive_uv(g, prim, n) if n else None
            have_uv |= uv is not None
            uvs.append(uv if uv is not None else np.zeros((n, 6)))
            mats.append(np.full(n, prim.get("material", -1), np.int32))
    if not tris:
        return dict(triangles=np.zeros((0, 9), np.float32), uv=None, material=np.zeros(0, np.int32), materials=[], images={})
    materials = g.j.get("materials", [])
    images = {}
    for m in materials:
        tex = m.get("pbrMetallicRoughness", {}).get("baseColorTexture")
        if tex is not None:
            src = g.j["textures"][tex["index"]].get("source")
            if src is not None and src not in images:
                images[src] = _decode_image(g, g.j["images"][src])
    return dict(triangles=np.ascontiguousarray(np.concatenate(tris).reshape(-1, 9), dtype=np.float32), uv=np.concatenate(uvs) if have_uv else None,
                material=np.concatenate(mats), materials=materials, images=images, textures=g.j.get("textures", []))


def mesh_scene(mesh, light=None, background_rgb=None, default_rgb=(0.8, 0.8, 0.8), name="gltf"):
    """A renderable Scene from `load_mesh` output, mapping glTF's metallic-roughness materials onto the reference's own
    material classes (src/material.h), since those are what the backend shades:
      emissiveFactor > 0                      -> diffuse_light(emissive * KHR_materials_emissive_strength)
      KHR_materials_transmission > 0.5        -> dielectric(KHR_materials_ior, default 1.5)
      metallicFactor >= 0.5                   -> metal(base colour, fuzz = roughnessFactor)
      otherwise                               -> lambertian(base colour)
    base colour = baseColorTexture as a picture_texture (nearest lookup, the reference's) when the mesh has TEXCOORD_0, else the
    solid baseColorFactor.  `light` = (corner, u, v, emission rgb) adds a quad light that is also passed as `light`."""
    from . import images as img_mod
    from . import scene as scene_mod

    mats = mesh["materials"]
    n_m = len(mats) + 1  # + the default material for primitives without one
    textures = np.zeros(n_m, dtype=abi.texture_dtype)
    materials = np.zeros(n_m, dtype=abi.material_dtype)
    materials["tex"] = np.arange(n_m)
    textures["kind"] = abi.TEX_SOLID
    textures["c0"][n_m - 1] = default_rgb
    materials["kind"][n_m - 1] = abi.MAT_LAMBERTIAN
    pending_pictures = []
    for i, m in enumerate(mats):
        pbr = m.get("pbrMetallicRoughness", {})
        base = pbr.get("baseColorFactor", [1, 1, 1, 1])[:3]
        ext = m.get("extensions", {})
        emissive = np.array(m.get("emissiveFactor", [0, 0, 0]), float) * ext.get("KHR_materials_emissive_strength", {}).get("emissiveStrength", 1.0)
        textures["c0"][i] = base
        if emissive.max() > 0:
            materials["kind"][i] = abi.MAT_DIFFUSE_LIGHT
            textures["c0"][i] = emissive
        elif ext.get("KHR_materials_transmission", {}).get("transmissionFactor", 0) > 0.5:
            materials["kind"][i] = abi.MAT_DIELECTRIC
            materials["p0"][i] = ext.get("KHR_materials_ior", {}).get("ior", 1.5)
        elif pbr.get("metallicFactor", 1.0) >= 0.5 and "metallicFactor" in pbr:
            materials["kind"][i] = abi.MAT_METAL
            materials["p0"][i] = min(1.0, max(0.0, pbr.get("roughnessFactor", 1.0)))
        else:
            materials["kind"][i] = abi.MAT_LAMBERTIAN
        tex = pbr.get("baseColorTexture")
        if tex is not None and mesh["uv"] is not None and emissive.max() == 0:
            src = mesh["textures"][tex["index"]].get("source")
            if src in mesh["images"]:
                pending_pictures.append((i, src))
    tri_mat = np.where(mesh["material"] < 0, n_m - 1, mesh["material"]).astype(np.int32)
    quad = None
    if light is not None:
        corner, u, v, rgb = light
        textures = np.concatenate([textures, np.zeros(1, abi.texture_dtype)])
        materials = np.concatenate([materials, np.zeros(1, abi.material_dtype)])
        textures["kind"][-1] = abi.TEX_SOLID
        textures["c0"][-1] = rgb
        materials["kind"][-1] = abi.MAT_DIFFUSE_LIGHT
        materials["tex"][-1] = len(textures) - 1
        quad = (corner, u, v, len(materials) - 1)
    bg = -1
    if background_rgb is not None:
        textures = np.concatenate([textures, np.zeros(1, abi.texture_dtype)])
        textures["kind"][-1] = abi.TEX_SOLID
        textures["c0"][-1] = background_rgb
        bg = len(textures) - 1
    s = scene_mod.triangle_scene(mesh["triangles"], materials, textures, tri_mat=tri_mat, light=quad, background_tex=bg, name=name)
    rows = {}
    for i, src in pending_pictures:
        if src not in rows:
            rows[src] = img_mod.add_image(s, mesh["images"][src])
        s.textures["kind"][i] = abi.TEX_PICTURE
        s.textures["aux"][i] = rows[src]
    if mesh["uv"] is not None:
        s.tri_uv = np.ascontiguousarray(mesh["uv"], dtype=np.float64)
    return s
