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
