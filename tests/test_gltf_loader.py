"""Host data path in front of the C ABI: glTF -> triangle rows -> scene description.
reference mode must give what the reference's own loader + copy loops give (src/gltf_loader.h, src/main.cc:346-393)."""
import base64
import json
import os
import struct

import numpy as np
import pytest
from conftest import ROOT, abi, pkg, same

gltf = pkg.gltf


def synthetic(tmp_path, binary=False):
    pos_a = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0]], np.float32)            # indexed, u16
    idx_a = np.array([0, 1, 2, 2, 1, 3], np.uint16)
    pos_b = np.array([[0, 0, 1], [2, 0, 1], [0, 2, 1], [5, 5, 5]], np.float32)            # not indexed: 1 triangle + 1 stray vertex
    pos_c = np.array([[0, 0, 2], [3, 0, 2], [0, 3, 2]], np.float32)                        # indexed, u32
    idx_c = np.array([2, 1, 0], np.uint32)
    blobs = [pos_a.tobytes(), idx_a.tobytes(), pos_b.tobytes(), pos_c.tobytes(), idx_c.tobytes()]
    views, off = [], 0
    for b in blobs:
        views.append({"buffer": 0, "byteOffset": off, "byteLength": len(b)})
        off += (len(b) + 3) // 4 * 4
    data = b"".join(b + b"\0" * ((-len(b)) % 4) for b in blobs)
    acc = [{"bufferView": 0, "componentType": 5126, "count": 4, "type": "VEC3"}, {"bufferView": 1, "componentType": 5123, "count": 6, "type": "SCALAR"},
           {"bufferView": 2, "componentType": 5126, "count": 4, "type": "VEC3"}, {"bufferView": 3, "componentType": 5126, "count": 3, "type": "VEC3"},
           {"bufferView": 4, "componentType": 5125, "count": 3, "type": "SCALAR"}]
    j = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}],
         "nodes": [{"translation": [10, 0, 0], "scale": [2, 2, 2], "children": [1], "mesh": 0}, {"mesh": 1, "translation": [0, 1, 0]}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0}, "indices": 1}, {"attributes": {"POSITION": 2}}]},
                    {"primitives": [{"attributes": {"POSITION": 3}, "indices": 4}]}],
         "accessors": acc, "bufferViews": views, "buffers": [{"byteLength": len(data)}]}
    if binary:
        js = json.dumps(j).encode()
        js += b" " * ((-len(js)) % 4)
        body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(data), 0x004E4942) + data
        path = tmp_path / "m.glb"
        path.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    else:
        j["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(data).decode()
        path = tmp_path / "m.gltf"
        path.write_text(json.dumps(j))
    return str(path), (pos_a, idx_a, pos_b, pos_c, idx_c)


@pytest.mark.parametrize("binary", [False, True])
def test_reference_mode_reads_what_the_reference_reads(tmp_path, binary):
    path, (pa, ia, pb, pc, ic) = synthetic(tmp_path, binary)
    t = gltf.load_triangles(path, "reference")
    want = np.concatenate([pa[ia].reshape(-1, 9), pb[:3].reshape(-1, 9)])  # u16-indexed + non-indexed; the u32 primitive gives nothing; no transforms
    assert t.dtype == np.float32 and np.array_equal(t, want)


def test_full_mode_applies_nodes_and_all_index_types(tmp_path):
    path, (pa, ia, pb, pc, ic) = synthetic(tmp_path)
    t = gltf.load_triangles(path, "full")
    root = lambda p: p.astype(np.float64) * 2 + [10, 0, 0]
    child = lambda p: (p.astype(np.float64) + [0, 1, 0]) * 2 + [10, 0, 0]
    want = np.concatenate([root(pa[ia]).reshape(-1, 9), root(pb[:3]).reshape(-1, 9), child(pc[ic]).reshape(-1, 9)]).astype(np.float32)
    assert np.allclose(t, want, rtol=0, atol=1e-6) and len(t) == 4


def test_triangle_scene_is_what_the_oracle_traces(tmp_path, build_oracle):
    """glTF -> rows -> scene description -> oracle: every triangle can be hit, ids are the reference's visiting order."""
    path, _ = synthetic(tmp_path)
    t = gltf.load_triangles(path, "full")
    tex = np.zeros(2, dtype=abi.texture_dtype)
    tex["kind"] = abi.TEX_SOLID
    tex["c0"][0], tex["c0"][1] = 0.7, 5.0
    mat = np.zeros(2, dtype=abi.material_dtype)
    mat["kind"], mat["tex"] = [abi.MAT_LAMBERTIAN, abi.MAT_DIFFUSE_LIGHT], [0, 1]
    sc = pkg.scene.triangle_scene(t, mat, tex, light=((0, 30, 0), (5, 0, 0), (0, 0, 5), 1))
    assert sc.n_primitives == len(t) + 1
    cent = t.reshape(-1, 3, 3).astype(np.float64).mean(axis=1)
    rays = np.zeros((len(cent), 7))
    rays[:, 0:3] = cent + [0.01, 0.02, 0.3]  # just above each triangle: nothing in between
    rays[:, 3:6] = cent - rays[:, 0:3]
    out = build_oracle.trace(sc, rays)
    assert (out["prim"] >= 0).all() and len(set(out["prim"].tolist())) == len(t)
    assert np.allclose(out["t"], 1.0, atol=1e-9)


@pytest.mark.skipif(not os.path.exists("/root/reference/assets/Fox/glTF/Fox.gltf"), reason="the Fox asset lives next to the reference checkout only")
def test_fox_triangles_equal_the_references_own_load():
    """tests/golden/fox.npz was flattened from the world glass_fox() builds with the reference's loader: same triangle set."""
    t = gltf.load_triangles("/root/reference/assets/Fox/glTF/Fox.gltf", "reference")
    z = np.load(os.path.join(ROOT, "tests", "golden", "fox.npz"))
    ref = pkg.scene.Scene.load(z).tris[:, :9].astype(np.float32)
    assert t.shape == ref.shape == (576, 9)
    key = lambda a: a[np.lexsort(a.T[::-1])]
    assert np.array_equal(key(t), key(ref))


# ------------------------------------------------------------------------------------------------ materials, textures, uv
FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "aux", "textured.gltf")


def textured_scene():
    m = pkg.gltf.load_mesh(FIXTURE)
    return m, pkg.gltf.mesh_scene(m, light=((-1, 6, -3), (3, 0, 0), (0, 0, 3), (8, 8, 8)), background_rgb=(0.2, 0.2, 0.3))


def test_load_mesh_reads_what_the_reference_drops():
    """TEXCOORD_0, per-primitive materials, the base-colour texture, u8 / u16 / u32 indices and the node transform
    (src/gltf_loader.h:65 `// Material material;`, :312 `// loadMaterials();`, :295-303; src/main.cc:367-375)."""
    m, s = textured_scene()
    assert m["triangles"].shape == (60, 9) and m["uv"].shape == (60, 6)
    assert np.bincount(m["material"]).tolist() == [32, 8, 2, 18]  # 4x4, 2x2, 1x1 and 3x3 grids of quads
    assert m["images"][0].shape == (8, 8, 3) and m["images"][0].dtype == np.uint8
    # the node's scale 1.5 and translation (0.5, 1, -2) are applied: the first grid spans x in [-3, -1] * 1.5 + 0.5
    x = m["triangles"][:32].reshape(-1, 3)[:, 0]
    assert np.isclose(x.min(), -4.0) and np.isclose(x.max(), -1.0)
    kinds = s.materials["kind"].tolist()
    assert kinds[:4] == [abi.MAT_LAMBERTIAN, abi.MAT_METAL, abi.MAT_DIFFUSE_LIGHT, abi.MAT_DIELECTRIC]
    assert s.textures["kind"][0] == abi.TEX_PICTURE and np.isclose(s.materials["p0"][1], 0.25) and np.isclose(s.materials["p0"][3], 1.45)
    assert np.allclose(s.textures["c0"][2], [6.0, 5.4, 4.8])
    # the reference's own behaviour stays available: positions only, file order, no transform
    ref = pkg.gltf.load_triangles(FIXTURE, mode="reference")
    assert len(ref) == 32 + 18  # only the u16-indexed primitives (src/main.cc:368)
    # linearised 8-bit texels, as stbi_loadf + image::convert_to_bytes would hold them
    assert m["images"][0][0, 0].tolist() == [255, 0, 0] and m["images"][0][7, 7].tolist() == [0, 0, 255]


def test_triangle_uv_interpolation_in_the_oracle(build_oracle):
    """b200rt_scene_desc::tri_uv: u, v of a triangle hit = the vertex coordinates weighted with the hit's barycentrics; without
    tri_uv the record carries 0, 0 as the reference leaves it (src/triangle.h:35-39)."""
    m, s = textured_scene()
    rng = np.random.RandomState(2)
    k = rng.randint(0, 60, 800)
    b = rng.dirichlet([1, 1, 1], 800)
    T = m["triangles"].astype(np.float64).reshape(-1, 3, 3)
    target = (b[:, :, None] * T[k]).sum(axis=1)
    rays = np.zeros((800, 7))
    rays[:, 0:3] = target + [0.1, 0.2, 5.0]
    rays[:, 3:6] = target - rays[:, 0:3]
    out = build_oracle.trace(s, rays)
    hit = out["prim"] >= 0
    assert hit.mean() > 0.95
    # which triangle was hit: nearest along the ray may be another layer; recompute the expectation from the hit point
    uv = s.tri_uv.reshape(-1, 3, 2)
    P = out["p"][hit]
    ok = 0
    for p, got in zip(P, out["uv"][hit]):
        for t in range(60):  # find a triangle containing p
            a, bb, c = T[t]
            n = np.cross(bb - a, c - a)
            if abs(np.dot(p - a, n)) > 1e-9 * np.linalg.norm(n):
                continue
            w1 = np.dot(np.cross(p - a, c - a), n) / np.dot(n, n)
            w2 = np.dot(np.cross(bb - a, p - a), n) / np.dot(n, n)
            if w1 >= -1e-9 and w2 >= -1e-9 and w1 + w2 <= 1 + 1e-9:
                want = (1 - w1 - w2) * uv[t, 0] + w1 * uv[t, 1] + w2 * uv[t, 2]
                if np.allclose(got, want, atol=1e-9):
                    ok += 1
                    break
    assert ok >= 0.99 * hit.sum(), (ok, hit.sum())
    s.tri_uv = np.zeros((0, 6))
    assert not build_oracle.trace(s, rays)["uv"].any()
