"""Flattened scene container: the host-side mirror of b200rt_scene_desc (include/b200rt.h).

A Scene owns numpy arrays laid out exactly as the C ABI expects; `desc()` hands out a
b200rt_scene_desc pointing into them.  Scenes come from three places:
  * the reference-side adaptor (include/b200rt_adaptor.hpp) walking live reference objects —
    `Scene.from_desc_pointer` copies such a description out of foreign memory;
  * committed fixtures under tests/golden/*.npz (`Scene.load`);
  * generators in this module for assets that are missing from the reference checkout
    (Sponza stand-in mesh).
"""
import ctypes as C

import numpy as np

from . import _abi as abi

_ARRAYS = [
    # name, dtype, trailing shape
    ("nodes", abi.node_dtype, ()),
    ("list_children", np.dtype("<u4"), ()),
    ("spheres", np.dtype("<f8"), (abi.SPHERE_STRIDE,)),
    ("sphere_mat", np.dtype("<i4"), ()),
    ("quads", np.dtype("<f8"), (abi.QUAD_STRIDE,)),
    ("quad_mat", np.dtype("<i4"), ()),
    ("tris", np.dtype("<f8"), (abi.TRI_STRIDE,)),
    ("tri_mat", np.dtype("<i4"), ()),
    ("volumes", abi.volume_dtype, ()),
    ("materials", abi.material_dtype, ()),
    ("textures", abi.texture_dtype, ()),
    ("images", abi.image_dtype, ()),
    ("image_bytes", np.dtype("u1"), ()),
    ("perlin_vec", np.dtype("<f8"), (256, 3)),
    ("perlin_perm", np.dtype("<i4"), (256,)),
    ("value_noise", abi.value_noise_dtype, ()),
    ("value_lattice", np.dtype("<f4"), ()),
]
_SCALARS = ["root", "n_primitives", "light_kind", "light_index", "background_tex"]


def _view(ptr, count, dtype, shape):
    """Copy `count` records of `dtype`+`shape` from raw address `ptr`."""
    n_items = int(count)
    full = (n_items,) + tuple(shape)
    if n_items == 0 or not ptr:
        return np.zeros(full, dtype=dtype)
    nbytes = int(np.prod(full)) * dtype.itemsize
    buf = (C.c_char * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(full).copy()


class Scene:
    def __init__(self):
        for name, dt, shape in _ARRAYS:
            setattr(self, name, np.zeros((0,) + shape, dtype=dt))
        self.root = 0
        self.n_primitives = 0
        self.light_kind = abi.LIGHT_NONE
        self.light_index = -1
        self.background_tex = -1
        self.name = ""
        self.tri_uv = np.zeros((0, 6))  # optional (ABI 2): per-triangle vertex texture coordinates, empty = the reference's behaviour

    # ---- construction -------------------------------------------------------------------------
    @classmethod
    def from_desc_pointer(cls, ptr):
        """Deep-copy a b200rt_scene_desc that lives in foreign (adaptor-owned) memory."""
        d = C.cast(ptr, C.POINTER(abi.SceneDesc)).contents
        if d.abi_version != abi.ABI_VERSION:
            raise ValueError(f"scene description has ABI version {d.abi_version}, expected {abi.ABI_VERSION}")
        s = cls()
        counts = {
            "nodes": d.n_nodes,
            "list_children": d.n_list_children,
            "spheres": d.n_spheres,
            "sphere_mat": d.n_spheres,
            "quads": d.n_quads,
            "quad_mat": d.n_quads,
            "tris": d.n_tris,
            "tri_mat": d.n_tris,
            "volumes": d.n_volumes,
            "materials": d.n_materials,
            "textures": d.n_textures,
            "images": d.n_images,
            "image_bytes": d.image_bytes_size,
            "perlin_vec": d.n_perlin,
            "perlin_perm": d.n_perlin,
            "value_noise": d.n_value_noise,
            "value_lattice": d.value_lattice_size,
        }
        for name, dt, shape in _ARRAYS:
            setattr(s, name, _view(getattr(d, name), counts[name], dt, shape))
        for k in _SCALARS:
            setattr(s, k, int(getattr(d, k)))
        if d.tri_uv:
            s.tri_uv = _view(d.tri_uv, d.n_tris, np.dtype("<f8"), (6,))
        return s

    def save(self, path, **extra):
        payload = {name: getattr(self, name) for name, _, _ in _ARRAYS}
        if len(self.tri_uv):
            payload["tri_uv"] = self.tri_uv
        payload["scalars"] = np.array([getattr(self, k) for k in _SCALARS], dtype=np.int64)
        payload.update(extra)
        np.savez_compressed(path, **payload)

    @classmethod
    def load(cls, path_or_npz):
        z = np.load(path_or_npz) if not hasattr(path_or_npz, "files") else path_or_npz
        s = cls()
        for name, dt, shape in _ARRAYS:
            arr = z[name]
            setattr(s, name, np.ascontiguousarray(arr.astype(dt, copy=False)).reshape((-1,) + shape))
        for k, v in zip(_SCALARS, z["scalars"].tolist()):
            setattr(s, k, int(v))
        if "tri_uv" in getattr(z, "files", []):
            s.tri_uv = np.ascontiguousarray(z["tri_uv"], dtype="<f8").reshape(-1, 6)
        return s

    # ---- C ABI view -----------------------------------------------------------------------------
    def desc(self):
        """-> (b200rt_scene_desc, keepalive).  The description borrows this Scene's arrays."""
        keep = []

        def p(name):
            a = np.ascontiguousarray(getattr(self, name))
            keep.append(a)
            return a.ctypes.data if a.size else None

        d = abi.SceneDesc()
        d.abi_version = abi.ABI_VERSION
        d.root = self.root
        d.n_nodes = len(self.nodes)
        d.n_list_children = len(self.list_children)
        d.nodes = p("nodes")
        d.list_children = p("list_children")
        d.n_spheres = len(self.sphere_mat)
        d.n_quads = len(self.quad_mat)
        d.n_tris = len(self.tri_mat)
        d.n_volumes = len(self.volumes)
        for name in ("spheres", "sphere_mat", "quads", "quad_mat", "tris", "tri_mat", "volumes", "materials", "textures",
                     "images", "image_bytes", "perlin_vec", "perlin_perm", "value_noise", "value_lattice"):
            setattr(d, name, p(name))
        d.n_materials = len(self.materials)
        d.n_textures = len(self.textures)
        d.n_images = len(self.images)
        d.n_perlin = len(self.perlin_perm)
        d.n_value_noise = len(self.value_noise)
        d.n_primitives = self.n_primitives
        d.image_bytes_size = self.image_bytes.size
        d.value_lattice_size = self.value_lattice.size
        d.light_kind = self.light_kind
        d.light_index = self.light_index
        d.background_tex = self.background_tex
        if len(self.tri_uv):
            if len(self.tri_uv) != len(self.tri_mat):
                raise ValueError("tri_uv must have one row per triangle")
            self.tri_uv = np.ascontiguousarray(self.tri_uv, dtype="<f8").reshape(-1, 6)
            d.tri_uv = p("tri_uv")
        return d, keep

    # ---- summaries ------------------------------------------------------------------------------
    def counts(self):
        return {
            "nodes": len(self.nodes),
            "spheres": len(self.sphere_mat),
            "quads": len(self.quad_mat),
            "tris": len(self.tri_mat),
            "volumes": len(self.volumes),
            "materials": len(self.materials),
            "textures": len(self.textures),
            "primitives": self.n_primitives,
        }


# --------------------------------------------------------------------------------------------------
# Sponza stand-in (assets/Sponza/glTF/Sponza.bin is missing from the reference checkout,
# /root/reference/.MISSING_LARGE_BLOBS).  A deterministic atrium with the real asset's triangle count
# (262 267, from Sponza.gltf's accessors) and raw bounds x[-1920.9,1799.9] y[-126.4,1429.4]
# z[-1182.8,1105.4]: floor, walls, two colonnade storeys of tessellated columns and
# arches, banners.  Vertices are float32 like glTF positions.  Labelled "Sponza-standin" wherever used.
SPONZA_TRIANGLES = 262_267
SPONZA_BOUNDS = ((-1920.9, 1799.9), (-126.4, 1429.4), (-1182.8, 1105.4))


def _grid_quad(o, u, v, nu, nv):
    """Tessellate parallelogram o + s u + t v into nu x nv cells -> (2*nu*nv, 3, 3) float64."""
    s = np.linspace(0.0, 1.0, nu + 1)
    t = np.linspace(0.0, 1.0, nv + 1)
    P = o[None, None, :] + s[:, None, None] * u[None, None, :] + t[None, :, None] * v[None, None, :]
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    t1 = np.stack([a, b, c], axis=2).reshape(-1, 3, 3)
    t2 = np.stack([a, c, d], axis=2).reshape(-1, 3, 3)
    return np.concatenate([t1, t2], axis=0)


def _column(cx, cz, y0, y1, r, seg, rings):
    """Fluted column as a tessellated cylinder -> (2*seg*rings, 3, 3)."""
    ang = np.linspace(0.0, 2.0 * np.pi, seg + 1)
    ys = np.linspace(y0, y1, rings + 1)
    rr = r * (1.0 + 0.06 * np.cos(8.0 * ang))
    X = cx + rr[None, :] * np.cos(ang)[None, :] + 0.0 * ys[:, None]
    Z = cz + rr[None, :] * np.sin(ang)[None, :] + 0.0 * ys[:, None]
    Y = ys[:, None] + 0.0 * ang[None, :]
    P = np.stack([X, Y, Z], axis=-1)
    a, b, c, d = P[:-1, :-1], P[:-1, 1:], P[1:, 1:], P[1:, :-1]
    t1 = np.stack([a, b, c], axis=2).reshape(-1, 3, 3)
    t2 = np.stack([a, c, d], axis=2).reshape(-1, 3, 3)
    return np.concatenate([t1, t2], axis=0)


def sponza_standin_triangles(n_triangles=SPONZA_TRIANGLES, seed=7):
    """-> (n_triangles, 9) float32, deterministic."""
    (x0, x1), (y0, y1), (z0, z1) = SPONZA_BOUNDS
    f = lambda *a: np.array(a, dtype=np.float64)
    parts = []
    # floor, ceiling band, four walls (coarse + fine tessellation like the real asset's mixed density)
    parts.append(_grid_quad(f(x0, y0, z0), f(x1 - x0, 0, 0), f(0, 0, z1 - z0), 96, 64))
    parts.append(_grid_quad(f(x0, y0, z0), f(x1 - x0, 0, 0), f(0, y1 - y0, 0), 80, 40))
    parts.append(_grid_quad(f(x0, y0, z1), f(x1 - x0, 0, 0), f(0, y1 - y0, 0), 80, 40))
    parts.append(_grid_quad(f(x0, y0, z0), f(0, 0, z1 - z0), f(0, y1 - y0, 0), 48, 40))
    parts.append(_grid_quad(f(x1, y0, z0), f(0, 0, z1 - z0), f(0, y1 - y0, 0), 48, 40))
    # gallery floors of the upper storey along both long sides
    for zz0, zz1 in ((z0, z0 + 420.0), (z1 - 420.0, z1)):
        parts.append(_grid_quad(f(x0, 560.0, zz0), f(x1 - x0, 0, 0), f(0, 0, zz1 - zz0), 96, 12))
    # two storeys of columns on both sides
    xs = np.linspace(x0 + 260.0, x1 - 260.0, 12)
    for storey, (cy0, cy1, r) in enumerate(((y0, 540.0, 60.0), (580.0, 1080.0, 45.0))):
        for zc in (z0 + 430.0, z1 - 430.0):
            for xc in xs:
                parts.append(_column(xc, zc, cy0, cy1, r, 48, 36))
    # hanging banners (thin, slightly curved sheets)
    for k, xc in enumerate(np.linspace(x0 + 500.0, x1 - 500.0, 6)):
        o = f(xc, 300.0, -160.0 + 50.0 * (k % 3))
        parts.append(_grid_quad(o, f(0, 0, 320.0), f(30.0, 620.0, 0), 40, 60))
    tris = np.concatenate(parts, axis=0)
    rng = np.random.RandomState(seed)
    if len(tris) < n_triangles:
        # fill up with small clutter triangles (ivy / vases / lion heads in the real asset) inside the bounds
        need = n_triangles - len(tris)
        centre = np.stack([rng.uniform(x0 + 50, x1 - 50, need), rng.uniform(y0 + 5, 1000.0, need), rng.uniform(z0 + 50, z1 - 50, need)], axis=1)
        # cluster them near walls / columns so the atrium's centre stays open
        centre[:, 2] = np.where(rng.rand(need) < 0.5, z0 + 60 + np.abs(rng.normal(0, 120, need)), z1 - 60 - np.abs(rng.normal(0, 120, need)))
        off = rng.normal(0.0, 9.0, (need, 3, 3))
        clutter = centre[:, None, :] + off
        tris = np.concatenate([tris, clutter], axis=0)
    tris = tris[:n_triangles]
    lo = np.array([x0, y0, z0])
    hi = np.array([x1, y1, z1])
    tris = np.clip(tris, lo, hi)
    return np.ascontiguousarray(tris.reshape(-1, 9).astype(np.float32))


def reference_like_bvh(boxes):
    """Node table of the tree the reference's bvh_node constructor builds over `boxes` ((n, 6) rows x0,x1,y0,y1,z0,z1):
    objects sorted by bounding-box x-min, ALWAYS axis 0 (src/bvh_node.h:20-31; a range of an array already sorted by the
    same key stays sorted, so one stable sort fixes every level), median split (begin+end)/2 (:42-44), two objects ->
    both leaves (:37-39), one object -> left_ == right_ (:34-36); box_ = aabb::enclose(left, right) through the interval
    constructor, i.e. padded to 1e-4 where thinner (src/aabb.h:7-12, 35-40, 81-86).
    -> (inner (m,) node_dtype rows whose a/b are: >= 0 inner-node index, < 0 ~object, root index)."""
    n = len(boxes)
    order = np.argsort(boxes[:, 0], kind="stable")
    inner = []  # [a, b, box]

    def pad(b):
        b = b.copy()
        for k in range(3):
            if b[2 * k + 1] - b[2 * k] < 0.0001:
                b[2 * k] -= 0.00005
                b[2 * k + 1] += 0.00005
        return b

    def enclose(b1, b2):
        out = np.empty(6)
        out[0::2] = np.minimum(b1[0::2], b2[0::2])
        out[1::2] = np.maximum(b1[1::2], b2[1::2])
        return pad(out)

    import sys

    sys.setrecursionlimit(max(10000, sys.getrecursionlimit()))

    def build(lo, hi):
        me = len(inner)
        inner.append(None)
        cnt = hi - lo
        if cnt == 1:
            a = b = ~int(order[lo])
            box = enclose(boxes[order[lo]], boxes[order[lo]])
        elif cnt == 2:
            a, b = ~int(order[lo]), ~int(order[lo + 1])
            box = enclose(boxes[order[lo]], boxes[order[lo + 1]])
        else:
            mid = (lo + hi) // 2
            a, ba = build(lo, mid)
            b, bb = build(mid, hi)
            box = enclose(ba, bb)
        inner[me] = (a, b, box)
        return me, box

    root, _ = build(0, n)
    return inner, root


def triangle_scene(t32, materials, textures, tri_mat=None, light=None, background_tex=-1, name="mesh"):
    """A scene as the reference's mesh functions build it (glass_fox src/main.cc:345-400, sponza src/main.cc:439-498): one
    `triangle` per row of `t32` (float32 (n, 9), e.g. from gltf.load_triangles), optionally one quad `light` = (corner, u, v,
    material index) that is both in the world and passed as `light`, ONE bvh_node over everything with the reference's
    x-sorted median splits (src/bvh_node.h:18-47).  `materials` / `textures` are tables in abi.material_dtype / abi.texture_dtype."""
    t = np.asarray(t32, dtype=np.float32).reshape(-1, 9).astype(np.float64)
    p0, p1, p2 = t[:, 0:3], t[:, 3:6], t[:, 6:9]
    e1, e2 = p1 - p0, p2 - p0
    c = np.stack([e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1], e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2], e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]], axis=1)
    with np.errstate(invalid="ignore", divide="ignore"):
        nrm = c / np.sqrt(c[:, 0] * c[:, 0] + c[:, 1] * c[:, 1] + c[:, 2] * c[:, 2])[:, None]  # unit_vector(cross(p1-p0, p2-p0)), src/triangle.h:24
    s = Scene()
    s.name = name
    s.tris = np.ascontiguousarray(np.concatenate([t, nrm], axis=1))
    s.tri_mat = np.zeros(len(t), dtype="<i4") if tri_mat is None else np.ascontiguousarray(tri_mat, dtype="<i4")
    s.textures = np.ascontiguousarray(textures, dtype=abi.texture_dtype)
    s.materials = np.ascontiguousarray(materials, dtype=abi.material_dtype)
    nt = len(t)
    n_obj = nt + (1 if light is not None else 0)
    # object boxes: triangle = aabb(min, max) unpadded (src/triangle.h:42-48); quad = enclose of its two diagonals (src/quad.h:19-21)
    boxes = np.empty((n_obj, 6))
    P = t.reshape(nt, 3, 3)
    boxes[:nt, 0::2] = P.min(axis=1)
    boxes[:nt, 1::2] = P.max(axis=1)
    if light is not None:
        # quad row: corner, u, v, unorm, w = n / dot(n, n), D = dot(unorm, corner), area (include/b200rt.h B200RT_QUAD_STRIDE)
        corner, u, v = (np.asarray(x, dtype=np.float64) for x in light[:3])
        nq = np.array([u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]])
        area = float(np.sqrt(nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2]))
        unorm = nq / area
        w = nq / (nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2])
        q = np.zeros((1, abi.QUAD_STRIDE))
        q[0, 0:3], q[0, 3:6], q[0, 6:9], q[0, 9:12], q[0, 12:15] = corner, u, v, unorm, w
        q[0, 15] = unorm[0] * corner[0] + unorm[1] * corner[1] + unorm[2] * corner[2]
        q[0, 16] = area
        s.quads = q
        s.quad_mat = np.array([int(light[3])], dtype="<i4")
        qb = np.stack([corner, corner + u, corner + v, corner + u + v])
        qbox = np.empty(6)
        qbox[0::2], qbox[1::2] = qb.min(axis=0), qb.max(axis=0)
        for k in range(3):
            if qbox[2 * k + 1] - qbox[2 * k] < 0.0001:
                qbox[2 * k] -= 0.00005
                qbox[2 * k + 1] += 0.00005
        boxes[nt] = qbox
    inner, root = reference_like_bvh(boxes)
    m = len(inner)
    nodes = np.zeros(m + n_obj, dtype=abi.node_dtype)
    # leaves first: node i (< nt) = triangle i, node nt = the quad; inner node k lives at n_obj + k
    nodes["kind"][:nt] = abi.NODE_TRIANGLE
    nodes["a"][:nt] = np.arange(nt)
    if light is not None:
        nodes["kind"][nt] = abi.NODE_QUAD
        nodes["a"][nt] = 0
    ia = np.array([x[0] for x in inner], dtype=np.int64)
    ib = np.array([x[1] for x in inner], dtype=np.int64)
    nodes["kind"][n_obj:] = abi.NODE_BVH
    nodes["a"][n_obj:] = np.where(ia >= 0, ia + n_obj, ~ia)
    nodes["b"][n_obj:] = np.where(ib >= 0, ib + n_obj, ~ib)
    nodes["f"][n_obj:] = np.array([x[2] for x in inner])
    # primitive ids = order of first visit in the reference's left-to-right depth-first walk
    ids = np.full(n_obj, -1, dtype=np.int64)
    nxt = 0
    stack = [root]
    while stack:
        k = stack.pop()
        if k < 0:
            o = ~k
            if ids[o] < 0:
                ids[o] = nxt
                nxt += 1
            continue
        stack.append(int(ib[k]))
        stack.append(int(ia[k]))
    nodes["c"][:n_obj] = ids
    s.nodes = nodes
    s.root = n_obj + root
    s.n_primitives = n_obj
    s.light_kind = abi.LIGHT_QUAD if light is not None else abi.LIGHT_NONE
    s.light_index = 0
    s.background_tex = background_tex
    return s


def sponza_standin_scene(n_triangles=SPONZA_TRIANGLES, seed=7):
    """BASELINE.json configs[4] as the reference's sponza() builds it (src/main.cc:439-489) over the stand-in mesh:
    every triangle lambertian white, quad light (0,1200,0) 500x500 emitting 10, one bvh_node over everything, the quad
    passed as `light`.  Labelled "Sponza-standin": the real Sponza.bin is missing from the reference checkout."""
    textures = np.zeros(2, dtype=abi.texture_dtype)
    textures["kind"] = abi.TEX_SOLID
    textures["c0"][0] = 1.0
    textures["c0"][1] = 10.0
    materials = np.zeros(2, dtype=abi.material_dtype)
    materials["kind"] = [abi.MAT_LAMBERTIAN, abi.MAT_DIFFUSE_LIGHT]
    materials["tex"] = [0, 1]
    return triangle_scene(sponza_standin_triangles(n_triangles, seed), materials, textures, light=((0.0, 1200.0, 0.0), (500.0, 0.0, 0.0), (0.0, 0.0, 500.0), 1),
                          name="sponza-standin")
