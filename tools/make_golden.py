#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference compiled here (oracle/_ref/libref_harness.so).

Run in the build container (needs /root/reference): `python tools/make_golden.py`.  Each file carries
  * the flattened scene exactly as include/b200rt_adaptor.hpp emits it from the live reference objects,
  * cameras (config resolution and a small test resolution) as b200rt_camera bytes,
  * reference outputs: generate_ray with the std::rand values it consumed, <world>.hit records on camera +
    secondary rays, texture::sample on random (u, v, p), and a multi-spp render (per-pixel sum and sum of squares).
The GPU box has no /root/reference; tests there compare against these files.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import refharness as R  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

# name -> (small width, render spp, has picture)
SCENES = {
    "cornell": (40, 512), "rtow": (48, 256), "rtnw": (40, 128), "fox": (64, 256),
    "three_ball": (48, 256), "three_ball_lens": (48, 256), "cornell_volume": (40, 256), "cornell_rotated": (40, 256),
    "cornell_specular": (40, 256), "perlin_ball": (40, 128), "light_earth": (48, 256), "skybox_fisheye": (40, 256),
    "fuzz_metal": (57, 256), "white_sphere": (40, 128), "motion_ball": (48, 64),
    "noise_perlin": (40, 16), "noise_value": (40, 16), "noise_worley": (40, 16), "noise_voronoi": (40, 16),
    # every material under each kind of `light` argument (none / quad / sphere: src/quad.h:66-78, src/sphere.h:76-81)
    "zoo_nolight": (48, 256), "zoo_quad": (48, 256), "zoo_sphere": (48, 256),
}


def secondary_rays(rs, cam_rays, rng):
    """Rays leaving first-hit points in random directions + a few axis-aligned and grazing ones."""
    h = rs.trace(cam_rays)
    hit = h["prim"] >= 0
    sec = cam_rays[hit].copy()
    sec[:, 0:3] = h["p"][hit]
    sec[:, 3:6] = rng.normal(size=(len(sec), 3))
    sec[:, 6] = rng.rand(len(sec))
    return sec


def aimed_rays(rs, rng, n=768):
    """Rays from random origins toward random interior points of random primitives (object-space coordinates)."""
    sc = rs.scene
    targets = []
    if len(sc.tri_mat):
        t = sc.tris[rng.randint(len(sc.tri_mat), size=n)]
        b = rng.dirichlet([1, 1, 1], size=n)
        targets.append(b[:, 0:1] * t[:, 0:3] + b[:, 1:2] * t[:, 3:6] + b[:, 2:3] * t[:, 6:9])
    if len(sc.sphere_mat):
        q = sc.spheres[rng.randint(len(sc.sphere_mat), size=n)]
        targets.append(q[:, 0:3] + rng.uniform(-0.6, 0.6, size=(n, 3)) * q[:, 3:4])
    if len(sc.quad_mat):
        q = sc.quads[rng.randint(len(sc.quad_mat), size=n)]
        targets.append(q[:, 0:3] + rng.rand(n, 1) * q[:, 3:6] + rng.rand(n, 1) * q[:, 6:9])
    tg = np.concatenate(targets)[rng.permutation(sum(len(t) for t in targets))[:n]]
    pos = np.array(list(rs.camera.pos))
    org = pos + rng.normal(size=(len(tg), 3)) * (0.2 * np.linalg.norm(tg - pos, axis=1, keepdims=True))
    rays = np.zeros((len(tg), 7))
    rays[:, 0:3] = org
    rays[:, 3:6] = (tg - org) * rng.uniform(0.3, 3.0, size=(len(tg), 1))  # unnormalised on purpose, like camera rays
    rays[:, 6] = rng.rand(len(tg))
    return rays


def scatter_vectors(rs, rays):
    """One level of ray_color below the hit test through the reference's own material / pdf objects, with the std::rand values
    it consumed (oracle/ref_harness.cc: ref_scatter_logged), on every third ray of the hit batch."""
    sub = rays[::3]
    g = rs.scatter_logged(sub, seed=5, max_draws=8)
    return dict(scatter_rays=sub, scatter_rec12=g["rec12"], scatter_out16=g["out16"], scatter_draws=g["draws"], scatter_counts=g["counts"])


def add_scatter(name, width):
    """Adds the scatter vectors to an existing golden file, leaving everything else in it untouched."""
    path = os.path.join(OUT, name + ".npz")
    z = dict(np.load(path))
    rs = R.RefScene(name, width=width)
    assert rs.scene.nodes.tobytes() == z["nodes"].tobytes(), "the scene built now must be the scene stored then"
    z.update(scatter_vectors(rs, z["trace_rays"]))
    np.savez_compressed(path, **z)
    m = z["scatter_out16"][:, 0]
    print(f"{name:20s} scatter vectors: {len(m)} rays, modes {np.bincount((m + 1).astype(int), minlength=4).tolist()} (miss / none / determined / random)")


def make(name, width, spp):
    import zlib

    rng = np.random.RandomState(zlib.crc32(name.encode()) % (2**31))
    full = R.RefScene(name)
    rs = R.RefScene(name, width=width)
    assert full.scene.nodes.tobytes() == rs.scene.nodes.tobytes(), "scene must not depend on the camera"
    extra = {"cam_full": np.frombuffer(full.camera.to_bytes(), np.uint8), "cam_small": np.frombuffer(rs.camera.to_bytes(), np.uint8)}
    rays, draws, counts = rs.generate_rays_logged(seed=4242, spp=2, max_draws=24)
    extra.update(ref_rays=rays, ref_draws=draws, ref_draw_counts=counts)
    finite = np.isfinite(rays).all(axis=1)
    tr = np.concatenate([rays[finite], secondary_rays(rs, rays[finite], rng), aimed_rays(rs, rng)])
    h = rs.trace(tr, seed=99)
    extra["trace_rays"] = tr
    for k, v in h.items():
        extra["trace_" + k] = v
    extra.update(scatter_vectors(rs, tr))
    # textures
    nt = len(rs.scene.textures)
    uvp = np.concatenate([rng.rand(64 * nt, 2), rng.uniform(-30, 600, size=(64 * nt, 3))], axis=1)
    if name.startswith("noise_value"):
        uvp[:, 2:5] = rng.uniform(0.0, 38.0, size=(64 * nt, 3))  # value_noise indexes its lattice without bounds checks (src/noise.h:109-116)
    rows = np.repeat(np.arange(nt), 64)
    rgb = np.zeros((len(uvp), 3))
    for t in range(nt):
        rgb[rows == t] = rs.texture_sample(t, uvp[rows == t])
    extra.update(tex_rows=rows, tex_uvp=uvp, tex_rgb=rgb)
    # render statistics (thread-local RNG streams per row: result independent of the thread count)
    s, q, st = rs.render(spp, seed=777, threads=0, rng_mode=1)
    extra.update(ref_sum=s, ref_sumsq=q, ref_spp=np.int64(spp), ref_samples=np.int64(st["samples"]), ref_rays_count=np.int64(st["rays"]),
                 ref_nan_samples=np.int64(st["nan_samples"]))
    path = os.path.join(OUT, name + ".npz")
    rs.scene.save(path, **extra)
    if name in ("cornell", "rtow", "rtnw", "fox"):  # BASELINE.json configs[0..3] as product scene files (scene + config camera only)
        pdir = os.path.join(ROOT, "cpu-ray-tracing-implementation_b200", "scenes")
        os.makedirs(pdir, exist_ok=True)
        full.scene.save(os.path.join(pdir, name + ".npz"), camera=extra["cam_full"])
    print(f"{name:20s} {os.path.getsize(path) / 1e3:9.1f} kB  prims={rs.scene.n_primitives} rays/sample={st['rays'] / st['samples']:.3f} "
          f"render={st['seconds']:.1f}s")


def make_write_color():
    """tests/golden/aux/write_color.npz: linear images and the bytes the reference's own output stage wrote for them.
    (a) a stock camera::render of a small Cornell box: cam.image_ and the P3 file it wrote (src/camera.h:149-151, 174);
    (b) crafted values through write_color (src/color.h:16-36): negative, zero, denormal, > 1, huge, inf, NaN, and values
        next to every integer boundary of int(255.999 * pow(c, 1/2.2))."""
    import tempfile

    rs = R.RefScene("cornell", width=20, spp=4, depth=6)
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "stock.ppm")
        _, img = rs.render_stock(spp=4, seed=1, ppm_path=path)
        stock_text = open(path, "rb").read()
    assert stock_text == R.write_colors(img), "camera::render's file is its header + write_color per pixel"
    rng = np.random.RandomState(7)
    k = np.arange(0, 300)
    edges = (k / 255.999) ** 2.2  # pow(c, 1/2.2) * 255.999 lands next to the integer k
    vals = np.concatenate([edges, np.nextafter(edges, 0), np.nextafter(edges, 1e9), edges * (1 - 1e-12), edges * (1 + 1e-12),
                           rng.rand(2000), 10 ** rng.uniform(-12, 6, 2000), -rng.rand(50),
                           [0.0, -0.0, 5e-324, 1e-310, 1.0, 2.0**31, 1e300, np.inf, -np.inf, np.nan, 255.0, 1e9, 8.4e6]])
    n = len(vals) // 3 * 3
    crafted = rng.permutation(vals[:n]).reshape(1, n // 3, 3)
    text = R.write_colors(crafted)
    out = os.path.join(OUT, "aux")
    os.makedirs(out, exist_ok=True)
    np.savez_compressed(os.path.join(out, "write_color.npz"), stock_image=img, stock_text=np.frombuffer(stock_text, np.uint8),
                        crafted_image=crafted, crafted_text=np.frombuffer(text, np.uint8))
    print(f"write_color: stock {img.shape} {len(stock_text)} B, crafted {crafted.shape} {len(text)} B")


# name -> (width, reference spp): sized so that the reference's own Monte-Carlo noise stays below ~6e-4 RMSE per channel
CONVERGED = {"cornell": (24, 1_500_000), "rtow": (24, 300_000), "rtnw": (16, 1_600_000), "fox": (32, 100_000), "sponza16k": (16, 400_000)}
SPONZA_SMALL = 16384  # triangles of the stand-in mesh used for the converged render: the reference's x-sorted tree needs 0.3 ms per ray on the full one


def append_converged(name, extra_spp):
    """More reference samples for an existing converged_<name>.npz (independent stream), summed into it."""
    path = os.path.join(OUT, "aux", f"converged_{name}.npz")
    z = dict(np.load(path))
    width = CONVERGED[name][0]
    rs = R.RefScene(name, width=width)
    assert rs.camera.to_bytes() == z["cam"].tobytes()
    s, q, st = rs.render(extra_spp, seed=20261105 + int(z["ref_spp"]), threads=0, rng_mode=1)
    z["ref_sum"], z["ref_sumsq"] = z["ref_sum"] + s, z["ref_sumsq"] + q
    z["ref_spp"] = np.int64(int(z["ref_spp"]) + extra_spp)
    z["ref_rays_count"] = np.int64(int(z["ref_rays_count"]) + st["rays"])
    z["ref_nan_samples"] = np.int64(int(z["ref_nan_samples"]) + st["nan_samples"])
    np.savez_compressed(path, **z)
    n = int(z["ref_spp"])
    m = z["ref_sum"] / n
    var = np.maximum(z["ref_sumsq"] / n - m * m, 0)
    print(f"converged {name}: now {n} spp; per-channel noise floor {np.sqrt(var.mean(axis=(0, 1)) / n)}", flush=True)


def make_converged(name):
    """tests/golden/aux/converged_<name>.npz: the reference's OWN high-spp render (its generate_ray + ray_color on all host threads)
    of a BASELINE.json config at thumbnail resolution: per-pixel sum and sum of squares.  The GPU test renders the same
    camera at 4x the samples and asserts the north-star criterion literally (per-channel RMSE < 1e-3, mean within 0.5 %)."""
    import importlib

    width, spp = CONVERGED[name]
    if name == "sponza16k":
        pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
        mesh = pkg.scene.sponza_standin_triangles(SPONZA_SMALL, 7)
        rs = R.RefScene("sponza", width=width, mesh=mesh)
    else:
        rs = R.RefScene(name, width=width)
    s, q, st = rs.render(spp, seed=20261105, threads=0, rng_mode=1)
    out = os.path.join(OUT, "aux")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, f"converged_{name}.npz")
    extra = dict(cam=np.frombuffer(rs.camera.to_bytes(), np.uint8), ref_sum=s, ref_sumsq=q, ref_spp=np.int64(spp), ref_rays_count=np.int64(st["rays"]),
                 ref_nan_samples=np.int64(st["nan_samples"]))
    if name == "sponza16k":
        np.savez_compressed(path, n_triangles=np.int64(SPONZA_SMALL), mesh_seed=np.int64(7), **extra)  # the scene is procedural (scene.sponza_standin_scene)
    else:
        np.savez_compressed(path, **extra)  # the scene is tests/golden/<name>.npz
    m = s / spp
    var = np.maximum(q / spp - m * m, 0)
    print(f"converged {name:10s} {rs.camera.image_width}x{rs.camera.image_height} {spp} spp in {st['seconds']:.0f} s: noise floor (RMSE of the reference mean) "
          f"{np.sqrt(var.mean() / spp):.2e}, mean {m.mean():.5f}", flush=True)


if __name__ == "__main__":
    if "--append-converged" in sys.argv:
        a = [x for x in sys.argv[1:] if not x.startswith("--")]
        append_converged(a[0], int(a[1]))
        sys.exit(0)
    if "--converged" in sys.argv:
        for n in ([a for a in sys.argv[1:] if not a.startswith("--")] or CONVERGED):
            make_converged(n)
        sys.exit(0)
    if "--write-color" in sys.argv:
        make_write_color()
        sys.exit(0)
    os.makedirs(OUT, exist_ok=True)
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    for n in (args or SCENES):
        if "--add-scatter" in sys.argv:
            add_scatter(n, SCENES[n][0])
        else:
            make(n, *SCENES[n])
