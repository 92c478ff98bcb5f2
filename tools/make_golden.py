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


def make(name, width, spp):
    rng = np.random.RandomState(abs(hash(name)) % (2**31))
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


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for n in (sys.argv[1:] or SCENES):
        make(n, *SCENES[n])
