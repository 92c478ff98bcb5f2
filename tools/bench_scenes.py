#!/usr/bin/env python
"""Mrays/s and Msamples/s of every BASELINE.json config on one B200 (bench.py measures configs[1] under the driver's
contract; this reports all five at full resolution on a bounded spp slice: per-sample cost is stationary).

  python tools/bench_scenes.py [--spp N] [--scenes cornell,rtow,rtnw,fox,sponza] [--cpu]

--cpu adds the reference's own CPU render (oracle/_ref) of the same scene at reduced spp on all host threads.
Sponza is the deterministic stand-in mesh (scene.sponza_standin_scene): the real Sponza.bin is absent from the reference
checkout.  The Fox environment (bathroom.exr) is absent too: the committed fox scene carries the stand-in used by the fixtures.
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = pkg._abi

CONFIGS = {
    "cornell": "C1 Cornell box 600x600, 100 spp, depth 50, perspective, quad light",
    "rtow": "C2 RTOW final 1200x675, 500 spp, depth 50, lens",
    "rtnw": "C3 RTNW final 800x800, 1000 spp, depth 40, volumes + perlin + earthmap",
    "fox": "C4 glTF Fox 1920x1080, 1024 spp, depth 5, fisheye, environment",
    "sponza": "C5 Sponza-standin 262,267 tris 3840x2160, 4096 spp, depth 5",
}


def load(name):
    if name == "sponza":
        sc = pkg.scene.sponza_standin_scene()
        cam = pkg.camera.initialize_perspective(3840, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 4096, 5)
        return sc, cam
    z = np.load(os.path.join(ROOT, "cpu-ray-tracing-implementation_b200", "scenes", name + ".npz"))
    return pkg.scene.Scene.load(z), abi.Camera.from_bytes(z["camera"].tobytes())


def cpu_baseline(name, sc, cam):
    """The reference's CPU code on this box's host threads, beside the GPU number (a reported baseline, not a target).
    cornell / rtow: the reference itself (oracle/_ref: its own ray_color / generate_ray on objects built by its constructors).
    rtnw / fox / sponza need assets that only exist next to the reference checkout, so there the oracle port (oracle/oracle.cc,
    same algorithm on the flattened scene, same reference-order tree) is timed instead; sponza at 480x270 (its x-sorted
    reference tree costs about a millisecond per ray on one core)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracleport
    import refharness

    threads = os.cpu_count() or 1
    if name in ("cornell", "rtow") and refharness.available():
        rs = refharness.RefScene(name, width=cam.image_width, spp=cam.samples_per_pixel, depth=cam.max_depth)
        _, _, st = rs.render(2, seed=12345, threads=threads, rng_mode=1, want_sums=False)
        rs.close()
        return {"kind": "reference", "cores": st["threads"], "mrays_per_s": st["rays"] / st["seconds"] / 1e6, "msamples_per_s": st["samples"] / st["seconds"] / 1e6,
                "sample": f"{cam.image_width}x{cam.image_height} at 2 spp in {st['seconds']:.2f} s"}
    if name == "sponza":
        cam = pkg.camera.initialize_perspective(480, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 1, 5)
    r = oracleport.render(sc, cam, 12345, 0, 1, threads=threads)
    st = r["stats"]
    return {"kind": "port", "cores": threads, "mrays_per_s": st["rays"] / st["seconds"] / 1e6, "msamples_per_s": st["samples"] / st["seconds"] / 1e6,
            "sample": f"{cam.image_width}x{cam.image_height} at 1 spp in {st['seconds']:.2f} s"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=0, help="spp slice per scene (default: sized to ~0.5 G samples, at most the config's)")
    ap.add_argument("--scenes", default="cornell,rtow,rtnw,fox,sponza")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--full", action="store_true", help="time every config at its full spp (C5 alone is 34 G samples: about a minute on one B200)")
    ap.add_argument("--cpu", action="store_true", help="add the CPU baseline of every scene (see cpu_baseline)")
    ap.add_argument("--builder", type=int, default=0, help="0 = host SAH BVH (default), 1 = device LBVH (csrc/bvh_gpu.cuh), 2 = device PLOC + host finishing pass (csrc/bvh_ploc.cuh)")
    ap.add_argument("--width", type=int, default=0, help="BVH width: 0 = automatic, 2, 4")
    ap.add_argument("--pipelines", type=int, default=2, help="independent wavefronts on two streams (1 also reports the extend share)")
    ap.add_argument("--tail", type=int, default=-1, help="B200RT_OPT_TAIL_PATHS (-1 = library default, 0 = off)")
    ap.add_argument("--opt", action="append", default=[], help="any b200rt_option as NUMBER=VALUE (repeatable), e.g. --opt 17=1 --opt 18=8")
    ap.add_argument("--mirror", action="store_true", help="also time the node-for-node mirror traversal (parity arbiter)")
    args = ap.parse_args()
    ctx = pkg.backend.Context(0)
    ctx.set_option(abi.OPT_BVH_BUILDER, args.builder)
    ctx.set_option(abi.OPT_BVH_WIDTH, args.width)
    ctx.set_option(abi.OPT_PIPELINES, args.pipelines)
    if args.tail >= 0:
        ctx.set_option(abi.OPT_TAIL_PATHS, args.tail)
    for kv in args.opt:
        k, v = kv.split("=")
        ctx.set_option(int(k), float(v))
    for name in args.scenes.split(","):
        t0 = time.perf_counter()
        sc, cam = load(name)
        t_load = time.perf_counter() - t0
        npix = cam.image_width * cam.image_height
        spp = cam.samples_per_pixel if args.full else (args.spp or max(1, min(cam.samples_per_pixel, int(5e8 // npix))))
        t0 = time.perf_counter()
        ctx.upload_scene(sc)
        t_up = time.perf_counter() - t0
        ctx.set_camera(cam)
        info = ctx.bvh_info()
        ctx.render(1, 0, max(1, spp // 8))  # warm-up (allocates the wave pools)
        best = None
        for _ in range(args.reps):
            _, st = ctx.render(1, 0, spp)
            if best is None or st["gpu_ms"] < best["gpu_ms"]:
                best = st
        line = {"scene": name, "opts": args.opt, "builder": ["host SAH", "device LBVH", "device PLOC"][args.builder], "config": CONFIGS[name], "width": cam.image_width, "height": cam.image_height, "spp_timed": spp,
                "spp_config": cam.samples_per_pixel, "max_depth": cam.max_depth, "mrays_per_s": best["rays"] / best["gpu_ms"] / 1e3,
                "msamples_per_s": best["samples"] / best["gpu_ms"] / 1e3, "rays_per_sample": best["rays"] / best["samples"],
                "gpu_ms": best["gpu_ms"], "extend_share": (best["extend_ms"] / best["gpu_ms"]) if best["extend_ms"] > 0 else None, "waves": best["waves"],
                "kernel_launches": best["kernel_launches"], "nan_samples": best["nan_samples"],
                "bvh": {k: info[k] for k in ("n_nodes", "n_leaves", "n_prims", "max_depth", "build_ms", "sah_cost")},
                "scene_load_s": t_load, "upload_s": t_up,
                "full_config_seconds_at_this_rate": cam.samples_per_pixel * npix / (best["samples"] / best["gpu_ms"] * 1e3)}
        if args.mirror:
            ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_MIRROR)
            _, sm = ctx.render(1, 0, max(1, spp // 16))
            ctx.set_option(abi.OPT_TRAVERSAL, abi.TRAVERSAL_FAST)
            line["mirror_mrays_per_s"] = sm["rays"] / sm["gpu_ms"] / 1e3
        if args.cpu:
            line["cpu_baseline"] = cpu_baseline(name, sc, cam)
        print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
