#!/usr/bin/env python
"""Render one of the BASELINE.json configs to a P3 .ppm exactly as camera::render would write it (src/camera.h:149-151, 174),
through the C ABI: upload -> render -> b200rt_format_ppm (text made on the device) -> file.

  python tools/render.py --scene cornell [--spp 100] [--width 600] [--out cornell.ppm] [--seed 24301]
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    from bench_scenes import CONFIGS, load, pkg

    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="cornell", choices=sorted(CONFIGS))
    ap.add_argument("--spp", type=int, default=0, help="samples per pixel (default: the config's)")
    ap.add_argument("--out", default="")
    ap.add_argument("--seed", type=int, default=0x5EED)
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args()
    sc, cam = load(a.scene)
    spp = a.spp or cam.samples_per_pixel
    out = a.out or f"{a.scene}.ppm"
    with pkg.backend.Context(a.device) as ctx:
        t0 = time.perf_counter()
        ctx.upload_scene(sc)
        ctx.set_camera(cam)
        rgb_sum, st = ctx.render(a.seed, 0, spp)
        t1 = time.perf_counter()
        text = ctx.format_ppm(rgb_sum, spp)
        t2 = time.perf_counter()
    with open(out, "wb") as f:
        f.write(text)
    print(f"{CONFIGS[a.scene]}: {st['samples']} samples, {st['rays']} rays in {t1 - t0:.3f} s ({st['rays'] / st['gpu_ms'] / 1e3:.0f} Mrays/s on the device); "
          f"{len(text)} bytes of P3 text in {t2 - t1:.3f} s -> {out}")


if __name__ == "__main__":
    main()
