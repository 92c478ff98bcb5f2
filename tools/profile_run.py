#!/usr/bin/env python
"""A minimal program to put under ncu: ONE frame of one config through the plain (host-driven, one-pipeline) wave loop, so that
launch k of k_extend_persistent is bounce k of a full wave.

  ncu --set full --clock-control none --import-source on -k regex:k_extend_persistent -s 0 -c 3 -o gpurun_out/prof \
      python tools/profile_run.py --scene rtow --spp 20
"""
import argparse
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = pkg._abi

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="rtow")
ap.add_argument("--spp", type=int, default=20)
ap.add_argument("--width", type=int, default=0, help="BVH width override (0 = automatic)")
ap.add_argument("--loop", type=int, default=0)
ap.add_argument("--pipelines", type=int, default=1)
a = ap.parse_args()
scene, cam, _ = bench.load_config(a.scene)
with pkg.backend.Context(0) as ctx:
    ctx.set_option(abi.OPT_PIPELINES, a.pipelines)
    try:
        ctx.set_option(abi.OPT_WAVE_LOOP, a.loop)
    except pkg.backend.B200RTError:
        pass  # an older build under B200RT_LIB (A/B runs): it only has the host-driven loop
    ctx.set_option(abi.OPT_BVH_WIDTH, a.width)
    ctx.upload_scene(scene)
    ctx.set_camera(cam)
    img, st = ctx.render(1, 0, a.spp)
    print({k: st[k] for k in ("samples", "rays", "waves", "kernel_launches", "gpu_ms", "extend_ms")})
