#!/usr/bin/env python
"""Sweep the tuning options of the wavefront (B200RT_OPT_*) on one scene, one process, CUDA-event timing.

  python tools/tune.py [--scene rtow] [--spp 32] [--wave 22,23] [--refill 32,24,16] [--blocks 4,5,6,8] [--leaf 2]

Prints one JSON line per combination (Mrays/s = rays of the step / device time of the step, best of --reps).
"""
import argparse
import importlib
import itertools
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = pkg._abi


def ints(s):
    return [int(x) for x in s.replace("m", "-").split(",") if x]


def main():
    import torch
    from bench_scenes import load

    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="rtow")
    ap.add_argument("--spp", type=int, default=32)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--wave", default="24")
    ap.add_argument("--refill", default="0")
    ap.add_argument("--blocks", default="4")
    ap.add_argument("--leaf", default="0")
    ap.add_argument("--stages", default="1", help="ray batches staged per warp: 1, 2 or 1,2")
    ap.add_argument("--carveout", default="-1", help="shared-memory carve-out percent of the extend kernel (-1 = automatic)")
    ap.add_argument("--pipelines", default="1", help="1, 2 or 1,2: independent wavefronts on two streams")
    ap.add_argument("--pctas", default="3", help="extend CTAs per SM in two-pipeline mode")
    ap.add_argument("--width", type=int, default=0, help="BVH width: 0 = automatic, 2, 4")
    ap.add_argument("--spec", default="0", help="speculative (postponed-leaf) traversal: 0, 1 or 0,1")
    a = ap.parse_args()
    scene, cam = load(a.scene)
    H, W = cam.image_height, cam.image_width
    accum = torch.zeros(H * W * 3, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for leaf in ints(a.leaf):
        ctx = pkg.backend.Context(0)
        ctx.set_option(abi.OPT_BVH_LEAF_SIZE, leaf)
        ctx.set_option(abi.OPT_BVH_WIDTH, a.width)
        ctx.upload_scene(scene)
        ctx.set_camera(cam)
        for wave, refill, blocks, spec, stages, carve, pipes, pctas in itertools.product(ints(a.wave), ints(a.refill), ints(a.blocks), ints(a.spec), ints(a.stages),
                                                                                         ints(a.carveout), ints(a.pipelines), ints(a.pctas)):
            if pipes == 1 and pctas != ints(a.pctas)[0]:
                continue
            ctx.set_option(abi.OPT_PIPELINES, pipes)
            ctx.set_option(abi.OPT_PIPE_EXTEND_CTAS, pctas)
            ctx.set_option(abi.OPT_SPECULATIVE_LEAF, spec)
            ctx.set_option(abi.OPT_EXTEND_STAGES, stages)
            ctx.set_option(abi.OPT_SMEM_CARVEOUT, carve)
            ctx.set_option(abi.OPT_WAVE_PATHS, 1 << wave)
            ctx.set_option(abi.OPT_REFILL_LANES, refill)
            ctx.set_option(abi.OPT_EXTEND_BLOCKS, blocks)
            best = None
            for _ in range(a.reps + 1):  # first one is the warm-up
                accum.zero_()
                st = ctx.render_device(0x5EED, 0, a.spp, accum.data_ptr(), stream, want_stats=True)
                if best is None or st["gpu_ms"] < best["gpu_ms"]:
                    best = st
            print(json.dumps({"scene": a.scene, "spp": a.spp, "leaf": leaf, "wave_log2": wave, "refill": refill, "blocks": blocks, "spec": spec, "width": a.width, "pipelines": pipes, "pctas": pctas, "stages": stages, "carveout": carve,
                              "mrays_per_s": best["rays"] / best["gpu_ms"] / 1e3, "gpu_ms": best["gpu_ms"], "extend_ms": best["extend_ms"],
                              "waves": best["waves"]}), flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
