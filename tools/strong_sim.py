#!/usr/bin/env python
"""What one rank of an N-GPU strong-scaling run does, timed on ONE GPU: the RTOW frame (1200x675, 500 spp) cut into N sample
ranges, rank 0's range rendered alone.  Ranks are independent until the final reduce, so t(500 / N spp) x N / t(500 spp) is the
scaling loss to expect from short frames (fixed per-wave costs, the depth-50 tail), before any NCCL time.

  python tools/strong_sim.py [--scene rtow] [--spp 500] [--loops 0,1] [--tails 32768] [--reps 5]
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
abi = pkg._abi


def main():
    import torch

    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="rtow")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--ns", default="1,2,4,8")
    ap.add_argument("--loops", default="0,1")
    ap.add_argument("--tails", default="32768")
    ap.add_argument("--waves", default=str(1 << 25))
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    scene, cam, spp = bench.load_config(a.scene)
    spp = a.spp or spp
    ctx = pkg.backend.Context(0)
    ctx.upload_scene(scene)
    ctx.set_camera(cam)
    accum = torch.zeros(cam.image_height * cam.image_width * 3, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream()
    for loop in (int(x) for x in a.loops.split(",")):
        for tail in (int(x) for x in a.tails.split(",")):
            for wave in (int(x) for x in a.waves.split(",")):
                ctx.set_option(abi.OPT_WAVE_LOOP, loop)
                ctx.set_option(abi.OPT_TAIL_PATHS, tail)
                ctx.set_option(abi.OPT_WAVE_PATHS, wave)
                base = None
                for n in (int(x) for x in a.ns.split(",")):
                    lo, hi = pkg.distributed.sample_range(0, n, spp)
                    for _ in range(2):
                        ctx.render_device(1, lo, hi, accum.data_ptr(), stream.cuda_stream)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(a.reps):
                        accum.zero_()
                        ctx.render_device(1, lo, hi, accum.data_ptr(), stream.cuda_stream)
                    e1.record()
                    torch.cuda.synchronize()
                    ms = e0.elapsed_time(e1) / a.reps
                    st = ctx.render_device(1, lo, hi, accum.data_ptr(), stream.cuda_stream, want_stats=True)
                    base = base or ms
                    print(json.dumps({"scene": a.scene, "loop": "graph" if loop else "host", "tail_paths": tail, "wave_paths": wave, "n": n, "spp_rank": hi - lo, "ms": round(ms, 3),
                                      "mrays_per_s": round(st["rays"] / ms / 1e3, 1), "waves": st["waves"], "launches": st["kernel_launches"],
                                      "strong_efficiency_estimate": round(base / (n * ms), 4)}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
