#!/usr/bin/env python
"""North-star image criterion, measured literally: per-channel RMSE and image-mean ratio of the GPU render against the
reference's OWN render (oracle/_ref, the unmodified reference code) of the same scene, both at high spp.

The reference side is the bottleneck (about 5 Msamples/s on 16 host threads), so the frame is small and the reference spp is
what the time budget allows; the RMSE reported is therefore an upper bound dominated by the REFERENCE's residual noise,
sqrt(var/N_ref), which is printed beside it.  The GPU render uses 8x more samples.

  python tools/converged_rmse.py [--scenes cornell:16:600000,rtow:32:100000]      (name:width:reference spp)

One JSON line per scene.  Test infrastructure: uses oracle/_ref as the checker, like tests/ do.
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
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", default="cornell:16:600000,rtow:32:100000,three_ball_lens:32:100000")
    ap.add_argument("--gpu-mult", type=int, default=8)
    a = ap.parse_args()
    pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
    import refharness

    ctx = pkg.backend.Context(0)
    for item in a.scenes.split(","):
        name, width, ref_spp = item.split(":")
        rs = refharness.RefScene(name, width=int(width))
        H, W = rs.camera.image_height, rs.camera.image_width
        t0 = time.perf_counter()
        s_ref, q_ref, st = rs.render(int(ref_spp), seed=4242, threads=0, rng_mode=1)
        t_ref = time.perf_counter() - t0
        n_ref = int(ref_spp)
        ctx.upload_scene(rs.scene)
        ctx.set_camera(rs.camera)
        n_gpu = a.gpu_mult * n_ref
        t0 = time.perf_counter()
        s_gpu, gst = ctx.render(777, 0, n_gpu)
        t_gpu = time.perf_counter() - t0
        fin = np.isfinite(s_gpu) & np.isfinite(s_ref)
        m_ref = np.where(fin, s_ref / n_ref, 0)
        m_gpu = np.where(fin, s_gpu / n_gpu, 0)
        var_ref = np.maximum(q_ref / n_ref - (s_ref / n_ref) ** 2, 0)
        d = (m_gpu - m_ref).reshape(-1, 3)
        rmse = np.sqrt((d ** 2).mean(axis=0))
        floor = np.sqrt((var_ref.reshape(-1, 3) / n_ref * (1 + 1.0 / a.gpu_mult)).mean(axis=0))  # expected RMSE from MC noise alone
        print(json.dumps({"scene": name, "frame": f"{W}x{H}", "ref_spp": n_ref, "gpu_spp": n_gpu,
                          "rmse_rgb": [float(x) for x in rmse], "mc_noise_floor_rgb": [float(x) for x in floor],
                          "rmse_over_floor": [float(x / max(y, 1e-300)) for x, y in zip(rmse, floor)],
                          "mean_ref": float(m_ref.mean()), "mean_gpu": float(m_gpu.mean()), "mean_ratio": float(m_gpu.mean() / m_ref.mean()),
                          "ref_nan_samples": int(st["nan_samples"]), "nonfinite_pixels_channels": int((~fin).sum()),
                          "ref_seconds": round(t_ref, 2), "ref_threads": int(st["threads"]), "gpu_seconds": round(t_gpu, 3)}), flush=True)
        rs.close()
    ctx.close()


if __name__ == "__main__":
    main()
