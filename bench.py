#!/usr/bin/env python
"""bench.py — Mrays/s and Msamples/s of the path-tracing hot path on BASELINE.json configs[1]
(RTOW final scene: ~485 spheres, lens camera, 1200x675, 500 spp, depth 50).

  python bench.py [--gpus N] [--steps K] [--warmup W]          our arm: libb200rt.so on N B200s
  python bench.py --impl reference ...                         the reference's own CPU code (oracle/_ref) on the host cores

A step = one full frame (W*H*spp samples) through the wavefront kernels.  One process per GPU (torchrun for N > 1):
rank r renders samples [r*spp, (r+1)*spp) of every pixel (weak scaling: per-GPU work fixed, decorrelated Philox
streams by global sample index) and the fp64 accumulation buffers are summed onto rank 0 with an NCCL reduce.
PyTorch is used for device buffers, streams/events and torch.distributed only.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENE = "rtow"
WORKLOAD = "RTOW final scene: 487 spheres (lambertian/metal/dielectric), lens DOF camera, 1200x675, 500 spp, max depth 50"
METRIC, UNIT = "Mrays/s on RTOW final scene (1200x675, 500 spp, depth 50); Msamples/s alongside", "Mrays/s"
TRAFFIC_BYTES_PER_RAY = 2415.63e6 / 16777216  # profiles/r01_p_extend_ncu_full.txt: dram__bytes_read.sum 2150.4 MB + dram__bytes_write.sum 265.2 MB of one full 16 Mi-ray launch = 144 B/ray (q7 shares q6's 32-byte sector)
EXTEND_BYTES_PER_RAY = 128  # k_extend_persistent: reads 7 of the 8 path quads (ray, time, box-filter constants, ids: 112 B), writes one HitRef (16 B)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 6 for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def load_scene():
    pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
    z = np.load(os.path.join(ROOT, "cpu-ray-tracing-implementation_b200", "scenes", SCENE + ".npz"))
    return pkg, pkg.scene.Scene.load(z), pkg._abi.Camera.from_bytes(z["camera"].tobytes())


# ------------------------------------------------------------------------------------------------ CPU arms
def cpu_reference_run(width, spp, threads):
    """The reference's own code (oracle/_ref, compiled from its unmodified sources) or, without it, the oracle port."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refharness

    if refharness.available():
        rs = refharness.RefScene(SCENE, width=width)  # same constructor calls and srand(1) as the committed scene file
        _, _, st = rs.render(spp, seed=12345, threads=threads, rng_mode=1, want_sums=False)
        return "reference", st["samples"], st["rays"], st["seconds"], st["threads"]
    import oracleport

    _, scene, cam = load_scene()
    if width != cam.image_width:
        raise RuntimeError("oracle port runs at the config resolution only")
    r = oracleport.render(scene, cam, 12345, 0, spp, threads=threads)
    return "port", r["stats"]["samples"], r["stats"]["rays"], r["stats"]["seconds"], threads


def cpu_baseline(budget_s=12.0):
    threads = os.cpu_count() or 1
    kind, s, r, secs, th = cpu_reference_run(1200, 1, threads)  # 1 spp probe, then a sample sized to the budget
    spp = int(max(1, min(64, budget_s / max(secs, 1e-3))))
    kind, s, r, secs, th = cpu_reference_run(1200, spp, threads)
    return {"value": r / secs / 1e6, "unit": UNIT, "msamples_per_s": s / secs / 1e6, "cores": th, "kind": kind,
            "sample": f"same scene and camera, 1200x675 at {spp} spp ({s} samples, {r} rays) in {secs:.2f} s on {th} host threads "
                      f"(reference ray_color/generate_ray; thread-local std::rand replacement so that rows run in parallel)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    spp_step = 4
    for _ in range(args.warmup):
        cpu_reference_run(1200, spp_step, threads)
    tot_s = tot_r = 0
    t0 = time.perf_counter()
    secs_in = 0.0
    for _ in range(args.steps):
        kind, s, r, secs, th = cpu_reference_run(1200, spp_step, threads)
        tot_s, tot_r, secs_in = tot_s + s, tot_r + r, secs_in + secs
    wall = time.perf_counter() - t0
    v = tot_r / secs_in / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "msamples_per_s": tot_s / secs_in / 1e6, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs_in / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "step": f"bounded sample: the full 1200x675 frame at {spp_step} spp per step"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": th, "kind": kind,
                             "sample": f"{args.steps} steps x {spp_step} spp of the 1200x675 frame, {th} threads, wall {wall:.1f} s"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this backend has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg, scene, cam = load_scene()
    abi = pkg._abi
    spp = args.spp or cam.samples_per_pixel
    H, W = cam.image_height, cam.image_width
    ctx = pkg.backend.Context(local)
    ctx.set_option(abi.OPT_WAVE_PATHS, args.wave)
    if args.refill:
        ctx.set_option(abi.OPT_REFILL_LANES, args.refill)
    if args.leaf:
        ctx.set_option(abi.OPT_BVH_LEAF_SIZE, args.leaf)
    if args.extend_blocks:
        ctx.set_option(abi.OPT_EXTEND_BLOCKS, args.extend_blocks)
    if args.stages:
        ctx.set_option(abi.OPT_EXTEND_STAGES, args.stages)
    if args.no_fuse_miss:
        ctx.set_option(abi.OPT_FUSE_MISS, 0)
    if args.pipelines:
        ctx.set_option(abi.OPT_PIPELINES, args.pipelines)
    if args.pctas:
        ctx.set_option(abi.OPT_PIPE_EXTEND_CTAS, args.pctas)
    ctx.upload_scene(scene)
    ctx.set_camera(cam)
    accum = torch.zeros(H * W * 3, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream()
    lo, hi = pkg.distributed.weak_sample_range(rank, world, spp)  # disjoint global sample ranges

    def step(want_stats=False):
        accum.zero_()
        st = ctx.render_device(args.seed, lo, hi, accum.data_ptr(), stream.cuda_stream, want_stats=want_stats)
        pkg.distributed.reduce_sums(accum)  # NCCL over NVLink: the one exchange step of the path (no-op at N = 1)
        return st

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    fence()
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    fence()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    clocks = sampler.stop() if sampler else None
    # per-kernel accounting from one extra, instrumented step (events around every k_extend launch)
    st_all = step(want_stats=True)  # as configured: rays, samples, launches of one timed step
    fence()
    ctx.set_option(abi.OPT_PIPELINES, 1)  # then the extend kernel timed alone: with two pipelines the halves' kernels overlap
    ctx.set_option(abi.OPT_WAVE_PATHS, max(1024, args.wave // 2))
    st = step(want_stats=True)
    fence()
    st["kernel_launches"] = st_all["kernel_launches"]
    counts = torch.tensor([st["samples"], st["rays"], st["kernel_launches"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    ms = float(ms.item())
    samples, rays, launches = (float(x) for x in counts.tolist())

    # end to end through the C ABI with host buffers: scene + camera upload (H2D), render, image read-back (D2H), every step
    e2e = None
    if rank == 0 or world > 1:
        out = torch.empty((H, W, 3), dtype=torch.float64, pin_memory=True).numpy()
        ctx2 = pkg.backend.Context(local)
        ctx2.set_option(abi.OPT_WAVE_PATHS, args.wave)
        if args.pipelines:
            ctx2.set_option(abi.OPT_PIPELINES, args.pipelines)
        if args.pctas:
            ctx2.set_option(abi.OPT_PIPE_EXTEND_CTAS, args.pctas)
        if args.stages:
            ctx2.set_option(abi.OPT_EXTEND_STAGES, args.stages)

        def e2e_step():
            ctx2.upload_scene(scene)
            ctx2.set_camera(cam)
            _, s2 = ctx2.render(args.seed, lo, hi, out=out)
            return s2

        e2e_step()
        fence()
        t0 = time.perf_counter()
        n_e2e = max(1, min(args.steps, 3))
        for _ in range(n_e2e):
            s2 = e2e_step()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": rays / (float(dt.item()) / n_e2e) / 1e6, "unit": UNIT, "msamples_per_s": samples / (float(dt.item()) / n_e2e) / 1e6,
               "h2d_bytes_per_step": s2["h2d_bytes"], "d2h_bytes_per_step": s2["d2h_bytes"],
               "what": "b200rt_upload_scene + b200rt_set_camera + b200rt_render with host buffers (BVH build, H2D, kernels, D2H), wall clock"}
        ctx2.close()
    if rank == 0:
        peak, peak_src = peaks()
        ext_s = st["extend_ms"] / 1e3
        achieved = st["rays"] * EXTEND_BYTES_PER_RAY / ext_s / 1e9
        line = {"metric": METRIC, "value": rays * args.steps / (ms / 1e3) / 1e6, "unit": UNIT, "msamples_per_s": samples * args.steps / (ms / 1e3) / 1e6,
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": WORKLOAD, "spp_per_gpu": spp, "wave_paths": args.wave, "pipelines": args.pipelines or 2, "seed": args.seed, "rays_per_sample": rays / samples,
                           "l2": "inputs larger than L2: the wavefront state streamed per wave (16 Mi paths x 128 B = 2 GB per pipeline) exceeds the 126 MB L2; the scene (~100 kB) is meant to stay cached",
                           "parallelism": f"sample-range split x{world}, NCCL reduce of the fp64 accumulators" if world > 1 else "single GPU"},
                "roofline": {"bound": "hbm", "kernel": "k_extend_persistent", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": TRAFFIC_BYTES_PER_RAY * st["rays"] / max(1, st["waves"]), "peak_source": peak_src, "bytes_per_ray": EXTEND_BYTES_PER_RAY,
                             "extend_share_of_step": st["extend_ms"] / st["gpu_ms"],
                             "algorithmic_bytes_per_launch": st["rays"] * EXTEND_BYTES_PER_RAY / max(1, st["waves"]), "launches": st["waves"],
                             "measured": "extend launches of one extra single-pipeline step (16 Mi-path waves), CUDA events around every launch", "note": "issue/latency bound, not HBM bound (fp32 box filter + binary64 reference-order leaf tests over an L1/L2-resident BVH); "
                                     "traffic = dram read+write per ray of a full launch from profiles/r01_p_extend_ncu_full.txt (ncu --set full) x mean rays per launch; see DESIGN.md"},
                "e2e": e2e, "gpu_launches": int(launches * args.steps), "clocks": clocks}
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp", type=int, default=0, help="samples per pixel per step (default: the config's 500)")
    ap.add_argument("--wave", type=int, default=1 << 25, help="paths resident per wave, shared by the pipelines")
    ap.add_argument("--seed", type=int, default=0x5EED)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--refill", type=int, default=0, help="tuning: idle lanes per warp that trigger a ray fetch in the persistent extend kernel")
    ap.add_argument("--stages", type=int, default=0, help="tuning: ray batches staged per warp in the extend kernel (1 or 2)")
    ap.add_argument("--extend-blocks", type=int, default=0, help="tuning: CTAs/SM of the persistent extend kernel (4, 5, 6, 8)")
    ap.add_argument("--pipelines", type=int, default=0, help="tuning: 1 or 2 independent wavefronts (two streams)")
    ap.add_argument("--pctas", type=int, default=0, help="tuning: extend CTAs per SM in two-pipeline mode")
    ap.add_argument("--no-fuse-miss", action="store_true", help="tuning: send missing rays through the miss queue and its shade launch")
    ap.add_argument("--leaf", type=int, default=0, help="tuning: max primitives per BVH leaf")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
