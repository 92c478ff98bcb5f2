#!/usr/bin/env python
"""bench.py — Mrays/s and Msamples/s of the path-tracing hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W]          our arm: libb200rt.so on N B200s
  python bench.py --impl reference ...                         the reference's own CPU code (oracle/_ref) on the host cores

Headline = BASELINE.json configs[1] (RTOW final scene: ~485 spheres, lens camera, 1200x675, 500 spp, depth 50).  A step = one
full frame (W*H*spp samples) through the wavefront kernels.  One process per GPU (torchrun for N > 1): the FIXED frame is split
by sample range (strong scaling): rank r renders samples distributed.sample_range(r, N, spp) of every pixel (Philox streams are
keyed by the global sample index, so the set of samples does not depend on N) and the fp64 accumulation buffers are summed
onto rank 0 with an NCCL reduce inside the timed region.  `per_scene` carries the other four configs (C1, C3, C4 at their full
size, C5 on a bounded slice), measured the same way.  At N > 1 the reduced image of a small frame is compared with rank 0's
own full-range render (`multi_gpu_parity`).  PyTorch is used for device buffers, streams/events and torch.distributed only.
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
OUT = sys.stdout  # replaced in __main__: the descriptor the single JSON line is written to

UNIT = "Mrays/s"
METRIC = "Mrays/s on RTOW final scene (1200x675, 500 spp, depth 50); Msamples/s alongside; other configs under per_scene"
# key -> (workload, spp timed per frame (0 = the config's own), note)
CONFIGS = {
    "rtow": ("C2 RTOW final scene: 487 spheres (lambertian/metal/dielectric), lens DOF camera, 1200x675, 500 spp, max depth 50", 0),
    "cornell": ("C1 Cornell box (quads + 2 boxes, diffuse_light) 600x600, 100 spp, max depth 50, perspective camera, quad-light sampling", 0),
    "rtnw": ("C3 RTNW final scene: constant_medium volumes, perlin noise, earthmap image texture, 800x800, 1000 spp, max depth 40", 0),
    "fox": ("C4 glTF Fox mesh (576 dielectric triangles), stand-in EXR environment, fisheye camera, 1920x1080, 1024 spp, max depth 5", 0),
    "sponza": ("C5 Sponza stand-in mesh (262 267 triangles + quad light) 3840x2160, max depth 5: a 128-spp slice of the 4096-spp frame", 128),
}
HEADLINE = "rtow"
EXTEND_BYTES_PER_RAY = 128  # k_extend_persistent: reads 7 of the 8 path quads (ray, time, box-filter constants, ids: 112 B), writes one HitRef (16 B)
# From the committed ncu --set full captures of k_extend_persistent (tools/ncu_summary.py); cold-cache single launches.
#   traffic: dram__bytes_read.sum + dram__bytes_write.sum per ray of one full 16 Mi-ray launch
#   issue_active: smsp__issue_active.avg.pct_of_peak_sustained_active / 100; threads_per_inst: smsp__thread_inst_executed_per_inst_executed.ratio
NCU = {
    # traffic from launch 0 (a full wave of camera rays: the ray count is known exactly), issue / lanes from launch 1 (the first secondary wave)
    "rtow": {"traffic_bytes_per_ray": 2279.28e6 / 16200000, "issue_active": 0.63, "threads_per_inst": 16.1, "source": "profiles/r02_as_extend_rtow_launch{0,1}_ncu_full.txt"},
    "rtnw": {"traffic_bytes_per_ray": 2347.15e6 / 16640000, "issue_active": 0.61, "threads_per_inst": 12.2, "source": "profiles/r02_as_extend_rtnw_launch{0,1}_ncu_full.txt (BVH4)"},
    "sponza": {"traffic_bytes_per_ray": 2340.10e6 / 16588800, "issue_active": 0.44, "threads_per_inst": 18.7, "source": "profiles/r02_as_extend_sponza_launch{0,1}_ncu_full.txt (BVH4, warp-uniform loop)"},
}


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


_pkg = None


def package():
    global _pkg
    if _pkg is None:
        _pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
    return _pkg


def load_config(key):
    """-> (scene, camera, spp per frame that is timed)."""
    pkg = package()
    if key == "sponza":
        sc = pkg.scene.sponza_standin_scene()  # the real Sponza.bin is absent from the reference checkout
        cam = pkg.camera.initialize_perspective(3840, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 4096, 5)
    else:
        z = np.load(os.path.join(ROOT, "cpu-ray-tracing-implementation_b200", "scenes", key + ".npz"))
        sc, cam = pkg.scene.Scene.load(z), pkg._abi.Camera.from_bytes(z["camera"].tobytes())
    return sc, cam, (CONFIGS[key][1] or cam.samples_per_pixel)


# ------------------------------------------------------------------------------------------------ CPU arms
def cpu_reference_run(key, spp, threads, scene=None, cam=None):
    """The reference's own code (oracle/_ref, compiled from its unmodified sources) where it can build the scene on this box
    (cornell, rtow: no asset files needed), else the oracle port (oracle/oracle.cc: the same algorithm and the same
    reference-order tree on the flattened scene).  Sponza: 480x270 (the x-sorted reference tree costs ~0.3 ms per ray per core).
    -> (kind, samples, rays, seconds, threads, what)"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refharness

    if key in ("cornell", "rtow") and refharness.available():
        rs = refharness.RefScene(key)  # same constructor calls and srand(1) as the committed scene file, config resolution
        _, _, st = rs.render(spp, seed=12345, threads=threads, rng_mode=1, want_sums=False)
        w, h = rs.camera.image_width, rs.camera.image_height
        rs.close()
        return "reference", st["samples"], st["rays"], st["seconds"], st["threads"], f"{w}x{h} at {spp} spp"
    import oracleport

    if scene is None:
        scene, cam, _ = load_config(key)
    if key == "sponza":
        cam = package().camera.initialize_perspective(480, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 1, 5)
    r = oracleport.render(scene, cam, 12345, 0, spp, threads=threads)
    return "port", r["stats"]["samples"], r["stats"]["rays"], r["stats"]["seconds"], threads, f"{cam.image_width}x{cam.image_height} at {spp} spp"


def cpu_baseline(key, budget_s, scene=None, cam=None):
    threads = os.cpu_count() or 1
    kind, s, r, secs, th, what = cpu_reference_run(key, 1, threads, scene, cam)  # 1 spp probe, then a sample sized to the budget
    spp = int(max(1, min(64, budget_s / max(secs, 1e-3))))
    if spp > 1:
        kind, s, r, secs, th, what = cpu_reference_run(key, spp, threads, scene, cam)
    return {"value": r / secs / 1e6, "unit": UNIT, "msamples_per_s": s / secs / 1e6, "cores": th, "kind": kind,
            "sample": f"same scene and camera, {what} ({s} samples, {r} rays) in {secs:.2f} s on {th} host threads"
                      + (" (reference ray_color/generate_ray; thread-local std::rand replacement so that rows run in parallel)" if kind == "reference" else " (oracle port)")}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    spp_step = 4
    for _ in range(args.warmup):
        cpu_reference_run(HEADLINE, spp_step, threads)
    tot_s = tot_r = 0
    t0 = time.perf_counter()
    secs_in = 0.0
    for _ in range(args.steps):
        kind, s, r, secs, th, what = cpu_reference_run(HEADLINE, spp_step, threads)
        tot_s, tot_r, secs_in = tot_s + s, tot_r + r, secs_in + secs
    wall = time.perf_counter() - t0
    v = tot_r / secs_in / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "msamples_per_s": tot_s / secs_in / 1e6, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs_in / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": CONFIGS[HEADLINE][0], "step": f"bounded sample: the full 1200x675 frame at {spp_step} spp per step (a rate: comparable with the 500-spp frame)"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": th, "kind": kind,
                             "sample": f"{args.steps} steps x {spp_step} spp of the 1200x675 frame, {th} threads, wall {wall:.1f} s"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), file=OUT, flush=True)


# ------------------------------------------------------------------------------------------------ our arm
class Arm:
    """One context per rank; measures frames of one config: device-resident (`value`) and end to end through the C ABI."""

    def __init__(self, args, torch, dist, world, rank, local):
        self.args, self.torch, self.dist, self.world, self.rank, self.local = args, torch, dist, world, rank, local
        self.pkg = package()
        self.abi = self.pkg._abi
        self.ctx = self.make_ctx()
        self.stream = torch.cuda.current_stream()

    def make_ctx(self):
        a, abi = self.args, self.abi
        ctx = self.pkg.backend.Context(self.local)
        ctx.set_option(abi.OPT_WAVE_PATHS, a.wave)
        for opt, val in ((abi.OPT_REFILL_LANES, a.refill), (abi.OPT_BVH_LEAF_SIZE, a.leaf), (abi.OPT_EXTEND_BLOCKS, a.extend_blocks), (abi.OPT_EXTEND_STAGES, a.stages),
                         (abi.OPT_PIPELINES, a.pipelines), (abi.OPT_PIPE_EXTEND_CTAS, a.pctas)):
            if val:
                ctx.set_option(opt, val)
        if a.no_fuse_miss:
            ctx.set_option(abi.OPT_FUSE_MISS, 0)
        if a.tail >= 0:
            ctx.set_option(abi.OPT_TAIL_PATHS, a.tail)
        if a.loop >= 0:
            ctx.set_option(abi.OPT_WAVE_LOOP, a.loop)
        return ctx

    def fence(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def sample_range(self, spp):
        d = self.pkg.distributed
        return d.weak_sample_range(self.rank, self.world, spp) if self.args.weak else d.sample_range(self.rank, self.world, spp)

    def measure(self, key, steps, warmup, want_e2e, want_clocks=False):
        torch, dist, abi, args = self.torch, self.dist, self.abi, self.args
        scene, cam, spp = load_config(key)
        if key == HEADLINE and args.spp:
            spp = args.spp
        H, W = cam.image_height, cam.image_width
        ctx = self.ctx
        ctx.upload_scene(scene)
        ctx.set_camera(cam)
        accum = torch.zeros(H * W * 3, dtype=torch.float64, device="cuda")
        lo, hi = self.sample_range(spp)

        def step(want_stats=False):
            accum.zero_()
            st = ctx.render_device(args.seed, lo, hi, accum.data_ptr(), self.stream.cuda_stream, want_stats=want_stats) if hi > lo else None
            self.pkg.distributed.reduce_sums(accum)  # NCCL over NVLink: the one exchange step of the path (no-op at N = 1)
            return st

        for _ in range(warmup):
            step()
        self.fence()
        sampler = ClockSampler(self.local) if (want_clocks and self.rank == 0) else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        self.fence()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        clocks = sampler.stop() if sampler else None
        # accounting from extra, instrumented steps: rays / samples / launches of one step as configured, then the extend kernel
        # timed alone (one pipeline, events around every k_extend launch: with two pipelines the halves' kernels overlap)
        st_all = step(want_stats=True) or {"samples": 0, "rays": 0, "kernel_launches": 0, "waves": 0}
        self.fence()
        ctx.set_option(abi.OPT_PIPELINES, 1)
        ctx.set_option(abi.OPT_WAVE_LOOP, 0)  # the host-driven loop is the one that can put events around single launches
        ctx.set_option(abi.OPT_WAVE_PATHS, max(1024, args.wave // 2))
        st1 = step(want_stats=True)
        self.fence()
        ctx.set_option(abi.OPT_PIPELINES, args.pipelines or 2)
        ctx.set_option(abi.OPT_WAVE_LOOP, args.loop if args.loop >= 0 else 1)
        ctx.set_option(abi.OPT_WAVE_PATHS, args.wave)
        counts = torch.tensor([st_all["samples"], st_all["rays"], st_all["kernel_launches"]], dtype=torch.float64, device="cuda")
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(counts, op=dist.ReduceOp.SUM)
        ms = float(ms.item())
        samples, rays, launches = (float(x) for x in counts.tolist())
        out = {"workload": CONFIGS[key][0], "value": rays * steps / (ms / 1e3) / 1e6, "unit": UNIT, "msamples_per_s": samples * steps / (ms / 1e3) / 1e6,
               "ms_per_step": ms / steps, "steps": steps, "spp_per_frame": spp * (self.world if args.weak else 1), "rays_per_sample": rays / max(1.0, samples),
               "samples_per_frame": samples, "rays_per_frame": rays, "gpu_launches_per_frame": launches, "waves_rank0": st_all["waves"]}
        if self.rank == 0 and st1 and st1["extend_ms"] > 0:
            peak, peak_src = peaks()
            achieved = st1["rays"] * EXTEND_BYTES_PER_RAY / (st1["extend_ms"] / 1e3) / 1e9
            ncu = NCU.get(key, {})
            roof = {"bound": "issue", "kernel": "k_extend_persistent", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": (ncu["traffic_bytes_per_ray"] * st1["rays"] / max(1, st1["waves"])) if ncu.get("traffic_bytes_per_ray") else None,
                    "peak_source": peak_src, "bytes_per_ray": EXTEND_BYTES_PER_RAY, "extend_share_of_step": st1["extend_ms"] / st1["gpu_ms"],
                    "algorithmic_bytes_per_launch": st1["rays"] * EXTEND_BYTES_PER_RAY / max(1, st1["waves"]), "launches": st1["waves"],
                    "measured": "k_extend_persistent launches of one extra single-pipeline step on rank 0, CUDA events around every launch on the launching stream",
                    "note": "achieved / peak / frac are the HBM figures the contract asks for (algorithmic 128 B per ray over the kernel's time vs the measured copy bandwidth); "
                            "the kernel is NOT HBM-bound: ncu shows DRAM at ~15 % while the SM issue slots and the lanes per instruction limit it "
                            "(fp32 box filter + binary64 reference-order leaf tests over an L1/L2-resident BVH), hence bound = issue"}
            if ncu.get("issue_active"):
                roof.update(issue_active=ncu["issue_active"], threads_per_inst=ncu["threads_per_inst"],
                            issue_roofline_frac=ncu["issue_active"] * ncu["threads_per_inst"] / 32.0, ncu_source=ncu["source"])
            out["roofline"] = roof
        if want_clocks:
            out["clocks"] = clocks
        if want_e2e:
            out["e2e"] = self.e2e(key, scene, cam, lo, hi, rays, samples, max(1, min(steps, 3)))
        del accum
        return out

    def e2e(self, key, scene, cam, lo, hi, rays, samples, n_steps):
        """Through the public C ABI with host buffers, every step: scene + camera hand-over (flatten -> BVH build -> H2D), render of
        this rank's sample range, NCCL reduce of the accumulators (N > 1), read-back of the image into pinned host memory (D2H)."""
        torch, dist = self.torch, self.dist
        H, W = cam.image_height, cam.image_width
        out = torch.empty((H, W, 3), dtype=torch.float64, pin_memory=True)
        ctx2 = self.make_ctx()
        accum = torch.zeros(H * W * 3, dtype=torch.float64, device="cuda") if self.world > 1 else None
        last = {}

        def e2e_step():
            ctx2.upload_scene(scene)
            ctx2.set_camera(cam)
            if self.world == 1:
                _, s2 = ctx2.render(self.args.seed, lo, hi, out=out.numpy())
                last.update(s2)
            else:
                accum.zero_()
                if hi > lo:
                    ctx2.render_device(self.args.seed, lo, hi, accum.data_ptr(), self.stream.cuda_stream)
                self.pkg.distributed.reduce_sums(accum)
                if self.rank == 0:
                    out.view(-1).copy_(accum, non_blocking=True)
                torch.cuda.synchronize()

        e2e_step()
        self.fence()
        t0 = time.perf_counter()
        for _ in range(n_steps):
            e2e_step()
        self.fence()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if self.world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        per = float(dt.item()) / n_steps
        ctx2.close()
        scene_bytes = float(sum(getattr(scene, n).nbytes for n in dir(scene) if isinstance(getattr(scene, n, None), np.ndarray)))
        return {"value": rays / per / 1e6, "unit": UNIT, "msamples_per_s": samples / per / 1e6, "ms_per_step": per * 1e3,
                "h2d_bytes_per_step": last.get("h2d_bytes", scene_bytes), "d2h_bytes_per_step": float(H * W * 3 * 8),
                "what": "b200rt_upload_scene + b200rt_set_camera + b200rt_render (N = 1) or b200rt_render_device + NCCL reduce + pinned read-back on rank 0 (N > 1), "
                        "host buffers, wall clock, max over ranks: BVH build, H2D, kernels, D2H inside"}

    def parity(self):
        """N > 1: the reduced image of a small frame equals rank 0's own render of the full sample range."""
        torch, pkg = self.torch, self.pkg
        scene, _, _ = load_config(HEADLINE)
        cam = pkg.camera.initialize_lens(240, 16.0 / 9.0, (13, 2, 3), (0, 0, 0), 0.6, 10, 20, 24, 50)
        self.ctx.upload_scene(scene)
        self.ctx.set_camera(cam)
        n = cam.image_height * cam.image_width * 3
        part = torch.zeros(n, dtype=torch.float64, device="cuda")
        lo, hi = pkg.distributed.sample_range(self.rank, self.world, 24)
        if hi > lo:
            self.ctx.render_device(self.args.seed, lo, hi, part.data_ptr(), self.stream.cuda_stream)
        pkg.distributed.reduce_sums(part)
        ok = None
        if self.rank == 0:
            whole = torch.zeros(n, dtype=torch.float64, device="cuda")
            self.ctx.render_device(self.args.seed, 0, 24, whole.data_ptr(), self.stream.cuda_stream)
            torch.cuda.synchronize()
            ok = bool(torch.allclose(part, whole, rtol=1e-12, atol=1e-12)) and bool(whole.abs().sum() > 0)
        self.fence()
        return ok


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
    arm = Arm(args, torch, dist, world, rank, local)
    head = arm.measure(HEADLINE, args.steps, args.warmup, want_e2e=True, want_clocks=True)
    per_scene = {}
    if not args.no_scenes:
        for key in CONFIGS:
            if key != HEADLINE:
                per_scene[key] = arm.measure(key, 2, 3, want_e2e=False)
    parity = arm.parity() if world > 1 else None
    if rank == 0:
        line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "msamples_per_s": head["msamples_per_s"], "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak" if args.weak else "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": CONFIGS[HEADLINE][0], "spp_per_frame": head["spp_per_frame"], "spp_per_gpu": head["spp_per_frame"] / world, "wave_paths": args.wave,
                           "pipelines": args.pipelines or 2, "seed": args.seed, "rays_per_sample": head["rays_per_sample"],
                           "l2": "inputs larger than L2: the wavefront state streamed per wave (16 Mi paths x 128 B = 2 GB per pipeline) exceeds the 126 MB L2; the scene (~100 kB) is meant to stay cached",
                           "parallelism": f"the fixed frame split by sample range x{world}, one NCCL reduce of the fp64 accumulators per frame" if world > 1 else "single GPU"},
                "roofline": head.get("roofline"), "e2e": head["e2e"], "gpu_launches": int(head["gpu_launches_per_frame"] * args.steps), "clocks": head.get("clocks"),
                "per_scene": per_scene}
        if world > 1:
            line["multi_gpu_parity"] = parity
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(HEADLINE, 12.0)
            for key, d in per_scene.items():
                d["cpu_baseline"] = cpu_baseline(key, 4.0)
        print(json.dumps(line), file=OUT, flush=True)
    arm.ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main_stdout_only_json():
    """Libraries write to stdout too (NCCL prints its version line there): fd 1 is pointed at stderr for the whole run and the
    JSON line goes to the real stdout, so that stdout carries exactly ONE line."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(real, "w")


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
    ap.add_argument("--tail", type=int, default=-1, help="tuning: B200RT_OPT_TAIL_PATHS (-1 = library default)")
    ap.add_argument("--loop", type=int, default=-1, help="tuning: B200RT_OPT_WAVE_LOOP, 0 = host-driven wave loop, 1 = device-driven CUDA graph (-1 = library default)")
    ap.add_argument("--weak", action="store_true", help="weak scaling instead: every rank renders the config's full spp (secondary figure)")
    ap.add_argument("--no-scenes", action="store_true", help="skip the per_scene measurements (C1, C3, C4, C5 slice)")
    a = ap.parse_args()
    OUT = main_stdout_only_json()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
