#!/usr/bin/env python
"""BASELINE.json configs[4]: the Sponza-sized mesh at 3840x2160 split by SAMPLE RANGE across the GPUs of one box, fp64
accumulators combined with one NCCL reduce.  Strong scaling: the total spp is fixed, rank r of N renders
`distributed.sample_range(r, N, spp)`.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_sponza_multi.py [--spp 64]

The mesh is the deterministic stand-in of scene.sponza_standin_scene() (the real Sponza.bin is absent from the reference
checkout); the full config is 4096 spp, a bounded slice is timed (per-sample cost is stationary).  One JSON line on rank 0.
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=64, help="total samples per pixel over all ranks")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--builder", type=int, default=0)
    a = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("cpu-ray-tracing-implementation_b200")
    sc = pkg.scene.sponza_standin_scene()
    cam = pkg.camera.initialize_perspective(3840, 16.0 / 9.0, (500, 320, 90), (0, 280, 0), 1, 45.0, 4096, 5)
    ctx = pkg.backend.Context(local)
    ctx.set_option(pkg._abi.OPT_BVH_BUILDER, a.builder)
    ctx.upload_scene(sc)
    ctx.set_camera(cam)
    H, W = cam.image_height, cam.image_width
    accum = torch.zeros(H * W * 3, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream()
    lo, hi = pkg.distributed.sample_range(rank, world, a.spp)

    def step(stats=False):
        accum.zero_()
        st = ctx.render_device(0x5EED, lo, hi, accum.data_ptr(), stream.cuda_stream, want_stats=stats)
        pkg.distributed.reduce_sums(accum)
        return st

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        step()
    fence()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    fence()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    st = step(stats=True)
    fence()
    cnt = torch.tensor([st["samples"], st["rays"]], dtype=torch.float64, device="cuda")
    mean = float((accum / a.spp).mean().item()) if rank == 0 else 0.0
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    if rank == 0:
        samples, rays = cnt.tolist()
        t = float(ms.item()) / a.steps / 1e3
        print(json.dumps({"scene": "C5 Sponza stand-in (262,268 primitives), 3840x2160, depth 5", "n_gpus": world, "scaling": "strong", "spp_total": a.spp,
                          "spp_per_gpu": hi - lo, "mrays_per_s": rays / t / 1e6, "msamples_per_s": samples / t / 1e6, "ms_per_frame_slice": t * 1e3,
                          "rays_per_sample": rays / samples, "image_mean": mean, "reduce_bytes": H * W * 3 * 8, "builder": "device LBVH" if a.builder else "host SAH",
                          "full_config_seconds_at_this_rate": 4096 * H * W / (samples / t)}))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
