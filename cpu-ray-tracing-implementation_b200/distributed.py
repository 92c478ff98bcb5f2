"""Frame split across the GPUs of one box: one process per GPU, no data-path collective except one reduce.

The reference's only parallel construct is the row loop of camera::render (src/camera.h:158); samples are independent
(:163-170), so rank g of G renders the GLOBAL sample indices [begin, end) of every pixel given by `sample_range`, with Philox
draws keyed by (seed, pixel, global sample index): the union over ranks is the same set of samples whatever G is, and the
streams are decorrelated by construction.  The per-rank fp64 accumulation buffers (sums, not means) are combined with one
`torch.distributed.reduce(SUM)` onto rank 0 (NCCL over NVLink on GPUs, gloo in the CPU tests); rank 0 divides by the total
sample count (b200rt_finalize).
"""


def sample_range(rank, world, total_spp):
    """Strong-scaling split of `total_spp` samples per pixel: contiguous, disjoint, covering, sizes differ by at most 1."""
    if not (0 <= rank < world) or total_spp < 0:
        raise ValueError(f"sample_range(rank={rank}, world={world}, total_spp={total_spp})")
    base, extra = divmod(total_spp, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def weak_sample_range(rank, world, spp_per_rank):
    """Weak-scaling split used by bench.py: every rank renders `spp_per_rank` samples per pixel, globally numbered."""
    if not (0 <= rank < world) or spp_per_rank < 0:
        raise ValueError(f"weak_sample_range(rank={rank}, world={world}, spp_per_rank={spp_per_rank})")
    return rank * spp_per_rank, (rank + 1) * spp_per_rank


def reduce_sums(accum, dst=0):
    """Sum the per-rank accumulation buffers onto `dst` (in place).  `accum` is a torch tensor on the device the process
    group's backend works on.  No-op without an initialised process group (single GPU)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def render_distributed(render_range, total_spp, accum, weak=False):
    """Drive one frame: `render_range(begin, end)` must ADD the sums of samples [begin, end) into `accum` (a torch tensor);
    returns (begin, end) of this rank.  With weak=True `total_spp` is per rank."""
    import torch.distributed as dist

    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0
    begin, end = (weak_sample_range if weak else sample_range)(rank, world, total_spp)
    if end > begin:
        render_range(begin, end)
    reduce_sums(accum)
    return begin, end
