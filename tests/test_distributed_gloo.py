"""N > 1 host logic on CPU: world_size-2 (and 3) gloo process groups drive cpu-ray-tracing-implementation_b200.distributed
with the ORACLE standing in for the GPU renderer (tests may use it as the checker; the product path never does).  What is
checked is the sharding contract the GPU bench relies on: disjoint covering sample ranges, Philox streams keyed by the global
sample index, one SUM reduce onto rank 0 == the single-process render."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, golden, pkg

D = pkg.distributed


def test_sample_ranges_partition():
    for world in (1, 2, 3, 8):
        for spp in (0, 1, 7, 8, 500, 4096):
            r = [D.sample_range(k, world, spp) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
            w = [D.weak_sample_range(k, world, spp) for k in range(world)]
            assert w[0][0] == 0 and w[-1][1] == world * spp and all(b - a == spp for a, b in w)
    with pytest.raises(ValueError):
        D.sample_range(2, 2, 8)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, spp, weak, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import oracleport
    from conftest import golden as g_

    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = g_("cornell")
    H, W = g.cam_small.image_height, g.cam_small.image_width
    accum = torch.zeros(H * W * 3, dtype=torch.float64)

    def render_range(begin, end):
        r = oracleport.render(g.scene, g.cam_small, 77, begin, end, threads=1)
        accum.add_(torch.from_numpy(r["sum"].reshape(-1)))

    rng = D.render_distributed(render_range, spp, accum, weak=weak)
    ranges = [None] * world
    dist.all_gather_object(ranges, rng)
    if rank == 0:
        np.savez(out_path, sum=accum.numpy().reshape(H, W, 3), ranges=np.array(ranges))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,weak", [(2, False), (3, False), (2, True)])
def test_split_render_equals_single_process(world, weak, build_oracle, tmp_path):
    spp = 7
    out = str(tmp_path / "out.npz")
    mp.spawn(_worker, args=(world, _free_port(), spp, weak, out), nprocs=world, join=True)
    z = np.load(out)
    total = spp * world if weak else spp
    g = golden("cornell")
    ref = build_oracle.render(g.scene, g.cam_small, 77, 0, total, threads=1)["sum"]
    assert np.allclose(z["sum"], ref, rtol=1e-12, atol=1e-12)
    r = z["ranges"]
    assert r[0][0] == 0 and r[-1][1] == total and all(r[k][1] == r[k + 1][0] for k in range(world - 1))
