"""N > 1 host logic on the CPU (gloo, world_size 2): the origin partition of mgb_partition_rows and the padded in-place
all-gather layout of mgb_allgather_results, with the oracle standing in for the per-rank GPU solve."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import util  # noqa: E402
from magritte_b200 import cabi  # noqa: E402


@pytest.mark.parametrize("npoints,nranks", [(10, 1), (10, 2), (11, 2), (7, 8), (112288, 8), (1, 4)])
def test_partition_rows_cover_once(npoints, nranks):
    seen = np.zeros(npoints, dtype=int)
    per0 = None
    for r in range(nranks):
        a, b, per = cabi.partition_rows(npoints, nranks, r)
        per0 = per if per0 is None else per0
        assert per == per0 and per * nranks >= npoints and b - a <= per and a == min(r * per, npoints)
        seen[a:b] += 1
    assert (seen == 1).all()


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc
    flat, ref = util.load_golden("velgrad_1d")
    N, F = int(flat["npoints"]), int(flat["nfreqs"])
    a, b, per = cabi.partition_rows(N, world, rank)
    o = orc.Oracle(flat, store_u=False, threads=1)
    Jfull = o.solve().copy()
    Jslice = o.solve(o_range=(a, b))[a:b].copy() if b > a else np.zeros((0, F))
    # padded in-place layout: every rank's slice sits at rank*per rows of a (per*world)-row buffer
    buf = torch.zeros(per * world, F, dtype=torch.float64)
    buf[a:b] = torch.from_numpy(Jslice)
    parts = [torch.zeros(per, F, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(parts, buf[rank * per:(rank + 1) * per].clone())
    J = torch.cat(parts)[:N].numpy()
    ok = np.array_equal(J, Jfull) and util.relerr(J, ref["ref_J"]) <= 1e-10
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_two_rank_gather_equals_single_rank():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert all(out[r] for r in range(world)), dict(out)
