#!/usr/bin/env python
"""Multi-GPU parity check (run under torchrun, one rank per GPU): partitioned solve + in-place ncclAllGather must
reproduce the single-GPU J and Lambda bit for bit on every rank, and match the golden reference J."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from magritte_b200 import cabi  # noqa: E402
from magritte_b200.flatten import approx_of  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    for name in ("wind3d_small", "c2_vanzadelhoff", "wind3d_spiral_2spec"):
        flat, ref = util.load_golden(name)
        approx = approx_of(flat["params"])
        ctx = cabi.Context(local)
        ctx.set_model(flat)
        ctx.run(approx)
        J1 = ctx.get_J(); L1 = ctx.get_lambda(); JL1 = ctx.get_J_lines()
        uid = [ctx.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(world, rank, uid[0])
        rng = ctx.partition(world, rank, balanced=True)
        if rng[1] > rng[0]:
            ctx.run(approx, rng)
        ctx.allgather()
        N = int(flat["npoints"])
        Jn = ctx.get_J((0, N)); Ln = ctx.get_lambda((0, N)); JLn = ctx.get_J_lines((0, N))
        same = np.array_equal(J1, Jn) and np.array_equal(JL1, JLn) and np.array_equal(L1[0], Ln[0]) and np.array_equal(L1[1], Ln[1])
        err = util.relerr(Jn, ref["ref_J"])
        print(f"rank {rank}/{world} {name}: range {rng} gathered==single {same}, max rel err vs reference {err:.2e}", flush=True)
        ok = ok and same and err <= 1e-10
        ctx.close()
    t = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.barrier(); dist.destroy_process_group()
    if t.item() != 1.0:
        raise SystemExit("multi-GPU parity FAILED")
    if rank == 0:
        print("multi-GPU parity ok")


if __name__ == "__main__":
    main()
