#!/usr/bin/env python
"""Where does the multi-GPU exchange time go?  Run under torchrun (one rank per GPU), NCCL_DEBUG=INFO recommended.

Times, on buffers of the configs[2] result size (J 395 MB + Lambda/J_lines 90 MB):
  (1) torch.distributed all_gather_into_tensor (NCCL's own ring/NVLS) as the yardstick,
  (2) a raw peer copy through a CUDA IPC handle (cudaIpcOpenMemHandle + cudaMemcpyAsync), i.e. what a kernel writing straight
      into peer memory over NVLink would see,
  (3) mgb_allgather_results of the library on a real (small-range) run.
Prints one JSON line per rank 0 measurement."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def ev_time(fn, n=5):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    out = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        out.append((a.elapsed_time(b), (time.perf_counter() - t0) * 1e3))
        dist.barrier()
    return out


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    say = (lambda **kw: print(json.dumps(kw), flush=True)) if rank == 0 else (lambda **kw: None)
    N, F, L = 112288, 440, 40
    nbytes = N * (F * 8 + L * 20)
    per = (nbytes // 8 // world) * 8
    full = torch.zeros(per * world, dtype=torch.uint8, device="cuda")
    mine = full[rank * per:(rank + 1) * per]
    mine.fill_(rank + 1)
    t = ev_time(lambda: dist.all_gather_into_tensor(full, mine))
    say(test="torch all_gather_into_tensor (in place)", bytes=per * world, ms_event_wall=t, gbps=per * world / (min(x[0] for x in t) * 1e-3) / 1e9)
    say(peer_access=[torch.cuda.can_device_access_peer(local, j) for j in range(torch.cuda.device_count()) if j != local])

    # (2) CUDA IPC peer copy
    try:
        from cuda.bindings import runtime as rt
        err, h = rt.cudaIpcGetMemHandle(full.data_ptr())
        assert int(err) == 0, err
        handles = [None] * world
        dist.all_gather_object(handles, bytes(h.reserved))
        peer = (rank + 1) % world
        hh = rt.cudaIpcMemHandle_t()
        hh.reserved = handles[peer]
        err, pptr = rt.cudaIpcOpenMemHandle(hh, rt.cudaIpcMemLazyEnablePeerAccess)
        assert int(err) == 0, err
        stream = torch.cuda.current_stream().cuda_stream

        def push():  # my slice -> the peer's buffer at my rows (what the fused reduce kernel's stores would do)
            e, = rt.cudaMemcpyAsync(pptr + rank * per, mine.data_ptr(), per, rt.cudaMemcpyKind.cudaMemcpyDeviceToDevice, stream)
            assert int(e) == 0, e
        t = ev_time(push)
        say(test="IPC peer push of one slice", bytes=per, ms_event_wall=t, gbps=per / (min(x[0] for x in t) * 1e-3) / 1e9)
        dist.barrier()
        rt.cudaIpcCloseMemHandle(pptr)
    except Exception as e:  # noqa: BLE001
        say(test="IPC peer push", error=repr(e))

    # (3) the library's own exchange
    from magritte_b200 import cabi, synth
    from magritte_b200 import modelbuild as mb
    from magritte_b200.flatten import approx_of
    flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=100000, nside_rays=4, nquads=11, seed=1234))
    ctx = cabi.Context(local)
    ctx.set_model(flat)
    uid = [ctx.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(world, rank, uid[0])
    rng = ctx.partition(world, rank, balanced=True)
    ctx.run(approx_of(flat["params"]), (rng[0], rng[0] + 128))
    for mode in (os.environ.get("MGB_GATHER_MODES", "0").split(",")):
        os.environ["MGB_GATHER_MODE"] = mode
        t = ev_time(lambda: ctx.allgather())
        say(test=f"mgb_allgather_results mode {mode}", ms_event_wall=t)
    ctx.close()
    dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
