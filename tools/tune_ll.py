#!/usr/bin/env python
"""Where the LL (flag-in-data) path should hand over to the barrier kernels, on real ranks (run
under torch.distributed.run): device-side time of back-to-back all_reduces / broadcasts of
4 KiB .. 1 MiB with the LL threshold at 32 KiB (old default) and at 256 KiB.  One JSON line per
measurement on rank 0; results are checked first."""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    W.init_worker(rank, world, os.environ.get("MASTER_ADDR", "127.0.0.1"), os.environ.get("MASTER_PORT", "29500"),
                  gpu_id=local, backend="b200")
    dev = torch.device("cuda", local)
    cpu = dist.new_group(backend="gloo")
    be = B.get_backend()

    def emit(obj):
        if rank == 0:
            print(json.dumps(obj), flush=True)

    def timed(fn, iters):
        fn(20)
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(iters)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu)
        return float(t.item())

    for nbytes in (4 << 10, 32 << 10, 48 << 10, 64 << 10, 96 << 10, 128 << 10, 192 << 10, 256 << 10, 512 << 10, 1 << 20):
        n = nbytes // 4
        for name, ll_max in (("ll<=32K", 32 << 10), ("ll<=256K", 256 << 10)):
            be.set_tunable("ll_max_bytes", ll_max)
            x = torch.full((n,), float(rank + 1), device=dev)
            dist.all_reduce(x)
            ok = bool((x == float(world * (world + 1) // 2)).all())
            algo = be.last_algo()
            x.zero_()
            us = timed(lambda it: be.repeat_all_reduce(x, it), 200)
            emit({"op": "all_reduce", "variant": name, "bytes": nbytes, "us": round(us, 2), "algo": algo, "ok": ok})
            y = torch.full((n,), float(rank + 1), device=dev)
            dist.broadcast(y, src=world - 1)
            okb = bool((y == float(world)).all())
            algob = be.last_algo()
            usb = timed(lambda it: be.repeat_broadcast(y, 0, it), 200)
            emit({"op": "broadcast", "variant": name, "bytes": nbytes, "us": round(usb, 2), "algo": algob, "ok": okb})
        be.set_tunable("ll_max_bytes", 0)
    dist.barrier(group=cpu)
    W.shutdown_worker()


if __name__ == "__main__":
    main()
