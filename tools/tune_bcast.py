#!/usr/bin/env python
"""Broadcast variants on real ranks (run under torch.distributed.run): the chunked scatter kernel,
the relay kernel (k_broadcast_relay) over tile size / slots / pushing groups, NCCL beside them.
One JSON line per measurement on rank 0; every variant's result is checked first."""
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
    nccl = None if "--no-nccl" in sys.argv else dist.new_group(backend="nccl")
    full = "--full" in sys.argv

    def emit(obj):
        if rank == 0:
            print(json.dumps(obj), flush=True)

    def timed(fn, iters, warmup=2):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu)
        return float(t.item())

    big = torch.empty((1 << 30) // 4, device=dev, dtype=torch.float32)
    variants = [("staged", {"pipe": 2})]
    tiles = (0, 512 << 10) if full else (0, 128 << 10)
    for tile in tiles:
        for depth in ((4, 8) if full else (4,)):
            for mode in (2, 4):
                variants.append((f"relay/tile={tile >> 10}K/depth={depth}/groups={2 if mode == 2 else 1}",
                                 {"pipe": 3, "pipe_tile_bytes": tile, "pipe_depth": depth, "pipe_mode": mode}))
    sizes = [64 << 20, 256 << 20, 1 << 30] if full else [1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30]
    for nbytes in sizes:
        x = big[: nbytes // 4]
        iters = 30 if nbytes <= (16 << 20) else (10 if nbytes <= (256 << 20) else 4)
        for name, kv in variants:
            for k, v in kv.items():
                be.set_tunable(k, v)
            root = world - 1
            x.fill_(float(rank + 1))
            x[-1] = float(rank + 7)
            dist.broadcast(x, src=root)
            ok = bool((x[:-1] == float(root + 1)).all()) and float(x[-1]) == float(root + 7)
            algo = be.last_algo()
            us = timed(lambda: be.repeat_broadcast(x, 0, iters), 1, warmup=1) / iters
            for k in kv:
                be.set_tunable(k, 0)
            emit({"op": "broadcast", "variant": name, "bytes": nbytes, "us": round(us, 2), "GBps": round(nbytes / us / 1e3, 1),
                  "algo": algo, "ok": ok})
        if nccl is not None:
            us = timed(lambda: dist.broadcast(x, src=0, group=nccl), iters)
            emit({"op": "broadcast", "variant": "nccl", "bytes": nbytes, "us": round(us, 2), "GBps": round(nbytes / us / 1e3, 1)})
    dist.barrier(group=cpu)
    del big
    W.shutdown_worker()


if __name__ == "__main__":
    main()
