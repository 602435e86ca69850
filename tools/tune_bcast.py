#!/usr/bin/env python
"""Broadcast pipeline-chunk sweep (run under torch.distributed.run): dist.broadcast from rank 0 for
several message sizes and "chunk_bytes" / "bcast_direct_max_bytes" settings.  JSON lines on rank 0."""
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
    big = torch.ones((1 << 30) // 4, device=dev)

    def measure(buf, iters):
        for _ in range(2):
            dist.broadcast(buf, src=0)
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            dist.broadcast(buf, src=0)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu)
        return float(t.item())

    for nbytes in (1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30):
        buf = big[: nbytes // 4]
        iters = 20 if nbytes <= (64 << 20) else 5
        for chunk in (0, 2 << 20, 4 << 20, 8 << 20, 16 << 20, 32 << 20):
            if chunk and chunk >= nbytes:
                continue
            for direct in ((0, 1 << 40) if nbytes <= (4 << 20) else (0,)):
                be.set_tunable("chunk_bytes", chunk)
                be.set_tunable("bcast_direct_max_bytes", direct)
                us = measure(buf, iters)
                be.set_tunable("chunk_bytes", 0)
                be.set_tunable("bcast_direct_max_bytes", 0)
                if rank == 0:
                    print(json.dumps({"bytes": nbytes, "chunk_bytes": chunk, "direct_max": direct, "us": round(us, 2),
                                      "busbw": round(nbytes / us / 1e3, 1), "algo": be.last_algo()}), flush=True)
    dist.barrier(group=cpu)
    W.shutdown_worker()


if __name__ == "__main__":
    main()
