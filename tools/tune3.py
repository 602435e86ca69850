#!/usr/bin/env python
"""Round-2 follow-up sweep (under torch.distributed.run): (a) the host-buffer all_reduce for several
pipeline chunk sizes next to the box's pinned-copy ceiling, (b) the zero-copy in-switch all_reduce
on multicast-pool memory for several grid sizes.  One JSON line per measurement on rank 0."""
from __future__ import annotations

import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    import bench
    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    bound = bench.bind_near_gpu(bench.physical_gpu_index(local))
    W.init_worker(rank, world, os.environ.get("MASTER_ADDR", "127.0.0.1"), os.environ.get("MASTER_PORT", "29500"),
                  gpu_id=local, backend="b200")
    dev = torch.device("cuda", local)
    cpu = dist.new_group(backend="gloo")
    be = B.get_backend()
    f_ar = 2.0 * (world - 1) / world

    def emit(obj):
        if rank == 0:
            print(json.dumps(obj), flush=True)

    S = 256 << 20
    h_in = torch.ones(S // 4, dtype=torch.float32).pin_memory()
    h_out = torch.empty(S // 4, dtype=torch.float32).pin_memory()
    ceil_ms = bench.host_copy_ceiling(torch, dist, dev, cpu, S)
    emit({"what": "host_copy_only", "ms": round(ceil_ms, 3), "cpus_bound": bound})
    for chunk in (16 << 20, 32 << 20, 64 << 20, 128 << 20):
        be.set_tunable("host_chunk_bytes", chunk)
        ms = bench._time_collective(torch, dist, lambda: B.all_reduce_host(h_in, out=h_out), 4, 1, cpu)
        emit({"what": "all_reduce_host", "chunk_MiB": chunk >> 20, "ms": round(ms, 3), "ok": bool(float(h_out[0]) == world),
              "frac_of_ceiling": round(ceil_ms / ms, 3)})
    be.set_tunable("host_chunk_bytes", 0)
    del h_in, h_out

    if be.nvls_status()[0]:
        be.set_tunable("nvls_reduce", 2)
        pool = B.empty_symmetric((256 << 20) // 4, torch.float32)
        for nbytes in (16 << 20, 64 << 20, 256 << 20):
            buf = pool[: nbytes // 4]
            for blocks in (0, 32, 64, 96):
                be.set_tunable("blocks", blocks)
                buf.fill_(1.0)
                dist.all_reduce(buf)
                ok = bool((buf == float(world)).all()) and be.last_algo() == "ar_nvls_zerocopy"
                buf.zero_()
                ms = bench._time_collective(torch, dist, lambda: dist.all_reduce(buf), 10, 3, cpu)
                emit({"what": "ar_nvls_zerocopy", "bytes": nbytes, "blocks": blocks, "us": round(ms * 1e3, 1),
                      "busbw": round(nbytes * f_ar / ms / 1e6, 1), "ok": ok})
        be.set_tunable("blocks", 0)
        del pool, buf
        be.set_tunable("nvls_reduce", 0)
    dist.barrier(group=cpu)
    W.shutdown_worker()


if __name__ == "__main__":
    main()
