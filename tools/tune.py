#!/usr/bin/env python
"""Tunable sweep for the b200 backend (run under torch.distributed.run, one process per GPU).

Times in-place fp32 all_reduce issued back to back from C++ (BackendB200.repeat_all_reduce) for
combinations of the library's tunables, on ordinary tensors (staged / auto-registered) and on
symmetric memory (zero-copy).  Prints one JSON line per measurement on rank 0.  Used to pick
the defaults in csrc/nbd_lib.cu; the numbers that justify them are kept under profiles/.
"""
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
    factor = 2.0 * (world - 1) / world

    def measure(buf, iters):
        be.repeat_all_reduce(buf, 3)
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        be.repeat_all_reduce(buf, iters)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=cpu)
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu)
        return float(t.item())

    def report(kind, nbytes, knobs, us):
        if rank == 0:
            print(json.dumps({"kind": kind, "bytes": nbytes, **knobs, "us": round(us, 2),
                              "busbw": round(nbytes * factor / us / 1e3, 1), "algo": be.last_algo()}), flush=True)

    def with_knobs(knobs, fn):
        for k, v in knobs.items():
            be.set_tunable(k, v)
        try:
            return fn()
        finally:
            for k in knobs:
                be.set_tunable(k, 0)

    sym = B.empty_symmetric((256 << 20) // 4, torch.float32)
    ordinary = torch.zeros((256 << 20) // 4, device=dev, dtype=torch.float32)
    sizes = [1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20]
    for nbytes in sizes:
        iters = 30 if nbytes <= (16 << 20) else 8
        o = ordinary[: nbytes // 4]
        s = sym[: nbytes // 4]
        # staged paths on ordinary memory
        for blocks in (0, 64, 32):
            for chunk in ((0, 8 << 20, 64 << 20) if nbytes >= (64 << 20) else (0,)):
                for ar1 in ((0, 2 << 20, 8 << 20) if nbytes <= (4 << 20) else (0,)):
                    knobs = {"auto_register_off": 1, "blocks": blocks, "chunk_bytes": chunk, "ar1_max_bytes": ar1}
                    report("staged", nbytes, knobs, with_knobs(knobs, lambda: measure(o, iters)))
        # auto-registered ordinary memory (agreement kernel + zero-copy)
        knobs = {"auto_min_bytes": 1 << 20}
        with_knobs(knobs, lambda: [be.repeat_all_reduce(o, 1), torch.cuda.synchronize(), be.repeat_all_reduce(o, 1),
                                   torch.cuda.synchronize()])
        report("auto", nbytes, knobs, with_knobs(knobs, lambda: measure(o, iters)))
        # zero-copy on symmetric memory
        for mode, tile in ((1, 0), (2, 0), (2, 256 << 10), (2, 2 << 20)):
            for blocks in (0, 64):
                knobs = {"zc_min_bytes": 1, "zc_ar_mode": mode, "zc_tile_bytes": tile, "blocks": blocks}
                report("zerocopy", nbytes, knobs, with_knobs(knobs, lambda: measure(s, iters)))
    dist.barrier(group=cpu)
    del sym
    W.shutdown_worker()


if __name__ == "__main__":
    main()
