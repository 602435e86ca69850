#!/usr/bin/env python
"""Round-2 sweep (run under torch.distributed.run, one process per GPU): every all_reduce and
broadcast path of the b200 backend on the same buffers, one JSON line per measurement on rank 0.

  all_reduce : staged (chunked two-shot) | auto (auto-registered zero-copy) | pipe (tile pipeline,
               peer-to-peer transport) | nvls (tile pipeline, in-switch reduction) | NCCL
  broadcast  : staged (scatter) | pipe | nvls | NCCL

Every variant is checked before it is timed: integer-valued fp32 inputs make the sum exact in
any order, so all paths must return the same bits.  Numbers that justify defaults go to profiles/.
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
    quick = "--quick" in sys.argv
    nccl = None if "--no-nccl" in sys.argv else dist.new_group(backend="nccl")

    def emit(obj):
        if rank == 0:
            print(json.dumps(obj), flush=True)

    nvls_only = "--nvls" in sys.argv
    warm = torch.zeros(1024, device=dev)
    dist.all_reduce(warm)  # creates the communicator (and the multicast object)
    torch.cuda.synchronize()
    st = be.nvls_status()
    gathered = [None] * world
    dist.all_gather_object(gathered, (bool(st[0]), st[1]), group=cpu)
    emit({"kind": "nvls_status", "world": world, "per_rank": gathered})
    have_nvls = all(g[0] for g in gathered)

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

    class knobs:
        def __init__(self, **kv):
            self.kv = kv

        def __enter__(self):
            for k, v in self.kv.items():
                be.set_tunable(k, v)

        def __exit__(self, *exc):
            for k in self.kv:
                be.set_tunable(k, 0)

    big = torch.empty((1 << 30) // 4, device=dev, dtype=torch.float32)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    f_ar = 2.0 * (world - 1) / world
    if nvls_only:
        # in-switch all_reduce: pipeline shape sweep (tile bytes, slots, reducing groups, loads in flight)
        for nbytes in (64 << 20, 256 << 20):
            n = nbytes // 4
            x = big[:n]
            src = torch.randint(-1000, 1000, (n,), generator=g, device=dev, dtype=torch.int32)
            want = src.clone()
            dist.all_reduce(want)
            want = want.to(torch.float32)
            per4 = -(-nbytes // (148 * 4) // 128) * 128
            per2 = -(-nbytes // (148 * 2) // 128) * 128
            per8 = -(-nbytes // (148 * 8) // 128) * 128
            for tile in ((256 << 10 if nbytes > (64 << 20) else per4),):
                for depth in (4,):
                    for mode in (0, 2):
                        kv = {"nvls_reduce": 2, "nvls_min_bytes": 1, "pipe_tile_bytes": tile, "pipe_depth": depth, "pipe_mode": mode}
                        with knobs(**kv):
                            x.copy_(src)
                            dist.all_reduce(x)
                            ok = bool(torch.equal(x, want))
                            algo = be.last_algo()
                            x.zero_()
                            us = timed(lambda: dist.all_reduce(x), 10)
                        emit({"op": "all_reduce", "variant": f"nvls/tile={tile}/depth={depth}/mode={mode}", "bytes": nbytes,
                              "us": round(us, 2), "busbw": round(nbytes * f_ar / us / 1e3, 1), "algo": algo, "ok": ok})
            # where the time goes: the same launch with copy-in (8), copy-out (16) or the in-switch
            # reduction (32) skipped (results are void; timing only)
            for mode in (8, 16, 24, 32, 40, 48, 56):
                kv = {"nvls_reduce": 2, "nvls_min_bytes": 1, "pipe_mode": mode}
                with knobs(**kv):
                    us = timed(lambda: dist.all_reduce(x), 10)
                emit({"op": "all_reduce", "variant": f"nvls-diag/skip={mode}", "bytes": nbytes, "us": round(us, 2),
                      "busbw": round(nbytes * f_ar / us / 1e3, 1)})
            with knobs(auto_min_bytes=1 << 20, pipe=2, nvls_off=1):
                for _ in range(3):
                    dist.all_reduce(x)
                us = timed(lambda: dist.all_reduce(x), 10)
            emit({"op": "all_reduce", "variant": "auto", "bytes": nbytes, "us": round(us, 2), "busbw": round(nbytes * f_ar / us / 1e3, 1)})
            if nccl is not None:
                us = timed(lambda: dist.all_reduce(x, group=nccl), 10)
                emit({"op": "all_reduce", "variant": "nccl", "bytes": nbytes, "us": round(us, 2), "busbw": round(nbytes * f_ar / us / 1e3, 1)})
            del src, want
        dist.barrier(group=cpu)
        del big
        W.shutdown_worker()
        return
    sizes = [1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20] + ([] if quick else [1 << 30])
    f_ar = 2.0 * (world - 1) / world

    ar_variants = [("staged", {"auto_register_off": 1, "pipe": 2, "nvls_off": 1}),
                   ("auto", {"auto_min_bytes": 1 << 20, "pipe": 2, "nvls_off": 1})]
    for tile in (0,):
        ar_variants.append((f"pipe/tile={tile >> 10}K", {"pipe": 1, "pipe_tile_bytes": tile, "nvls_off": 1}))
    if have_nvls:
        for tile in (0, 128 << 10):
            for blocks in (0, 96):
                ar_variants.append((f"nvls/tile={tile >> 10}K/blocks={blocks}",
                                    {"nvls_reduce": 2, "nvls_min_bytes": 1, "pipe_tile_bytes": tile, "blocks": blocks}))
    for nbytes in sizes:
        n = nbytes // 4
        x = big[:n]
        src = torch.randint(-1000, 1000, (n,), generator=g, device=dev, dtype=torch.int32)
        want = src.clone()
        dist.all_reduce(want)  # int32: exact on any path
        want = want.to(torch.float32)
        iters = 30 if nbytes <= (16 << 20) else (10 if nbytes <= (256 << 20) else 4)
        for name, kv in ar_variants:
            with knobs(**kv):
                x.copy_(src)
                dist.all_reduce(x)
                if name == "auto":  # first touches export/import the allocation
                    for _ in range(2):
                        x.copy_(src)
                        dist.all_reduce(x)
                ok = bool(torch.equal(x, want))
                algo = be.last_algo()
                x.zero_()
                us = timed(lambda: dist.all_reduce(x), iters)
            emit({"op": "all_reduce", "variant": name, "bytes": nbytes, "us": round(us, 2),
                  "busbw": round(nbytes * f_ar / us / 1e3, 1), "algo": algo, "ok": ok})
        if nccl is not None:
            x.zero_()
            us = timed(lambda: dist.all_reduce(x, group=nccl), iters)
            emit({"op": "all_reduce", "variant": "nccl", "bytes": nbytes, "us": round(us, 2),
                  "busbw": round(nbytes * f_ar / us / 1e3, 1)})
        del src, want

    # does the content of the buffers matter?  (constant buffers are what most sweeps time)
    nbytes = 256 << 20
    x = big[: nbytes // 4]
    dvariants = [("auto", {"auto_min_bytes": 1 << 20, "pipe": 2, "nvls_off": 1})]
    if have_nvls:
        dvariants.append(("nvls", {"nvls_reduce": 2, "nvls_min_bytes": 1}))
    for data in ("zeros", "ones", "randn"):
        def fill():
            if data == "zeros":
                x.zero_()
            elif data == "ones":
                x.fill_(1.0)
            else:
                x.normal_(generator=g)
                x.mul_(1e-30)  # in-place sums grow by world per iteration: stay finite
        for name, kv in dvariants:
            with knobs(**kv):
                fill()
                us = timed(lambda: dist.all_reduce(x), 10)
                algo = be.last_algo()
            emit({"op": "all_reduce", "variant": f"{name}/data={data}", "bytes": nbytes, "us": round(us, 2),
                  "busbw": round(nbytes * f_ar / us / 1e3, 1), "algo": algo})
        if nccl is not None:
            fill()
            us = timed(lambda: dist.all_reduce(x, group=nccl), 10)
            emit({"op": "all_reduce", "variant": f"nccl/data={data}", "bytes": nbytes, "us": round(us, 2),
                  "busbw": round(nbytes * f_ar / us / 1e3, 1)})

    bc_variants = [("staged", {"pipe": 2, "nvls_off": 1}), ("pipe", {"pipe": 1, "nvls_off": 1})]
    if have_nvls:
        for tile in (0,):
            bc_variants.append((f"nvls/tile={tile >> 10}K", {"nvls_min_bytes": 1, "pipe_tile_bytes": tile}))
    for nbytes in sizes:
        n = nbytes // 4
        x = big[:n]
        iters = 30 if nbytes <= (16 << 20) else (10 if nbytes <= (256 << 20) else 4)
        for name, kv in bc_variants:
            with knobs(**kv):
                root = world - 1
                x.fill_(float(rank + 1))
                dist.broadcast(x, src=root)
                ok = bool((x == float(root + 1)).all())
                algo = be.last_algo()
                us = timed(lambda: dist.broadcast(x, src=0), iters)
            emit({"op": "broadcast", "variant": name, "bytes": nbytes, "us": round(us, 2),
                  "busbw": round(nbytes / us / 1e3, 1), "algo": algo, "ok": ok})
        if nccl is not None:
            us = timed(lambda: dist.broadcast(x, src=0, group=nccl), iters)
            emit({"op": "broadcast", "variant": "nccl", "bytes": nbytes, "us": round(us, 2), "busbw": round(nbytes / us / 1e3, 1)})
    dist.barrier(group=cpu)
    del big
    W.shutdown_worker()


if __name__ == "__main__":
    main()
