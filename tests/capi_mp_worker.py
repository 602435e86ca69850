"""Per-rank body of the C-ABI multi-process test: real ranks (one process per GPU), handles
exchanged through multiprocessing queues instead of a c10d store — torch.distributed is not
involved, only include/nbdist_b200.h through ctypes."""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main(rank: int, world: int, qs, result_q) -> None:
    import faulthandler

    faulthandler.enable(all_threads=True)
    try:
        _body(rank, world, qs)
        result_q.put((rank, None))
    except Exception as e:  # noqa: BLE001
        import traceback

        result_q.put((rank, f"{e}\n{traceback.format_exc()}"))


def _body(rank: int, world: int, qs) -> None:
    import torch

    from nbdistributed_b200 import capi
    from oracle import collectives_oracle as O
    from tests import helpers as H

    lib = capi.load()
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    comm = C.c_void_p()
    capi.check(lib.nbd_comm_create(C.byref(comm), rank, world, rank, 64 << 20))
    capi.check(lib.nbd_comm_set_timeout_ms(comm, 30000))
    mine = (C.c_ubyte * capi.NBD_HANDLE_BYTES)()
    capi.check(lib.nbd_comm_get_handle(comm, mine))
    for r in range(world):  # all-gather of handles over the queues
        if r != rank:
            qs[r].put((rank, bytes(mine)))
    handles = {rank: bytes(mine)}
    while len(handles) < world:
        src, blob = qs[rank].get(timeout=120)
        handles[src] = blob
    allh = b"".join(handles[r] for r in range(world))
    capi.check(lib.nbd_comm_connect(comm, allh))
    stream = torch.cuda.current_stream().cuda_stream

    def sync():
        torch.cuda.synchronize()
        capi.check(lib.nbd_comm_async_error(comm))

    for dtype in ("f32", "bf16", "i32"):
        for numel in (1, 1000, (1 << 20) + 3, 24 << 20):
            x = O.make_inputs(world, numel, dtype)
            t = H.to_cuda(x[rank], dtype, dev)
            capi.check(lib.nbd_all_reduce(comm, t.data_ptr(), t.data_ptr(), numel, capi.torch_dtype_code(t.dtype),
                                          capi.SUM, stream))
            sync()
            H.assert_same(H.to_numpy(t, dtype), O.reduce_rank_order(x, dtype), dtype,
                          f"all_reduce {dtype} {numel} {lib.nbd_last_algo(comm).decode()}")
    for nbytes in (40, (1 << 20) + 5, (70 << 20) + 1):
        for root in (0, world - 1):
            x = O.make_inputs(world, nbytes, "u8", seed=root)
            t = H.to_cuda(x[rank], "u8", dev)
            capi.check(lib.nbd_broadcast(comm, t.data_ptr(), nbytes, capi.BYTES, root, stream))
            sync()
            H.assert_same(H.to_numpy(t, "u8"), x[root], "u8", f"broadcast {nbytes} root {root}")
    for numel in (25, (3 << 20) + 8):
        x = O.make_inputs(world, numel, "bf16")
        t = H.to_cuda(x[rank], "bf16", dev)
        out = torch.empty(world * numel, dtype=torch.bfloat16, device=dev)
        capi.check(lib.nbd_all_gather(comm, t.data_ptr(), out.data_ptr(), numel, capi.BF16, stream))
        sync()
        H.assert_same(H.to_numpy(out, "bf16"), O.all_gather(x)[0], "bf16", f"all_gather {numel}")
        xs = O.make_inputs(world, world * numel, "bf16", seed=5)
        s = H.to_cuda(xs[rank], "bf16", dev)
        o = torch.empty(numel, dtype=torch.bfloat16, device=dev)
        capi.check(lib.nbd_reduce_scatter(comm, s.data_ptr(), o.data_ptr(), numel, capi.BF16, capi.SUM, stream))
        sync()
        H.assert_same(H.to_numpy(o, "bf16"), O.reduce_scatter(xs, "bf16")[rank], "bf16", f"reduce_scatter {numel}")
    capi.check(lib.nbd_barrier(comm, stream))
    sync()
    # nobody unmaps while a peer may still pull: rendezvous over the queues before destroy
    for r in range(world):
        if r != rank:
            qs[r].put((rank, b"bye"))
    seen = 0
    while seen < world - 1:
        qs[rank].get(timeout=120)
        seen += 1
    capi.check(lib.nbd_comm_destroy(comm))
