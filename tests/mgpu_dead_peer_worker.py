"""Per-rank body of the dead-peer test (spawned by test_backend_multigpu.py, 2 processes).

The reference hangs a cell forever when a rank never reaches a collective (default timeout None,
ref:src/nbdistributed/communication.py:263-269).  Here rank 1 drops out right after the group's
first collective (it stays idle and never joins another one); rank 0 then issues an all_reduce
nobody answers.  Its kernel must give up after
the group's timeout, work.wait(timeout) must return control to the host with an exception, and
the next call must fail with DistBackendError instead of hanging."""
from __future__ import annotations

import datetime
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main(rank: int, world: int, port: int) -> None:
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import worker_init as W

    os.environ["NBDIST_B200_SHUTDOWN_WAIT_MS"] = "200"
    W.init_worker(rank, world, "127.0.0.1", port, gpu_id=rank, backend="b200", timeout=datetime.timedelta(seconds=3))
    dev = torch.device("cuda", rank)
    t = torch.ones(1 << 20, device=dev)
    dist.all_reduce(t)  # creates and connects the communicator on both ranks
    torch.cuda.synchronize()
    assert float(t[0]) == float(world)
    if rank == 1:
        # never shows up for the next collective.  (It stays alive, idle, until the survivor is
        # done: freeing its arena under the survivor's running kernel would test the driver's
        # IPC lifetime rules, not this library.)
        time.sleep(8.0)
        os._exit(0)
    time.sleep(0.5)
    t0 = time.perf_counter()
    work = dist.all_reduce(t, async_op=True)
    try:
        work.wait(datetime.timedelta(seconds=1))  # host-blocking wait with a deadline
        raise AssertionError("wait(timeout) returned although the peer is gone")
    except RuntimeError as e:
        assert "expired" in str(e) or "timed out" in str(e) or "aborted" in str(e), e
    assert time.perf_counter() - t0 < 10.0, "wait(timeout) did not bound the wait"
    torch.cuda.synchronize()  # the kernel left its flag wait (abort), the stream drains
    try:
        dist.all_reduce(t)
        raise AssertionError("a dead communicator accepted another collective")
    except RuntimeError as e:
        assert "abort" in str(e).lower() or "timed out" in str(e).lower(), e
    print("[dead-peer] survivor got errors instead of a hang", flush=True)
    os._exit(0)  # the store's other client is gone: skip the orderly teardown
