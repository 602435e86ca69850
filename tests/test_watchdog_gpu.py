"""The flag-wait watchdog and the abort path, on one GPU with emulated ranks.

The reference hangs forever when a rank never reaches a collective (default timeout None,
ref:src/nbdistributed/communication.py:263-269).  Here a peer that never signals
is reproduced with the "debug_mute_rank" tunable (its signals are redirected, as if its process
had died); the survivors must give up after the configured timeout, report NBD_ERR_TIMEOUT,
drain quickly even though the launch has many more flag waits ahead, and leave the communicator
unusable.  nbd_comm_abort must free kernels that are already waiting.
"""
import time

import pytest

pytestmark = pytest.mark.gpu

TIMEOUT, ABORTED = -4, -5


def _tensors(torch, world, numel):
    return [torch.ones(numel, device="cuda:0") for _ in range(world)]


@pytest.mark.parametrize("path", [{"zero_copy_off": 1}, {"zc_min_bytes": 1}, {"pipe": 1}], ids=["staged", "zerocopy", "pipe"])
def test_missing_peer_times_out_and_kills_the_communicator(path):
    import torch

    from nbdistributed_b200 import capi

    world = 4
    grp = capi.EmuGroup(world, device=0, stage_bytes=8 << 20)
    try:
        for k, v in path.items():
            grp.set_tunable(k, v)
        t = _tensors(torch, world, 3 << 20)  # 12 MiB through 8 MiB of staging: several chunks, many flag waits
        grp.all_reduce(t, t)                 # healthy first
        torch.cuda.synchronize()
        grp.async_error()
        grp.set_timeout_ms(300)
        grp.set_tunable("debug_mute_rank", 2 + 1)
        t0 = time.perf_counter()
        grp.all_reduce(t, t)
        torch.cuda.synchronize()
        elapsed = time.perf_counter() - t0
        with pytest.raises(capi.NbdError) as ei:
            grp.async_error()
        assert ei.value.code == TIMEOUT, ei.value
        assert "timed out" in str(ei.value)
        # one timeout, then every later wait of the launch drains at once
        assert 0.25 < elapsed < 3.0, f"{elapsed:.2f} s"
        # the communicator is dead: the next call fails on the host, nothing is launched
        n0 = capi.launch_count()
        with pytest.raises(capi.NbdError):
            grp.all_reduce(t, t)
        assert capi.launch_count() == n0
    finally:
        grp.close()


def test_abort_releases_waiting_kernels():
    import torch

    from nbdistributed_b200 import capi

    world = 3
    grp = capi.EmuGroup(world, device=0, stage_bytes=8 << 20)
    try:
        grp.set_timeout_ms(60_000)
        grp.set_tunable("debug_mute_rank", 0 + 1)
        t = _tensors(torch, world, 1 << 20)
        grp.all_reduce(t, t)       # ranks 1 and 2 now wait for rank 0, for up to a minute
        time.sleep(0.2)
        t0 = time.perf_counter()
        grp.abort()
        torch.cuda.synchronize()
        assert time.perf_counter() - t0 < 5.0
        with pytest.raises(capi.NbdError) as ei:
            grp.async_error()
        assert ei.value.code in (ABORTED, TIMEOUT), ei.value
    finally:
        grp.close()
