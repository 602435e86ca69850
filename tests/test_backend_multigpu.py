"""The c10d backend end to end, one process per GPU, through torch.distributed.

Needs >= 2 GPUs: skipped on the single-GPU box (where tests/test_emu_gpu.py covers the same
kernels with emulated ranks) and run with ``gpurun --gpus N``.  NBD_TEST_WORLD limits the
world size (default: all visible GPUs, at most 8)."""
import os

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _world():
    import torch

    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    want = int(os.environ.get("NBD_TEST_WORLD", "0"))
    return min(n, want) if want else min(n, 8)


def test_backend_against_gloo_and_oracle():
    import torch.multiprocessing as mp

    from oracle.gloo_runner import free_port
    from tests import mgpu_worker

    world = _world()
    if world < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    mp.spawn(mgpu_worker.main, args=(world, free_port()), nprocs=world, join=True)


def test_dead_peer_gives_an_error_not_a_hang():
    import torch.multiprocessing as mp

    from oracle.gloo_runner import free_port
    from tests import mgpu_dead_peer_worker

    if _world() < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    port = free_port()
    procs = [ctx.Process(target=mgpu_dead_peer_worker.main, args=(r, 2, port)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
    alive = [p.is_alive() for p in procs]
    for p in procs:
        if p.is_alive():
            p.kill()
    assert not any(alive), "a rank hung after its peer died"
    assert [p.exitcode for p in procs] == [0, 0], [p.exitcode for p in procs]
