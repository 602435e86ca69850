"""Real ranks (one process per GPU) through the C ABI only — no torch.distributed."""
import os

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def test_real_ranks_through_c_abi():
    import torch
    import torch.multiprocessing as mp

    from tests import capi_mp_worker

    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    want = int(os.environ.get("NBD_TEST_WORLD", "0"))
    world = min(n, want) if want else min(n, 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    qs = [ctx.Queue() for _ in range(world)]
    result_q = ctx.Queue()
    procs = [ctx.Process(target=capi_mp_worker.main, args=(r, world, qs, result_q)) for r in range(world)]
    for p in procs:
        p.start()
    errors = []
    try:
        for _ in range(world):
            rank, err = result_q.get(timeout=600)
            if err:
                errors.append(f"rank {rank}: {err}")
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    assert not errors, "\n".join(errors)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
