"""The c10d plugin on the single-GPU box: registration, world_size == 1 through
torch.distributed (every collective degenerates to a local copy), error behaviour."""
import os

import pytest

pytestmark = pytest.mark.gpu


def _run_world1(port):
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    ns = W.init_worker(0, 1, "127.0.0.1", port, gpu_id=0, backend="b200")
    try:
        assert dist.get_backend() in ("cpu:gloo,cuda:b200", "b200") or "b200" in str(dist.get_backend())
        x = torch.arange(1000, device="cuda:0", dtype=torch.float32)
        y = x.clone()
        dist.all_reduce(y)
        assert torch.equal(x, y)
        dist.broadcast(y, src=0)
        out = torch.empty_like(x)
        dist.all_gather_into_tensor(out, x)
        assert torch.equal(out, x)
        out2 = torch.empty_like(x)
        dist.reduce_scatter_tensor(out2, x)
        assert torch.equal(out2, x)
        dist.barrier()
        be = B.get_backend()
        assert type(be).__name__ == "BackendB200"
        assert be.device_index() == 0
        assert ns["device"] == 0
    finally:
        W.shutdown_worker()


def test_world_size_one_through_torch_distributed():
    from oracle.gloo_runner import free_port

    _run_world1(free_port())
