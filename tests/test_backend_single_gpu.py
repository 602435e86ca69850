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


def _run_world1_cuda_only(port):
    """Backend string without a CPU half: dist.barrier() is then routed to BackendB200::barrier
    (with "cpu:gloo,cuda:b200" Python sends a device-less barrier to gloo)."""
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    ns = W.init_worker(0, 1, "127.0.0.1", port, gpu_id=0, backend="cuda:b200")
    try:
        assert ns["__backend__"] == "cuda:b200"
        dist.barrier()                      # -> our backend (current CUDA device)
        dist.barrier(device_ids=[0])
        x = torch.ones(4096, device="cuda:0")
        dist.all_reduce(x)
        assert float(x.sum()) == 4096.0
        h = torch.arange(1 << 20, dtype=torch.float32).pin_memory()
        out = torch.empty_like(h).pin_memory()
        B.all_reduce_host(h, out=out)       # host-buffer pipeline at world size 1: a copy through the GPU
        assert torch.equal(out, h)
        sym = B.empty_symmetric(1 << 16, torch.bfloat16)
        sym.fill_(2.0)
        dist.all_reduce(sym)
        assert float(sym[0]) == 2.0
        del sym
    finally:
        W.shutdown_worker()


def test_cuda_only_backend_string_routes_barrier_to_b200():
    import multiprocessing as mp

    from oracle.gloo_runner import free_port

    # a fresh process: the default process group of this interpreter was already used above
    ctx = mp.get_context("spawn")
    p = ctx.Process(target=_run_world1_cuda_only, args=(free_port(),))
    p.start()
    p.join(timeout=180)
    assert p.exitcode == 0


def test_topology_discovery_on_real_hardware():
    """topology.discover() through NVML on the GPU box (the fake-matrix CPU test covers the logic):
    square matrix over every visible GPU, every pair reachable peer to peer on an NVSwitch box,
    and check_assignment accepts one rank per GPU (ref:src/nbdistributed/process_manager.py:104-115
    assigns GPUs blindly)."""
    import torch

    from nbdistributed_b200 import topology

    n = torch.cuda.device_count()
    topo = topology.discover()
    assert topo["source"] == "nvml", topo["source"]  # pynvml ships in the image: no torch fallback needed
    assert len(topo["gpus"]) == n and len(topo["p2p"]) == n and all(len(row) == n for row in topo["p2p"])
    assert all("B200" in g["name"] for g in topo["gpus"]), [g["name"] for g in topo["gpus"]]
    assert all(g["nvlinks_up"] == 18 for g in topo["gpus"]), [g["nvlinks_up"] for g in topo["gpus"]]
    if n >= 2:
        assert topo["all_p2p"] and topo["all_nvlink"], topology.format_matrix(topo)
    rep = topology.check_assignment(n, list(range(n)), topo=topo)
    assert rep["map"] == list(range(n)) and rep["warnings"] == []
    text = topology.format_matrix(topo)
    assert "GPU0" in text and "source: nvml" in text
    with pytest.raises(ValueError):
        topology.check_assignment(2, [0, 0], topo=topo)  # two ranks on one GPU
