"""Host-side logic on CPU: the backend-selection seam (mirror of worker.py:128-151), the
plugin registration, the store leg of the handle exchange, topology checks, and the N > 1
path with world_size-2 gloo (BASELINE config 1: the README example)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_select_backend_keeps_reference_default():
    from nbdistributed_b200 import worker_init as W

    os.environ.pop(W.ENV_BACKEND, None)
    assert W.select_backend(None, True) == "nccl"      # ref:src/nbdistributed/worker.py:145
    assert W.select_backend(None, False) == "gloo"     # ref:src/nbdistributed/worker.py:147
    assert W.select_backend("b200", True) == "cpu:gloo,cuda:b200"
    assert W.select_backend("cpu:gloo,cuda:b200", True) == "cpu:gloo,cuda:b200"
    assert W.select_backend("nccl", True) == "nccl"
    with pytest.raises(RuntimeError):
        W.select_backend("b200", False)
    assert W.wants_b200("cpu:gloo,cuda:b200") and W.wants_b200("b200") and not W.wants_b200("nccl")
    os.environ[W.ENV_BACKEND] = "b200"
    try:
        assert W.select_backend(None, True) == "cpu:gloo,cuda:b200"
    finally:
        os.environ.pop(W.ENV_BACKEND, None)


def test_plugin_registers_and_refuses_to_run_without_cuda():
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B

    assert B.register() == "b200"
    assert B.register() == "b200"  # idempotent
    assert "b200" in dist.Backend.backend_list
    assert B.backend_string() == "cpu:gloo,cuda:b200"
    if torch.cuda.is_available():
        return
    ext = B._load_ext()
    store = dist.HashStore()
    import datetime

    with pytest.raises(Exception) as ei:
        ext.create_backend(store, 0, 1, datetime.timedelta(seconds=5))
    assert "CUDA" in str(ei.value) or "cuda" in str(ei.value)


def test_store_leg_of_the_handle_exchange():
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from oracle.gloo_runner import free_port

    ext = B._load_ext()
    tcp = dist.TCPStore("127.0.0.1", free_port(), 1, True)
    prefixed = dist.PrefixStore("cuda/", dist.PrefixStore("default_pg/", tcp))
    assert ext.store_roundtrip(prefixed, "nbd_b200/handle/0", 256)
    assert ext.nbd_version() == 100 and ext.launch_count() >= 0


def test_topology_assignment_rules():
    from nbdistributed_b200 import topology as T

    n = 4
    full = [[True] * n for _ in range(n)]
    topo = {"source": "fake", "gpus": [{"index": i} for i in range(n)], "p2p": full, "nvlink": full}
    assert T.check_assignment(4, None, topo)["map"] == [0, 1, 2, 3]
    assert T.check_assignment(2, [3, 1], topo)["map"] == [3, 1]
    with pytest.raises(ValueError, match="distinct"):
        T.check_assignment(4, [0, 1], topo)          # the reference would cycle gpu_ids: two ranks per GPU
    with pytest.raises(ValueError, match="at most 8"):
        T.check_assignment(9, None, topo)
    no_p2p = [[a == b or (a < 2) == (b < 2) for b in range(n)] for a in range(n)]
    with pytest.raises(ValueError, match="P2P"):
        T.check_assignment(4, None, {**topo, "p2p": no_p2p})
    no_nvl = [[a == b for b in range(n)] for a in range(n)]
    rep = T.check_assignment(2, None, {**topo, "nvlink": no_nvl})
    assert rep["warnings"] and "NVLink" in rep["warnings"][0]
    assert T.rank_gpu_map(3, None, 2) == [0, 1, 0]  # ref:src/nbdistributed/worker.py:141-143
    assert "NV" in T.format_matrix(topo)
    assert isinstance(T.discover(), dict)


def _readme_rank(rank, world, port, q):
    try:
        from nbdistributed_b200 import worker_init as W

        ns = W.init_worker(rank, world, "127.0.0.1", port)  # no GPU here -> "gloo", as in the reference
        assert ns["__backend__"] == "gloo" and ns["rank"] == rank and ns["world_size"] == world
        assert ns["gpu_id"] is None and str(ns["device"]) == "cpu"
        W.run_cell(ns, "import torch\ntorch.manual_seed(1234 + rank)\nx = torch.randn(100, 100)\nxin = x.clone()\n"
                       "dist.all_reduce(x)\n")
        mean = W.run_cell(ns, "x.mean().item()")
        W.run_cell(ns, "torch.manual_seed(100 + rank)\nmodel = torch.nn.Linear(10, 10)\n"
                       "for param in model.parameters():\n    dist.broadcast(param.data, src=0)\n")
        W.run_cell(ns, "dist.barrier()")
        q.put((rank, ns["xin"].numpy(), ns["x"].numpy(), mean, ns["model"].weight.data.numpy(), None))
        W.shutdown_worker()
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, None, None, None, None, f"{e}\n{traceback.format_exc()}"))


def test_readme_example_world_size_2_gloo_matches_golden_and_oracle():
    """BASELINE config 1 through init_worker + run_cell; same seeds as the golden fixture that
    the reference's own worker.py produced, so outputs must be bit-identical to it."""
    import multiprocessing as mp

    import torch

    if torch.cuda.is_available():
        pytest.skip("CPU plumbing test (the GPU variant is tests/test_backend_multigpu.py)")
    from oracle import collectives_oracle as O
    from oracle.gloo_runner import free_port
    from tests import golden_cases as G

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_readme_rank, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        rank, xin, x, mean, w, err = q.get(timeout=180)
        assert err is None, err
        got[rank] = (xin, x, mean, w)
    for p in procs:
        p.join(timeout=60)
    data = G.load(2)
    for r in range(2):
        assert np.array_equal(got[r][0], data[f"readme_allreduce__in_r{r}"])
        assert np.array_equal(got[r][1], data["readme_allreduce__out"])
        assert np.array_equal(got[r][3], data["linear_weight__out"])
    assert got[0][2] == got[1][2]
    want = O.reduce_rank_order([got[0][0], got[1][0]], "f32")
    assert np.array_equal(got[0][1], want)


def test_product_code_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under nbdistributed_b200/ may import it."""
    bad = []
    pkg = os.path.join(ROOT, "nbdistributed_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M) or "oracle/" in text:
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad
