"""Host-side mirror of the reference's backend-selection seam.

``ref:src/nbdistributed/worker.py:128-151`` sets the rendezvous environment,
picks the GPU and hard-codes ``backend = "nccl" if cuda else "gloo"`` before
``dist.init_process_group``; ``:160-177`` then injects ``torch``, ``dist``,
``rank``, ``world_size``, ``gpu_id`` and ``device`` into the namespace cell code
runs in.  ``init_worker`` does the same with ONE addition: the backend string
may be chosen (``%dist_init --backend`` -> ``ProcessManager.start_workers(backend=)``
-> optional 7th argv slot / ``NBDIST_BACKEND`` env).  With no
choice the behaviour is the reference's, so existing notebooks are unaffected.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Optional

from . import BACKEND_NAME

ENV_BACKEND = "NBDIST_BACKEND"


def select_backend(requested: Optional[str], cuda_available: bool) -> str:
    """The reference's rule (worker.py:135-147) unless a backend was requested.

    ``requested`` may be a plain name (``"b200"``, ``"nccl"``, ``"gloo"``) or a
    device map (``"cpu:gloo,cuda:b200"``).  ``"b200"`` alone is widened to the
    device map so CPU tensors keep working through gloo.
    """
    req = (requested or os.environ.get(ENV_BACKEND) or "").strip()
    if not req:
        return "nccl" if cuda_available else "gloo"
    if req == BACKEND_NAME:
        if not cuda_available:
            raise RuntimeError("backend 'b200' needs CUDA GPUs; there is no CPU fallback (use 'gloo')")
        return f"cpu:gloo,cuda:{BACKEND_NAME}"
    return req


def wants_b200(backend: str) -> bool:
    return any(part.split(":")[-1].strip() == BACKEND_NAME for part in backend.split(","))


def init_worker(rank: int, world_size: int, master_addr: str, master_port: str | int,
                gpu_id: Optional[int] = None, backend: Optional[str] = None,
                timeout=None) -> Dict[str, Any]:
    """Environment + device + process group, then the namespace the reference injects.

    Mirrors DistributedWorker.__init__ (ref:src/nbdistributed/worker.py:94-179) minus the
    ZMQ socket, which is control plane and out of scope here.
    """
    import torch
    import torch.distributed as dist

    os.environ["RANK"] = str(rank)
    os.environ["LOCAL_RANK"] = str(rank)  # sic: the reference sets LOCAL_RANK = rank
    os.environ["WORLD_SIZE"] = str(world_size)
    os.environ["MASTER_ADDR"] = master_addr
    os.environ["MASTER_PORT"] = str(master_port)

    cuda = torch.cuda.is_available()
    device_id = None
    if cuda:
        device_id = gpu_id if gpu_id is not None else rank % torch.cuda.device_count()
        torch.cuda.set_device(device_id)
    chosen = select_backend(backend, cuda)
    if wants_b200(chosen):
        from . import backend as _backend

        _backend.register()
    kwargs = {}
    if timeout is not None:
        kwargs["timeout"] = timeout
    dist.init_process_group(backend=chosen, rank=rank, world_size=world_size, **kwargs)

    namespace: Dict[str, Any] = {
        "torch": torch,
        "dist": dist,
        "rank": rank,
        "world_size": world_size,
        "__rank__": rank,
        "__world_size__": world_size,
        "__backend__": chosen,
    }
    if cuda:
        namespace["gpu_id"] = device_id
        namespace["device"] = torch.cuda.current_device()  # an int, as in the reference
    else:
        namespace["gpu_id"] = None
        namespace["device"] = torch.device("cpu")
    return namespace


def run_cell(namespace: Dict[str, Any], code: str) -> Any:
    """Execute cell text the way the worker does: exec the body, eval a trailing expression
    (ref:src/nbdistributed/worker.py:282-367, simplified: no stdout streaming)."""
    import ast

    tree = ast.parse(code, mode="exec")
    last = None
    if tree.body and isinstance(tree.body[-1], ast.Expr):
        last = ast.Expression(tree.body.pop().value)
    exec(compile(tree, "<cell>", "exec"), namespace)
    if last is not None:
        return eval(compile(last, "<cell>", "eval"), namespace)
    return None


def shutdown_worker() -> None:
    """ref:src/nbdistributed/worker.py:578."""
    import torch.distributed as dist

    if dist.is_initialized():
        dist.destroy_process_group()
