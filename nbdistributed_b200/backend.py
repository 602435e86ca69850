"""Registration of the ``b200`` backend with torch.distributed.

``register()`` is what a worker process calls before ``init_process_group``
(ref:src/nbdistributed/worker.py:151): it loads the c10d glue (``_c10d_b200.so``,
which in turn loads the C-ABI library with the sm_100a kernels) and registers
its creator through torch's backend-extension API
(``torch.distributed.Backend.register_backend``,
$TORCH/distributed/distributed_c10d.py:328-390).  Afterwards
``dist.init_process_group("b200")`` or ``"cpu:gloo,cuda:b200"`` selects it.

No fallback: if the native libraries are missing this raises ImportError.
"""
from __future__ import annotations

import os

from . import BACKEND_NAME

_HERE = os.path.dirname(os.path.abspath(__file__))
_registered = False


def _load_ext():
    import torch  # noqa: F401  (libtorch must be loaded before the extension)

    so = os.path.join(_HERE, "_c10d_b200.so")
    if not os.path.exists(so) or not os.path.exists(os.path.join(_HERE, "libnbdist_b200.so")):
        raise ImportError(
            f"{so} (or libnbdist_b200.so) is missing: build with "
            "`python -c 'import __graft_entry__ as g; g.build()'`. "
            "The b200 backend has no CPU or NCCL fallback.")
    from . import _c10d_b200

    return _c10d_b200


def register(name: str = BACKEND_NAME) -> str:
    """Register the backend under ``name`` (idempotent). Returns the name."""
    global _registered
    import torch.distributed as dist

    ext = _load_ext()
    if _registered or hasattr(dist.Backend, name.upper()):
        _registered = True
        return name
    dist.Backend.register_backend(name, ext.create_backend, extended_api=False, devices=["cuda"])
    _registered = True
    return name


def backend_string(name: str = BACKEND_NAME, with_cpu: bool = True) -> str:
    """The string a notebook passes instead of "nccl": CPU tensors keep going through gloo
    (object collectives, the CPU oracle side by side), CUDA tensors through this backend."""
    return f"cpu:gloo,cuda:{name}" if with_cpu else name


def get_backend(group=None):
    """The BackendB200 object behind a process group (tunables, last_algo, ...)."""
    import torch
    import torch.distributed as dist

    pg = group if group is not None else dist.group.WORLD
    return pg._get_backend(torch.device("cuda"))


def empty_symmetric(numel: int, dtype=None, group=None):
    """A 1-D CUDA tensor in symmetric memory (mapped by every rank of ``group``): collectives on
    it — or on views of it, at the same offset on every rank — run zero-copy over NVLink instead
    of through the staging arena.  COLLECTIVE: every rank must call it in the same order with
    the same size (it is ``Backend.allocate_tensor``, which FSDP also uses).  Keep the process
    group alive for as long as the tensor is used."""
    import torch

    be = get_backend(group)
    return be.allocate_tensor(int(numel), dtype=dtype or torch.float32, device=torch.device("cuda", torch.cuda.current_device()))


def all_reduce_host(tensor, out=None, op: int = 0, group=None):
    """all_reduce of a CPU tensor (pinned memory recommended) through the GPUs of ``group``.

    The data stays host-resident for the caller: the library streams it through the GPU in
    chunks, overlapping the H2D copy, the NVLink collective and the D2H copy.  ``op`` is the
    c10d ReduceOp number (0 = SUM).  In place unless ``out`` is given.  Blocks until done."""
    be = get_backend(group)
    be.all_reduce_host(tensor, tensor if out is None else out, int(op))
    return tensor if out is None else out


def broadcast_tensors(tensors, src: int = 0, group=None) -> None:
    """``for t in tensors: dist.broadcast(t, src)`` (the reference's parameter-sync idiom,
    ref:README.md:122-123) as one call: tensors of any sizes and dtypes travel as one pipelined
    launch per 256 tensors.  ``src`` is the global rank, as in ``dist.broadcast``.  Stream-ordered
    on the current CUDA stream; the same result as the loop, bit for bit."""
    import torch.distributed as dist

    be = get_backend(group)
    root = dist.get_group_rank(group, src) if group is not None else int(src)
    be.broadcast_many(list(tensors), root)


def launch_count() -> int:
    """Kernels launched by the native library in this process (bench.py's gpu_launches)."""
    return int(_load_ext().launch_count())
