"""Live gloo oracle: the reference's real CPU path.

TEST INFRASTRUCTURE ONLY (see collectives_oracle.py).  When CUDA is absent the
reference selects ``backend="gloo"`` (ref:src/nbdistributed/worker.py:146-147)
and runs the very same cell code on CPU tensors.  This module starts
``world`` local processes with the same environment contract as worker.py
(``RANK/LOCAL_RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT``,
ref:src/nbdistributed/worker.py:128-132), runs a list of collective jobs on
given per-rank inputs and returns every rank's outputs — for parity tests — or
timings — for bench.py's ``cpu_baseline`` / ``--impl reference`` legs.
"""
from __future__ import annotations

import os
import socket
import time
from typing import Any, Dict, List

import numpy as np


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _to_torch(a: np.ndarray, dtype: str):
    import torch

    if dtype == "bf16":
        return torch.from_numpy(np.ascontiguousarray(a).view(np.int16).copy()).view(torch.bfloat16)
    return torch.from_numpy(np.ascontiguousarray(a).copy())


def _to_numpy(t, dtype: str) -> np.ndarray:
    import torch

    if dtype == "bf16":
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


def _reduce_op(name: str):
    import torch.distributed as dist

    return {"sum": dist.ReduceOp.SUM, "avg": dist.ReduceOp.AVG, "prod": dist.ReduceOp.PRODUCT,
            "min": dist.ReduceOp.MIN, "max": dist.ReduceOp.MAX, "band": dist.ReduceOp.BAND,
            "bor": dist.ReduceOp.BOR, "bxor": dist.ReduceOp.BXOR}[name]


def _run_job(job: Dict[str, Any], rank: int, world: int):
    import torch
    import torch.distributed as dist

    kind, dtype = job["kind"], job.get("dtype", "f32")
    x = _to_torch(job["input"], dtype) if job.get("input") is not None else None
    if kind == "all_reduce":
        dist.all_reduce(x, op=_reduce_op(job.get("op", "sum")))
        return _to_numpy(x, dtype)
    if kind == "broadcast":
        dist.broadcast(x, src=job.get("root", 0))
        return _to_numpy(x, dtype)
    if kind == "all_gather":
        out = torch.empty(world * x.numel(), dtype=x.dtype)
        dist.all_gather_into_tensor(out, x)
        return _to_numpy(out, dtype)
    if kind == "reduce_scatter":
        # gloo has no reduce_scatter_tensor; the reference's CPU path would do
        # all_reduce + slice (what a notebook user falls back to) — same result set.
        y = x.clone()
        dist.all_reduce(y, op=_reduce_op(job.get("op", "sum")))
        n = x.numel() // world
        return _to_numpy(y[rank * n:(rank + 1) * n].clone(), dtype)
    if kind == "barrier":
        dist.barrier()
        return None
    raise ValueError(kind)


def _time_job(job: Dict[str, Any], rank: int, world: int) -> float:
    """Seconds per iteration of one collective on CPU tensors, barrier-fenced, max over ranks."""
    import torch
    import torch.distributed as dist

    kind, dtype = job["kind"], job.get("dtype", "f32")
    numel = int(job["numel"])
    tdt = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16, "f64": torch.float64,
           "i32": torch.int32, "i64": torch.int64}[dtype]
    g = torch.Generator().manual_seed(1234 + rank)
    x = torch.randn(numel, generator=g).to(tdt) if tdt.is_floating_point else torch.randint(-2 ** 30, 2 ** 30, (numel,), generator=g).to(tdt)
    out = torch.empty(world * numel, dtype=tdt) if kind == "all_gather" else None

    def step():
        if kind == "all_reduce":
            dist.all_reduce(x)
        elif kind == "broadcast":
            dist.broadcast(x, src=job.get("root", 0))
        elif kind == "all_gather":
            dist.all_gather_into_tensor(out, x)
        elif kind == "reduce_scatter":
            dist.all_reduce(x)
        else:
            raise ValueError(kind)

    for _ in range(int(job.get("warmup", 1))):
        step()
    dist.barrier()
    t0 = time.perf_counter()
    iters = int(job.get("iters", 3))
    for _ in range(iters):
        step()
    dist.barrier()
    dt = torch.tensor([(time.perf_counter() - t0) / iters], dtype=torch.float64)
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item())


def _worker(rank: int, world: int, port: int, jobs: List[Dict[str, Any]], mode: str, q) -> None:
    # a parent launched by torchrun exports TORCHELASTIC_* (agent store, restart counts, ...);
    # this group is its own rendezvous, exactly like a worker.py process
    for k in [k for k in os.environ if k.startswith("TORCHELASTIC_") or k in ("GROUP_RANK", "ROLE_RANK", "LOCAL_WORLD_SIZE",
                                                                              "ROLE_WORLD_SIZE", "GROUP_WORLD_SIZE")]:
        os.environ.pop(k, None)
    os.environ["RANK"] = str(rank)
    os.environ["LOCAL_RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    try:
        import torch
        import torch.distributed as dist

        threads = int(os.environ.get("NBD_GLOO_THREADS", "0"))
        if threads > 0:
            torch.set_num_threads(threads)
        dist.init_process_group(backend="gloo", rank=rank, world_size=world)
        res = []
        for job in jobs:
            res.append(_run_job(job, rank, world) if mode == "run" else _time_job(job, rank, world))
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, res, None))
    except Exception as e:  # pragma: no cover - surfaced to the parent
        import traceback

        q.put((rank, None, f"{e}\n{traceback.format_exc()}"))


def _spawn(world: int, per_rank_jobs: List[List[Dict[str, Any]]], mode: str, timeout: float):
    import multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, per_rank_jobs[r], mode, q), daemon=True)
             for r in range(world)]
    for p in procs:
        p.start()
    results: Dict[int, Any] = {}
    deadline = time.time() + timeout
    try:
        while len(results) < world:
            remaining = deadline - time.time()
            if remaining <= 0:
                raise TimeoutError("gloo oracle timed out")
            rank, res, err = q.get(timeout=remaining)
            if err:
                raise RuntimeError(f"gloo oracle rank {rank} failed: {err}")
            results[rank] = res
    finally:
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world)]


def run_gloo(world: int, jobs: List[Dict[str, Any]], timeout: float = 300.0) -> List[List[Any]]:
    """Run ``jobs`` on a fresh gloo group of ``world`` local processes.

    Each job: ``{"kind": "all_reduce"|"broadcast"|"all_gather"|"reduce_scatter"|"barrier",
    "dtype": name, "op": name, "root": int, "inputs": [per-rank np arrays]}``.
    Returns ``out[rank][job]`` (np arrays; bf16 as uint16 bit patterns).
    """
    per_rank = []
    for r in range(world):
        per_rank.append([{**{k: v for k, v in j.items() if k != "inputs"},
                          "input": (j["inputs"][r] if j.get("inputs") is not None else None)} for j in jobs])
    return _spawn(world, per_rank, "run", timeout)


def time_gloo(world: int, jobs: List[Dict[str, Any]], timeout: float = 900.0) -> List[float]:
    """Seconds per iteration (max over ranks) for each job ``{"kind", "dtype", "numel", "iters", "warmup"}``."""
    per_rank = [list(jobs) for _ in range(world)]
    out = _spawn(world, per_rank, "time", timeout)
    return list(out[0])
