"""CPU restatement of the collective semantics on nbdistributed's tensor data path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``nbdistributed_b200/`` may import
this module; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs do, and only as the checker.

What is restated.  The reference (a pure-Python IPython extension) has no
collective arithmetic of its own: cell code calls ``dist.all_reduce`` etc. on
the process group that ``ref:src/nbdistributed/worker.py:151`` creates, and
the arithmetic lives in torch's c10d backends (``torch>=1.12.0``, un-pinned,
``ref:pyproject.toml:30``; resolved here to torch 2.11.0+cu128 with bundled
gloo).  The functions below restate the *published semantics* of those
collectives, anchored on the reference's call sites:

* ``all_reduce``      ref:README.md:109            (dist.all_reduce(x))
* ``broadcast``       ref:README.md:122-123        (dist.broadcast(param.data, src=0))
* ``all_gather``      ref:00_accelerate.ipynb:1146 (gather_for_metrics -> all_gather)
* ``reduce_scatter``  BASELINE.json configs[4]     (ZeRO-style flat buffer)
* ``barrier``         ref:src/nbdistributed/worker.py:214

Pinning.  The reference ships no tests or golden vectors, so the
oracle is pinned against outputs of the reference itself run in the build
container: ``tests/golden/make_golden.py`` drives the reference's own
``ProcessManager`` + ``CommunicationManager`` + ``worker.py`` (gloo, N=2 and 4)
and stores the results in ``tests/golden/*.npz``; ``tests/test_oracle_golden.py``
checks this file against them, and ``tests/test_oracle_gloo_live.py`` checks it
against a live gloo group on every run (gloo is present on the GPU box too).

Reduction orders
----------------
``order="rank"``  — the order the CUDA kernels implement: ascending rank
    0..N-1, accumulator in fp32 for bf16/f16 (f64 for f64, wrap-around integers
    for ints), ONE rounding to the tensor dtype at the end.
``order="gloo"``  — gloo's CPU ring as characterised against the live library
    (see ``gloo_chunk_elems``): the tensor is cut into N chunks; chunk c is the
    chain x[s] + x[s-1] + ... (ranks descending mod N, starting at s = c-1),
    rounded to the tensor dtype after every add.  At N=2 both orders coincide
    bit-for-bit (a+b is commutative).

bf16 has no numpy dtype: bf16 arrays are carried as ``uint16`` bit patterns and
``dtype`` is passed by name.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

FLOAT_DTYPES = ("f32", "bf16", "f16", "f64")
INT_DTYPES = ("i32", "i64", "u8", "i8")
NP_DTYPE = {"f32": np.float32, "bf16": np.uint16, "f16": np.float16, "f64": np.float64,
            "i32": np.int32, "i64": np.int64, "u8": np.uint8, "i8": np.int8}
OPS = ("sum", "avg", "prod", "min", "max", "band", "bor", "bxor")


# --------------------------------------------------------------------------- bf16
def bf16_to_f32(u16: np.ndarray) -> np.ndarray:
    """Exact widening of bf16 bit patterns to float32."""
    return (u16.astype(np.uint32) << 16).view(np.float32)


def f32_to_bf16(f32: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even float32 -> bf16 bit patterns (NaN stays NaN)."""
    f32 = np.ascontiguousarray(f32, dtype=np.float32)
    bits = f32.view(np.uint32)
    rounded = ((bits.astype(np.uint64) + 0x7FFF + ((bits >> 16) & 1)) >> 16).astype(np.uint16)
    nan = np.isnan(f32)
    if nan.any():
        rounded = np.where(nan, np.uint16(0x7FC0), rounded)
    return rounded


def _widen(x: np.ndarray, dtype: str) -> np.ndarray:
    """Tensor dtype -> accumulator dtype."""
    if dtype == "bf16":
        return bf16_to_f32(x)
    if dtype == "f16":
        return x.astype(np.float32)
    return x


def _narrow(acc: np.ndarray, dtype: str) -> np.ndarray:
    """Accumulator -> tensor dtype, one round-to-nearest-even."""
    if dtype == "bf16":
        return f32_to_bf16(acc)
    if dtype == "f16":
        return acc.astype(np.float16)
    return acc


def _combine(a: np.ndarray, b: np.ndarray, op: str) -> np.ndarray:
    if op in ("sum", "avg"):
        return a + b  # integers wrap (numpy is modular), floats round to nearest even
    if op == "prod":
        return a * b
    if op == "min":
        if a.dtype.kind == "f":
            return np.where(np.isnan(a), a, np.where(np.isnan(b), b, np.minimum(a, b)))
        return np.minimum(a, b)
    if op == "max":
        if a.dtype.kind == "f":
            return np.where(np.isnan(a), a, np.where(np.isnan(b), b, np.maximum(a, b)))
        return np.maximum(a, b)
    if op == "band":
        return a & b
    if op == "bor":
        return a | b
    if op == "bxor":
        return a ^ b
    raise ValueError(f"unknown op {op}")


def check_op(dtype: str, op: str) -> None:
    """Mirror of the dtype/op validation at the C ABI (and of gloo's: AVG is rejected on ints)."""
    if dtype not in NP_DTYPE:
        raise ValueError(f"unknown dtype {dtype}")
    if op not in OPS:
        raise ValueError(f"unknown op {op}")
    if dtype in FLOAT_DTYPES and op in ("band", "bor", "bxor"):
        raise TypeError("bitwise reduce ops need an integer dtype")
    if dtype in INT_DTYPES and op == "avg":
        raise TypeError("AVG needs a floating-point dtype")


# --------------------------------------------------------------------------- reductions
def reduce_rank_order(inputs: Sequence[np.ndarray], dtype: str, op: str = "sum") -> np.ndarray:
    """acc = x_0; acc = op(acc, x_k) for k = 1..N-1 in the accumulator dtype; round once."""
    check_op(dtype, op)
    with np.errstate(all="ignore"):
        acc = _widen(np.asarray(inputs[0]), dtype).copy()
        for k in range(1, len(inputs)):
            acc = _combine(acc, _widen(np.asarray(inputs[k]), dtype), op)
        if op == "avg":
            acc = acc / acc.dtype.type(len(inputs))
        return _narrow(acc, dtype)


def gloo_chunk_elems(numel: int, elsize: int, world: int) -> int:
    """Elements per ring chunk in gloo's allreduce (published ring algorithm, gloo/allreduce.cc):
    segment = ceil(total_bytes / (2*world)) rounded up to the element size, the number of
    segments rounded up to a multiple of world, chunk = segments_per_rank * segment.
    Characterised against the live library in this container (torch 2.11.0 bundled gloo) for
    4 B .. 256 MiB messages: no segment-size cap is observable, so a chunk is always two
    segments.  Valid when the buffer spans at least ``world`` segments (``gloo_order_is_pinned``)."""
    total = numel * elsize
    if total == 0:
        return 0
    seg = max(1, -(-total // (world * 2)))
    seg = -(-seg // elsize) * elsize
    nseg = -(-total // seg)
    nseg = -(-nseg // world) * world
    return (nseg // world) * seg // elsize


def gloo_order_is_pinned(numel: int, elsize: int, world: int) -> bool:
    """True where reduce_gloo_order was verified bit-exact against live gloo: every rank owns a
    non-empty chunk prefix (tiny buffers with fewer segments than ranks take a path that this
    restatement does not cover; they are checked by tolerance only)."""
    total = numel * elsize
    if total == 0:
        return True
    seg = max(1, -(-total // (world * 2)))
    seg = -(-seg // elsize) * elsize
    return -(-total // seg) >= world


def reduce_gloo_order(inputs: Sequence[np.ndarray], dtype: str, op: str = "sum") -> np.ndarray:
    """gloo ring restatement: see module docstring.  Per-step rounding in the tensor dtype.

    Segments are laid out round-robin over ranks: segment j of the buffer belongs to chunk
    (j mod ... ) as computed by gloo; we follow the characterised rule that element e lives in
    chunk c = (e // chunk_elems) and is reduced along ranks c-1, c-2, ..., c (mod N).
    """
    check_op(dtype, op)
    if op == "avg":
        raise TypeError("gloo does not implement AVG")
    world = len(inputs)
    x = [np.asarray(a).reshape(-1) for a in inputs]
    numel = x[0].size
    out = np.empty_like(x[0])
    elsize = x[0].dtype.itemsize
    ce = gloo_chunk_elems(numel, elsize, world)
    if ce == 0:
        return out.reshape(np.asarray(inputs[0]).shape)
    with np.errstate(all="ignore"):
        for c in range(world):
            lo, hi = c * ce, min((c + 1) * ce, numel)
            if lo >= hi:
                continue
            order = [(c - 1 - k) % world for k in range(world)]
            acc = x[order[0]][lo:hi]
            for r in order[1:]:
                # round to the tensor dtype after every add, as gloo's typed kernels do
                acc = _narrow(_combine(_widen(acc, dtype), _widen(x[r][lo:hi], dtype), op), dtype)
            out[lo:hi] = acc
    return out.reshape(np.asarray(inputs[0]).shape)


def all_reduce(inputs: Sequence[np.ndarray], dtype: str, op: str = "sum", order: str = "rank") -> List[np.ndarray]:
    """dist.all_reduce: every rank ends with the same reduction of all ranks' tensors."""
    if len(inputs) == 1:
        return [np.asarray(inputs[0]).copy()]
    red = reduce_rank_order(inputs, dtype, op) if order == "rank" else reduce_gloo_order(inputs, dtype, op)
    return [red.copy() for _ in inputs]


def reduce_scatter(inputs: Sequence[np.ndarray], dtype: str, op: str = "sum") -> List[np.ndarray]:
    """dist.reduce_scatter_tensor: rank r receives op_k inputs[k][r*n:(r+1)*n] (rank order)."""
    world = len(inputs)
    flat = [np.asarray(a).reshape(-1) for a in inputs]
    if flat[0].size % world:
        raise ValueError("reduce_scatter input must split evenly over ranks")
    n = flat[0].size // world
    return [reduce_rank_order([f[r * n:(r + 1) * n] for f in flat], dtype, op) for r in range(world)]


# --------------------------------------------------------------------------- copies
def broadcast(inputs: Sequence[np.ndarray], root: int) -> List[np.ndarray]:
    """dist.broadcast: every rank ends with root's bytes."""
    if not 0 <= root < len(inputs):
        raise ValueError("root out of range")
    return [np.asarray(inputs[root]).copy() for _ in inputs]


def all_gather(inputs: Sequence[np.ndarray]) -> List[np.ndarray]:
    """dist.all_gather_into_tensor: every rank ends with the concatenation in rank order."""
    cat = np.concatenate([np.asarray(a).reshape(-1) for a in inputs])
    return [cat.copy() for _ in inputs]


# --------------------------------------------------------------------------- tolerances
def sum_abs(inputs: Sequence[np.ndarray], dtype: str) -> np.ndarray:
    acc = np.zeros(np.asarray(inputs[0]).shape, dtype=np.float64)
    for a in inputs:
        acc += np.abs(_widen(np.asarray(a), dtype).astype(np.float64))
    return acc


def float_sum_tolerance(inputs: Sequence[np.ndarray], dtype: str) -> np.ndarray:
    """Element-wise bound on |ours - reference| for floating-point SUM under a different
    (fixed) association, as north_star states it: 1e-6 relative for fp32; for bf16/f16 half an
    ulp (unit roundoff 2^-8 / 2^-11) per rounding step — gloo rounds N-1 times, the kernels
    once — i.e. "ulp-scaled"; all scaled by sum_k |x_k|, the magnitude the partial sums can reach."""
    world = len(inputs)
    scale = sum_abs(inputs, dtype)
    rel = {"f32": 1e-6, "f64": 1e-14, "bf16": world * 2.0 ** -8, "f16": world * 2.0 ** -11}[dtype]
    return rel * scale + np.finfo(np.float32).tiny


def to_f64(x: np.ndarray, dtype: str) -> np.ndarray:
    return _widen(np.asarray(x), dtype).astype(np.float64)


# --------------------------------------------------------------------------- seeded inputs
def make_inputs(world: int, numel: int, dtype: str, seed: int = 1234, scale: float = 1.0) -> List[np.ndarray]:
    """The synthetic inputs of SURVEY.md §8(d): per-rank generator seeded with seed + rank;
    N(0,1) floats, integers uniform in [-2^30, 2^30) (full range for 8-bit)."""
    out = []
    for r in range(world):
        g = np.random.default_rng(seed + r)
        if dtype in ("f32", "f64", "f16"):
            out.append((g.standard_normal(numel) * scale).astype(NP_DTYPE[dtype]))
        elif dtype == "bf16":
            out.append(f32_to_bf16((g.standard_normal(numel) * scale).astype(np.float32)))
        elif dtype in ("i32", "i64"):
            out.append(g.integers(-2 ** 30, 2 ** 30, size=numel, dtype=np.int64).astype(NP_DTYPE[dtype]))
        elif dtype == "u8":
            out.append(g.integers(0, 256, size=numel, dtype=np.int64).astype(np.uint8))
        elif dtype == "i8":
            out.append(g.integers(-128, 128, size=numel, dtype=np.int64).astype(np.int8))
        else:
            raise ValueError(dtype)
    return out
