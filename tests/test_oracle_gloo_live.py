"""The oracle against a LIVE gloo group (the reference's CPU path), every run, N = 2, 4 and 8."""
import numpy as np
import pytest

from oracle import collectives_oracle as O
from oracle.gloo_runner import run_gloo
from tests import helpers as H


@pytest.mark.parametrize("world", [2, 4, 8])
def test_live_gloo_matches_oracle(world):
    jobs, meta = [], []
    for dtype, n in (("f32", 1), ("f32", 4096 + 5), ("bf16", 65536 + 13), ("f16", 1000), ("f64", 333), ("i64", 4097),
                     ("i32", 10), ("u8", 257)):
        x = O.make_inputs(world, n, dtype)
        jobs.append({"kind": "all_reduce", "dtype": dtype, "op": "sum", "inputs": x})
        meta.append(("all_reduce", dtype, x, None))
    x = O.make_inputs(world, 1001, "u8", seed=3)
    jobs.append({"kind": "broadcast", "dtype": "u8", "root": world - 1, "inputs": x})
    meta.append(("broadcast", "u8", x, world - 1))
    x = O.make_inputs(world, 77, "f32", seed=4)
    jobs.append({"kind": "all_gather", "dtype": "f32", "inputs": x})
    meta.append(("all_gather", "f32", x, None))
    x = O.make_inputs(world, world * 50, "i32", seed=5)
    jobs.append({"kind": "reduce_scatter", "dtype": "i32", "inputs": x})
    meta.append(("reduce_scatter", "i32", x, None))
    x = O.make_inputs(world, 999, "f32", seed=6)
    jobs.append({"kind": "all_reduce", "dtype": "f32", "op": "max", "inputs": x})
    meta.append(("all_reduce_max", "f32", x, None))
    jobs.append({"kind": "barrier"})
    meta.append(("barrier", None, None, None))

    res = run_gloo(world, jobs)
    for j, (kind, dtype, x, root) in enumerate(meta):
        if kind == "barrier":
            continue
        got = [res[r][j] for r in range(world)]
        if kind == "all_reduce":
            for r in range(1, world):  # gloo leaves identical bits on every rank
                H.assert_same(got[r], got[0], dtype, f"job {j}: ranks differ")
            if dtype in O.INT_DTYPES or world == 2:
                H.assert_same(got[0], O.reduce_rank_order(x, dtype), dtype, f"job {j} rank order")
            if dtype in O.FLOAT_DTYPES:
                n, es = x[0].size, x[0].dtype.itemsize
                if O.gloo_order_is_pinned(n, es, world):
                    H.assert_same(got[0], O.reduce_gloo_order(x, dtype), dtype, f"job {j} gloo order")
                err = np.abs(O.to_f64(got[0], dtype) - O.to_f64(O.reduce_rank_order(x, dtype), dtype))
                assert (err <= O.float_sum_tolerance(x, dtype)).all(), f"job {j}: rank order outside tolerance"
        elif kind == "all_reduce_max":
            H.assert_same(got[0], O.reduce_rank_order(x, dtype, "max"), dtype, f"job {j} max")
        elif kind == "broadcast":
            for r in range(world):
                H.assert_same(got[r], O.broadcast(x, root)[r], dtype, f"job {j} rank {r}")
        elif kind == "all_gather":
            for r in range(world):
                H.assert_same(got[r], O.all_gather(x)[r], dtype, f"job {j} rank {r}")
        elif kind == "reduce_scatter":
            want = O.reduce_scatter(x, dtype)
            for r in range(world):
                H.assert_same(got[r], want[r], dtype, f"job {j} rank {r}")
