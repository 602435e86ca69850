"""Access to the golden fixtures produced by the reference itself (tests/golden/make_golden.py)."""
from __future__ import annotations

import os
from typing import Dict, List, NamedTuple

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


class Case(NamedTuple):
    name: str
    kind: str      # all_reduce | broadcast | all_gather
    dtype: str     # oracle dtype name
    op: str
    root: int      # broadcast only; -1 = world-1


CASES = [
    Case("readme_allreduce", "all_reduce", "f32", "sum", 0),
    Case("linear_weight", "broadcast", "f32", "", 0),
    Case("linear_bias", "broadcast", "f32", "", 0),
    Case("ar_bf16", "all_reduce", "bf16", "sum", 0),
    Case("ar_f16", "all_reduce", "f16", "sum", 0),
    Case("ar_f64", "all_reduce", "f64", "sum", 0),
    Case("ar_i64", "all_reduce", "i64", "sum", 0),
    Case("ar_i32", "all_reduce", "i32", "sum", 0),
    Case("ar_max_f32", "all_reduce", "f32", "max", 0),
    Case("ar_prod_f32", "all_reduce", "f32", "prod", 0),
    Case("bc_u8", "broadcast", "u8", "", -1),
    Case("ag_f32", "all_gather", "f32", "", 0),
    Case("ar_odd_f32", "all_reduce", "f32", "sum", 0),
]


def load(world: int) -> Dict[str, np.ndarray]:
    with np.load(os.path.join(GOLDEN_DIR, f"ref_gloo_n{world}.npz")) as z:
        return {k: z[k] for k in z.files}


def inputs_and_output(data: Dict[str, np.ndarray], case: Case, world: int):
    """Per-rank inputs and the (rank-identical) output of the reference; bf16 as uint16 bits."""
    ins: List[np.ndarray] = [data[f"{case.name}__in_r{r}"] for r in range(world)]
    out = data[f"{case.name}__out"]
    if case.dtype == "bf16":
        ins = [a.view(np.uint16) for a in ins]
        out = out.view(np.uint16)
    return ins, out
