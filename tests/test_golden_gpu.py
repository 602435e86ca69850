"""The CUDA kernels against the reference's own outputs (golden fixtures), on one GPU with
emulated ranks: bit-exact for copies, integers, order-free ops and any N = 2 float sum; within
the north_star tolerance for N = 4 float sums (gloo's ring order differs from the kernels'
fixed rank order)."""
import numpy as np
import pytest

from oracle import collectives_oracle as O
from tests import golden_cases as G
from tests import helpers as H

pytestmark = pytest.mark.gpu
OPCODE = {"sum": 0, "avg": 1, "prod": 2, "min": 3, "max": 4}


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("case", G.CASES, ids=[c.name for c in G.CASES])
def test_cuda_reproduces_reference_outputs(world, case):
    import torch

    from nbdistributed_b200 import capi

    data = G.load(world)
    ins, ref = G.inputs_and_output(data, case, world)
    flat = [np.ascontiguousarray(a.reshape(-1)) for a in ins]
    ref = ref.reshape(-1)
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        t = [H.to_cuda(a, case.dtype) for a in flat]
        if case.kind == "broadcast":
            root = case.root if case.root >= 0 else world - 1
            grp.broadcast(t, root=root)
            out = t
        elif case.kind == "all_gather":
            out = [torch.empty(world * t[0].numel(), dtype=t[0].dtype, device="cuda:0") for _ in range(world)]
            grp.all_gather(t, out)
        else:
            grp.all_reduce(t, t, op=OPCODE[case.op])
            out = t
        torch.cuda.synchronize()
        grp.async_error()
        got = [H.to_numpy(o, case.dtype) for o in out]
    for r in range(1, world):
        H.assert_same(got[r], got[0], case.dtype, f"{case.name}: rank {r} differs from rank 0")
    exact = (case.kind != "all_reduce" or case.dtype in O.INT_DTYPES or case.op in ("max", "min") or world == 2)
    if exact:
        H.assert_same(got[0], ref, case.dtype, f"{case.name} vs reference")
    elif case.op == "sum":
        err = np.abs(O.to_f64(got[0], case.dtype) - O.to_f64(ref, case.dtype))
        tol = O.float_sum_tolerance(flat, case.dtype)
        assert (err <= tol).all(), f"{case.name}: max err/tol {np.max(err / tol)}"
    else:  # product at N = 4: relative, a few fp32 ulps
        rel = np.abs(O.to_f64(got[0], case.dtype) - O.to_f64(ref, case.dtype)) / (np.abs(O.to_f64(ref, case.dtype)) + 1e-30)
        assert rel.max() < 1e-6
