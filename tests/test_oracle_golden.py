"""The oracle against the reference's own outputs (golden fixtures, gloo path, N = 2, 4 and 8)."""
import numpy as np
import pytest

from oracle import collectives_oracle as O
from tests import golden_cases as G
from tests import helpers as H


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("case", G.CASES, ids=[c.name for c in G.CASES])
def test_oracle_reproduces_reference_outputs(world, case):
    data = G.load(world)
    assert int(data["__world__"][0]) == world
    ins, ref = G.inputs_and_output(data, case, world)
    flat = [a.reshape(-1) for a in ins]
    ref = ref.reshape(-1)
    if case.kind == "broadcast":
        root = case.root if case.root >= 0 else world - 1
        H.assert_same(O.broadcast(flat, root)[0], ref, case.dtype, case.name)
        return
    if case.kind == "all_gather":
        H.assert_same(O.all_gather(flat)[0], ref, case.dtype, case.name)
        return
    rank_order = O.reduce_rank_order(flat, case.dtype, case.op)
    is_float = case.dtype in O.FLOAT_DTYPES
    if not is_float or case.op in ("max", "min") or world == 2:
        # integers, order-free ops, and any two-rank float sum: bit-exact with the reference
        H.assert_same(rank_order, ref, case.dtype, f"{case.name} (rank order)")
    if is_float and case.op == "sum":
        # gloo's ring order restated: bit-exact with the reference at every world size
        elsize = flat[0].dtype.itemsize
        assert O.gloo_order_is_pinned(flat[0].size, elsize, world)
        H.assert_same(O.reduce_gloo_order(flat, case.dtype, "sum"), ref, case.dtype, f"{case.name} (gloo order)")
        # and the kernels' rank order stays inside the stated tolerance of it
        err = np.abs(O.to_f64(rank_order, case.dtype) - O.to_f64(ref, case.dtype))
        assert (err <= O.float_sum_tolerance(flat, case.dtype)).all()
    if is_float and case.op == "prod" and world > 2:
        rel = np.abs(O.to_f64(rank_order, case.dtype) - O.to_f64(ref, case.dtype)) / (np.abs(O.to_f64(ref, case.dtype)) + 1e-30)
        assert rel.max() < 1e-6


def test_bf16_helpers_round_trip_and_round_to_nearest_even():
    x = np.array([1.0, 1.00390625, 1.01171875, -2.5, 3.0e38, np.inf, np.nan, 1e-40], dtype=np.float32)
    b = O.f32_to_bf16(x)
    back = O.bf16_to_f32(b)
    assert back[0] == 1.0 and back[3] == -2.5 and np.isinf(back[5]) and np.isnan(back[6])
    # 1 + 2^-8 is a tie between 1.0 and 1 + 2^-7: round to even (1.0); 1 + 3*2^-8 ties up to 1 + 2^-6
    assert back[1] == 1.0
    assert back[2] == np.float32(1.015625)
    import torch

    t = torch.from_numpy(x[:5].copy()).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(t, b[:5])


def test_oracle_rejects_what_the_reference_rejects():
    x = O.make_inputs(2, 8, "i32")
    with pytest.raises(TypeError):
        O.reduce_rank_order(x, "i32", "avg")  # gloo raises on integer AVG too
    with pytest.raises(TypeError):
        O.reduce_rank_order(O.make_inputs(2, 8, "f32"), "f32", "bxor")
    with pytest.raises(ValueError):
        O.broadcast(x, 2)
    with pytest.raises(ValueError):
        O.reduce_scatter(O.make_inputs(2, 7, "f32"), "f32")
