"""nbd_local_all_reduce_noalloc — the all-reduce for callers that may not allocate.

Some hosts of this library cannot call cudaMalloc when they need a collective: the GPU's
memory already belongs to another allocator, or the loop they sit in cannot absorb the
device-wide stall a driver allocation causes.  For them the library carries a fixed scratch
array and this one entry point that uses it instead of a communicator's arena.

The test calls it through the C ABI (ctypes) and compares, bit for bit (the bar every other
reduction test here uses), with nbd_emu_all_reduce — the existing entry point for the same
operation — on two calls in a row, since the second call runs on whatever the first one
left in the shared array.  This file sorts last on purpose: a run that stops at the first
failure has then already executed every other test.
"""
import pytest

from oracle import collectives_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _reference(torch, capi, x, dtype):
    """The same all-reduce through the existing, arena-backed entry point."""
    world = len(x)
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        t = [H.to_cuda(a, dtype) for a in x]
        out = [torch.empty_like(a) for a in t]
        grp.all_reduce(t, out)
        torch.cuda.synchronize()
        grp.async_error()
        return [H.to_numpy(o, dtype) for o in out]


@pytest.mark.parametrize("world,dtype,numel", [(4, "f32", 100003), (8, "bf16", 4099), (2, "i32", 1 << 18)])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(world, dtype, numel):
    import torch

    from nbdistributed_b200 import capi

    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    assert numel * O.NP_DTYPE[dtype]().itemsize <= capi.NOALLOC_MAX_BYTES
    # two different inputs, two consecutive calls on the same stream, no host sync between
    xs = [O.make_inputs(world, numel, dtype, seed=11), O.make_inputs(world, numel, dtype, seed=12)]
    sends = [[H.to_cuda(a, dtype) for a in x] for x in xs]
    outs = [[torch.empty_like(a) for a in s] for s in sends]
    capi.local_all_reduce_noalloc(sends[0], outs[0])
    capi.local_all_reduce_noalloc(sends[1], outs[1])
    torch.cuda.synchronize()
    for call, x in enumerate(xs):
        want = _reference(torch, capi, x, dtype)
        for r in range(world):
            H.assert_same(H.to_numpy(outs[call][r], dtype), want[r], dtype, f"call {call} rank {r}")
            H.assert_same(H.to_numpy(sends[call][r], dtype), x[r], dtype, f"call {call}: input of rank {r} modified")
    # and both agree with the oracle's fixed rank order
    H.assert_same(H.to_numpy(outs[1][0], dtype), O.reduce_rank_order(xs[1], dtype), dtype)


def test_noalloc_world_size_changes_between_calls():
    """The shared control block is reset per call, so 8 ranks, then 3, then 8 again works."""
    import torch

    from nbdistributed_b200 import capi

    torch.cuda.set_device(0)
    for world in (8, 3, 8):
        x = O.make_inputs(world, 20011, "f32", seed=world)
        t = [H.to_cuda(a, "f32") for a in x]
        capi.local_all_reduce_noalloc(t)  # in place
        torch.cuda.synchronize()
        want = O.reduce_rank_order(x, "f32")
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], "f32"), want, "f32", f"world {world} rank {r}")


def test_noalloc_rejects_oversize_with_a_clear_error():
    import torch

    from nbdistributed_b200 import capi

    torch.cuda.set_device(0)
    n = capi.NOALLOC_MAX_BYTES // 4 + 1
    t = [torch.zeros(n, device="cuda:0") for _ in range(2)]
    with pytest.raises(capi.NbdError) as ei:
        capi.local_all_reduce_noalloc(t)
    assert "NBD_NOALLOC_MAX_BYTES" in str(ei.value) and "exceeds" in str(ei.value)
    # exactly at the limit is accepted
    t = [torch.ones(capi.NOALLOC_MAX_BYTES // 4, device="cuda:0") for _ in range(2)]
    capi.local_all_reduce_noalloc(t)
    torch.cuda.synchronize()
    assert float(t[0][0]) == 2.0 and float(t[1][-1]) == 2.0
