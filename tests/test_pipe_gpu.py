"""The tile-pipeline kernels (csrc/nbd_pipe.cuh) on ONE GPU: emulated ranks, peer-to-peer
transport (tunable "pipe" = 1).  Same bar as tests/test_emu_gpu.py: bit-exact against the
oracle's ascending-rank restatement for every dtype, broadcast bit-exact.

What only this file exercises: the warp-specialised CTA (TMA copy-in ring, reducing warps, TMA
copy-out ring), per-tile A/B flags with slot reuse after kPipeDepth tiles, the opening and
closing flag barriers that hand the arena staging over to and from the chunked kernels, and the
byte-wise path for misaligned / ragged user buffers.  The NVSwitch-multicast transport of the
same kernels needs >= 2 GPUs and is covered by tests/mgpu_worker.py.
"""
import numpy as np
import pytest

from oracle import collectives_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu

OPCODE = {"sum": 0, "avg": 1, "prod": 2, "min": 3, "max": 4, "band": 5, "bor": 6, "bxor": 7}


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available()
    return torch


@pytest.fixture(scope="module")
def groups(torch_cuda):
    from nbdistributed_b200 import capi

    made = {}

    def get(world, stage_bytes=16 << 20):
        key = (world, stage_bytes)
        if key not in made:
            made[key] = capi.EmuGroup(world, device=0, stage_bytes=stage_bytes)
            made[key].set_tunable("pipe", 1)
        return made[key]

    yield get
    for g in made.values():
        g.close()


def _all_reduce(torch, grp, x, dtype, op="sum", inplace=True):
    sends = [H.to_cuda(a, dtype) for a in x]
    recvs = sends if inplace else [torch.empty_like(s) for s in sends]
    grp.all_reduce(sends, recvs, op=OPCODE[op])
    torch.cuda.synchronize()
    grp.async_error()
    return [H.to_numpy(r, dtype) for r in recvs], [H.to_numpy(s, dtype) for s in sends]


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype", ["f32", "bf16", "f16", "f64", "i32", "i64", "u8", "i8"])
@pytest.mark.parametrize("numel", [(32 << 10) + 16, (1 << 18) + 1, (3 << 20) + 5])
def test_pipe_all_reduce_matches_oracle(torch_cuda, groups, world, dtype, numel):
    x = O.make_inputs(world, numel, dtype)
    grp = groups(world)
    got, _ = _all_reduce(torch_cuda, grp, x, dtype)
    assert grp.last_algo() == "ar_pipe", grp.last_algo()
    got2, sends = _all_reduce(torch_cuda, grp, x, dtype, inplace=False)
    want = O.reduce_rank_order(x, dtype)
    for r in range(world):
        H.assert_same(got[r], want, dtype, f"in place, rank {r}")
        H.assert_same(got2[r], want, dtype, f"out of place, rank {r}")
        H.assert_same(sends[r], x[r], dtype, f"send buffer of rank {r} was modified")


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("dtype,op", [("f32", "max"), ("f32", "min"), ("bf16", "avg"), ("f32", "prod"), ("i32", "bxor"),
                                      ("i64", "band"), ("u8", "bor"), ("f64", "avg")])
def test_pipe_all_reduce_other_ops(torch_cuda, groups, world, dtype, op):
    x = O.make_inputs(world, (1 << 17) + 3, dtype)
    got, _ = _all_reduce(torch_cuda, groups(world), x, dtype, op)
    want = O.reduce_rank_order(x, dtype, op)
    for r in range(world):
        H.assert_same(got[r], want, dtype, f"{op} rank {r}")


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("tile", [128, 4096, 64 << 10])
def test_pipe_many_rounds_reuse_slots(torch_cuda, groups, world, tile):
    """Tiny tiles force many rounds per CTA: every slot is reused several times."""
    grp = groups(world)
    grp.set_tunable("pipe_tile_bytes", tile)
    grp.set_tunable("blocks", 3)
    try:
        numel = 40_000 + 13 if tile < 4096 else (1 << 20) + 77
        x = O.make_inputs(world, numel, "f32")
        got, _ = _all_reduce(torch_cuda, grp, x, "f32")
        want = O.reduce_rank_order(x, "f32")
        for r in range(world):
            H.assert_same(got[r], want, "f32", f"rank {r}")
    finally:
        grp.set_tunable("pipe_tile_bytes", 0)
        grp.set_tunable("blocks", 0)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("depth", [2, 3, 4])
@pytest.mark.parametrize("mode", [0, 2])
def test_pipe_shapes_slots_and_reducing_groups(torch_cuda, groups, world, depth, mode):
    """Every pipeline shape the planner can be tuned to: 2-4 slots per CTA, one or two groups
    of reducing warps (mode bit 1; the groups take alternate tiles and must still signal in tile
    order), many short rounds."""
    grp = groups(world)
    knobs = {"pipe_depth": depth, "pipe_mode": mode, "pipe_tile_bytes": 8192, "blocks": 4}
    for k, v in knobs.items():
        grp.set_tunable(k, v)
    try:
        for dtype, numel in (("bf16", (1 << 20) + 9), ("f32", 300_001)):
            x = O.make_inputs(world, numel, dtype, seed=7)
            got, _ = _all_reduce(torch_cuda, grp, x, dtype)
            want = O.reduce_rank_order(x, dtype)
            for r in range(world):
                H.assert_same(got[r], want, dtype, f"{dtype} rank {r}")
        x = O.make_inputs(world, (1 << 20) + 3, "u8", seed=8)
        t = [H.to_cuda(a, "u8") for a in x]
        grp.broadcast(t, root=world - 1)
        torch_cuda.cuda.synchronize()
        grp.async_error()
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], "u8"), x[world - 1], "u8", f"broadcast rank {r}")
    finally:
        for k in knobs:
            grp.set_tunable(k, 0)


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("offset", [1, 2, 4, 8])
def test_pipe_misaligned_views(torch_cuda, groups, world, offset):
    """storage_offset views: the user side goes byte-wise, staging stays unit-aligned."""
    torch = torch_cuda
    n = (1 << 16) + 3
    x = O.make_inputs(world, n, "u8")
    backing = [torch.zeros(n + 64, dtype=torch.uint8, device="cuda:0") for _ in range(world)]
    views = [b[offset:offset + n] for b in backing]
    for v, a in zip(views, x):
        v.copy_(H.to_cuda(a, "u8"))
    grp = groups(world)
    grp.all_reduce(views, views)
    torch.cuda.synchronize()
    grp.async_error()
    want = O.reduce_rank_order(x, "u8")
    for r in range(world):
        H.assert_same(H.to_numpy(views[r], "u8"), want, "u8", f"rank {r}")
        assert int(backing[r][:offset].sum()) == 0 and int(backing[r][offset + n:].sum()) == 0, "wrote outside the view"


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("nbytes", [(32 << 10) + 16, (1 << 20) + 17, (5 << 20) + 3])
def test_pipe_broadcast_every_root(torch_cuda, groups, world, nbytes):
    grp = groups(world)
    for root in range(world):
        x = O.make_inputs(world, nbytes, "u8", seed=31 + root)
        t = [H.to_cuda(a, "u8") for a in x]
        grp.broadcast(t, root=root)
        assert grp.last_algo() == "bcast_pipe", grp.last_algo()
        torch_cuda.cuda.synchronize()
        grp.async_error()
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], "u8"), x[root], "u8", f"root {root} rank {r}")


def test_pipe_interleaves_with_chunked_kernels(torch_cuda, groups):
    """40 back-to-back collectives alternating pipeline and chunked kernels on one stream, no
    host sync in between: the opening/closing barriers must keep the shared arena consistent."""
    torch = torch_cuda
    world = 4
    grp = groups(world)
    rng = np.random.default_rng(5)
    cases, outs = [], []
    for it in range(40):
        dtype = ("f32", "bf16", "i32")[it % 3]
        numel = int(rng.integers(40_000, 900_000))
        x = O.make_inputs(world, numel, dtype, seed=100 + it)
        t = [H.to_cuda(a, dtype) for a in x]
        grp.set_tunable("pipe", 1 if it % 2 == 0 else 2)
        if it % 5 == 4:
            grp.broadcast(t, root=it % world)
            cases.append(("bc", dtype, x, it % world))
        else:
            grp.all_reduce(t, t)
            cases.append(("ar", dtype, x, None))
        outs.append(t)
    grp.set_tunable("pipe", 1)
    torch.cuda.synchronize()
    grp.async_error()
    for (kind, dtype, x, root), t in zip(cases, outs):
        want = O.reduce_rank_order(x, dtype) if kind == "ar" else x[root]
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], dtype), want, dtype, f"{kind} {dtype} rank {r}")


def test_pipe_256mib_properties(torch_cuda):
    """BASELINE-size message (256 MiB per rank, 4 ranks) through size-independent properties:
    an fp32 sum of small integers is exact in any order, and every rank ends bit-identical."""
    from nbdistributed_b200 import capi

    torch = torch_cuda
    world, numel = 4, 64 << 20
    with capi.EmuGroup(world, device=0, stage_bytes=128 << 20) as grp:
        grp.set_tunable("pipe", 1)
        g = torch.Generator(device="cuda:0")
        ints = []
        for r in range(world):
            g.manual_seed(77 + r)
            ints.append(torch.randint(-1000, 1000, (numel,), generator=g, device="cuda:0", dtype=torch.int32))
        want = ints[0].clone()
        for r in range(1, world):
            want += ints[r]
        t = [a.to(torch.float32) for a in ints]
        del ints
        grp.all_reduce(t, t)
        torch.cuda.synchronize()
        grp.async_error()
        assert grp.last_algo() == "ar_pipe"
        wf = want.to(torch.float32)
        for r in range(world):
            assert torch.equal(t[r], wf), f"rank {r}"


# ---- relay broadcast (k_broadcast_relay): scatter + relay with per-tile flags --------------------

@pytest.fixture(scope="module")
def relay_groups(torch_cuda):
    from nbdistributed_b200 import capi

    made = {}

    def get(world):
        if world not in made:
            made[world] = capi.EmuGroup(world, device=0, stage_bytes=16 << 20)
            made[world].set_tunable("pipe", 3)
        return made[world]

    yield get
    for g in made.values():
        g.close()


def _relay_bcast(torch, grp, world, nbytes, root, seed, offset=0):
    x = O.make_inputs(world, nbytes, "u8", seed=seed)
    backing = [torch.zeros(nbytes + 64, dtype=torch.uint8, device="cuda:0") for _ in range(world)]
    views = [b[offset:offset + nbytes] for b in backing]
    for v, a in zip(views, x):
        v.copy_(H.to_cuda(a, "u8"))
    grp.broadcast(views, root=root)
    assert grp.last_algo() == "bcast_relay", grp.last_algo()
    return x, backing, views


def _relay_check(torch, world, nbytes, root, x, backing, views, offset, what):
    for r in range(world):
        H.assert_same(H.to_numpy(views[r], "u8"), x[root], "u8", f"{what}: root {root} rank {r}")
        assert int(backing[r][:offset].sum()) == 0 and int(backing[r][offset + nbytes:].sum()) == 0, f"{what}: rank {r} wrote outside the view"


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("nbytes", [(32 << 10) + 16, 100_001, (1 << 20) + 17, (5 << 20) + 3])
def test_relay_broadcast_every_root(torch_cuda, relay_groups, world, nbytes):
    grp = relay_groups(world)
    for root in range(world):
        x, backing, views = _relay_bcast(torch_cuda, grp, world, nbytes, root, 31 + root)
        torch_cuda.cuda.synchronize()
        grp.async_error()
        _relay_check(torch_cuda, world, nbytes, root, x, backing, views, 0, "aligned")


@pytest.mark.parametrize("world", [3, 8])
@pytest.mark.parametrize("offset", [1, 4, 8])
def test_relay_broadcast_misaligned_views(torch_cuda, relay_groups, world, offset):
    grp = relay_groups(world)
    nbytes = (1 << 18) + 5
    x, backing, views = _relay_bcast(torch_cuda, grp, world, nbytes, world - 1, 5, offset)
    torch_cuda.cuda.synchronize()
    grp.async_error()
    _relay_check(torch_cuda, world, nbytes, world - 1, x, backing, views, offset, f"offset {offset}")


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("depth", [2, 4])
@pytest.mark.parametrize("mode", [2, 4])
@pytest.mark.parametrize("tile", [128, 8192])
def test_relay_broadcast_shapes(torch_cuda, relay_groups, world, depth, mode, tile):
    """Many short rounds: every slot reused tens of times; one or two pushing groups (mode 2 /
    mode 4 = one group); sub-shards of a few units, some of them empty in the last tile."""
    grp = relay_groups(world)
    knobs = {"pipe_depth": depth, "pipe_mode": mode, "pipe_tile_bytes": tile, "blocks": 3}
    for k, v in knobs.items():
        grp.set_tunable(k, v)
    try:
        nbytes = 40_013 if tile < 4096 else (1 << 20) + 77
        for root in (0, world - 1):
            x, backing, views = _relay_bcast(torch_cuda, grp, world, nbytes, root, 70 + root)
            torch_cuda.cuda.synchronize()
            grp.async_error()
            _relay_check(torch_cuda, world, nbytes, root, x, backing, views, 0, f"depth {depth} mode {mode} tile {tile}")
    finally:
        for k in knobs:
            grp.set_tunable(k, 0)


def test_relay_broadcast_interleaves_with_other_kernels(torch_cuda, relay_groups):
    """Back-to-back relay broadcasts, chunked all_reduces and LL-size broadcasts on one stream,
    roots rotating, no host sync in between."""
    torch = torch_cuda
    world = 4
    grp = relay_groups(world)
    rng = np.random.default_rng(9)
    done = []
    for it in range(30):
        if it % 3 == 2:
            dtype, numel = "f32", int(rng.integers(40_000, 600_000))
            x = O.make_inputs(world, numel, dtype, seed=300 + it)
            t = [H.to_cuda(a, dtype) for a in x]
            grp.all_reduce(t, t)
            done.append(("ar", dtype, x, None, t))
        else:
            nbytes = int(rng.integers(1_000, 3_000_000))
            x = O.make_inputs(world, nbytes, "u8", seed=300 + it)
            t = [H.to_cuda(a, "u8") for a in x]
            grp.broadcast(t, root=it % world)
            done.append(("bc", "u8", x, it % world, t))
    torch.cuda.synchronize()
    grp.async_error()
    for kind, dtype, x, root, t in done:
        want = O.reduce_rank_order(x, dtype) if kind == "ar" else x[root]
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], dtype), want, dtype, f"{kind} rank {r}")


def test_relay_broadcast_256mib(torch_cuda):
    """BASELINE-size message (256 MiB, 4 ranks, default tile and depth): every rank ends with the
    root's bytes."""
    from nbdistributed_b200 import capi

    torch = torch_cuda
    world, nbytes = 4, 256 << 20
    with capi.EmuGroup(world, device=0, stage_bytes=128 << 20) as grp:
        grp.set_tunable("pipe", 3)
        g = torch.Generator(device="cuda:0")
        t = []
        for r in range(world):
            g.manual_seed(11 + r)
            t.append(torch.randint(0, 256, (nbytes,), generator=g, device="cuda:0", dtype=torch.uint8))
        want = t[2].clone()
        grp.broadcast(t, root=2)
        assert grp.last_algo() == "bcast_relay"
        torch.cuda.synchronize()
        grp.async_error()
        for r in range(world):
            assert torch.equal(t[r], want), f"rank {r}"
