"""Point-to-point batches (csrc/nbd_p2p.cuh) on ONE GPU with emulated ranks: the primitive under
dist.send / dist.recv / batch_isend_irecv, all_to_all, gather and scatter.  Pure byte moves:
every check is bit-exact.  Covered: ring exchange, full all-to-all with a different size per pair,
messages longer than the 4-slot window of an ordered pair (slot reuse + acks), more messages per
lane than one launch carries, misaligned and empty buffers, back-to-back batches whose pair
counters must carry over between launches.  Real ranks: tests/mgpu_worker.py.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available()
    return torch


def _rand(torch, n, seed):
    g = torch.Generator(device="cuda:0")
    g.manual_seed(seed)
    return torch.randint(0, 256, (n,), generator=g, device="cuda:0", dtype=torch.uint8)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("nbytes", [1, 4099, (1 << 20) + 3, (6 << 20) + 17])
def test_ring_exchange(torch_cuda, world, nbytes):
    from nbdistributed_b200 import capi

    torch = torch_cuda
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        src = [_rand(torch, nbytes, 10 + r) for r in range(world)]
        dst = [torch.zeros(nbytes, dtype=torch.uint8, device="cuda:0") for _ in range(world)]
        # recv posted before send on even ranks, after on odd ranks: order must not matter
        ops = []
        for r in range(world):
            s, rc = ("send", (r + 1) % world, src[r]), ("recv", (r - 1) % world, dst[r])
            ops.append([rc, s] if r % 2 == 0 else [s, rc])
        for _ in range(3):  # pair counters carry over between launches
            grp.p2p_batch(ops)
        torch.cuda.synchronize()
        grp.async_error()
        assert grp.last_algo() == "p2p_batch"
        for r in range(world):
            assert torch.equal(dst[r], src[(r - 1) % world]), f"rank {r}"


@pytest.mark.parametrize("world", [2, 4, 8])
def test_all_to_all_with_a_size_per_pair(torch_cuda, world):
    from nbdistributed_b200 import capi

    torch = torch_cuda
    rng = np.random.default_rng(3)
    size = {(s, d): int(rng.integers(1, 3 << 20)) for s in range(world) for d in range(world) if s != d}
    size[(0, world - 1)] = 0  # an empty message is skipped on both sides
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        out = {k: _rand(torch, max(v, 1), 100 + 17 * k[0] + k[1])[:v] for k, v in size.items()}
        inn = {k: torch.zeros(v, dtype=torch.uint8, device="cuda:0") for k, v in size.items()}
        ops = [[("send", d, out[(r, d)]) for d in range(world) if d != r] +
               [("recv", s, inn[(s, r)]) for s in range(world) if s != r] for r in range(world)]
        grp.p2p_batch(ops)
        torch.cuda.synchronize()
        grp.async_error()
        for k in size:
            assert torch.equal(inn[k], out[k]), f"pair {k} ({size[k]} bytes)"


def test_many_messages_per_lane_and_misaligned_views(torch_cuda):
    from nbdistributed_b200 import capi

    torch = torch_cuda
    world = 3
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        sizes = [7, 70001, (2 << 20) + 1, 16, 5 << 20, 333, 1 << 20, 9, 123457, 64]  # 10 messages: 3 launches
        backing_s = [torch.zeros(n + 32, dtype=torch.uint8, device="cuda:0") for n in sizes]
        backing_d = [torch.zeros(n + 32, dtype=torch.uint8, device="cuda:0") for n in sizes]
        src = [b[1 + k % 5:1 + k % 5 + n] for k, (b, n) in enumerate(zip(backing_s, sizes))]
        dst = [b[3:3 + n] for b, n in zip(backing_d, sizes)]
        for k, s in enumerate(src):
            s.copy_(_rand(torch, sizes[k], 500 + k))
        back = torch.zeros(4099, dtype=torch.uint8, device="cuda:0")
        token = _rand(torch, 4099, 999)
        ops = [[("send", 2, s) for s in src] + [("recv", 2, back)],
               [],
               [("recv", 0, d) for d in dst] + [("send", 0, token)]]
        grp.p2p_batch(ops)
        torch.cuda.synchronize()
        grp.async_error()
        for k in range(len(sizes)):
            assert torch.equal(dst[k], src[k]), f"message {k}"
            assert int(backing_d[k][:3].sum()) == 0 and int(backing_d[k][3 + sizes[k]:].sum()) == 0, "wrote outside the view"
        assert torch.equal(back, token)


def test_p2p_interleaves_with_collectives(torch_cuda):
    """Point-to-point traffic uses its own area and counters: collectives before and after are
    unaffected, with no host synchronisation in between."""
    from nbdistributed_b200 import capi

    torch = torch_cuda
    world = 4
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        x = [torch.full((1 << 20,), float(r + 1), device="cuda:0") for r in range(world)]
        a = [_rand(torch, (3 << 20) + 5, 40 + r) for r in range(world)]
        b = [torch.zeros_like(t) for t in a]
        grp.all_reduce(x, x)
        grp.p2p_batch([[("send", (r + 2) % world, a[r]), ("recv", (r + 2) % world, b[r])] for r in range(world)])
        grp.all_reduce(x, x)
        torch.cuda.synchronize()
        grp.async_error()
        want = float(sum(range(1, world + 1))) * world
        for r in range(world):
            assert float(x[r][0]) == want and float(x[r][-1]) == want
            assert torch.equal(b[r], a[(r + 2) % world])


def test_invalid_arguments():
    from nbdistributed_b200 import capi
    import torch

    with capi.EmuGroup(2, device=0, stage_bytes=8 << 20) as grp:
        t = torch.zeros(8, dtype=torch.uint8, device="cuda:0")
        with pytest.raises(capi.NbdError):
            grp.p2p_batch([[("send", 0, t)], []])   # self
        with pytest.raises(capi.NbdError):
            grp.p2p_batch([[("send", 5, t)], []])   # no such rank
