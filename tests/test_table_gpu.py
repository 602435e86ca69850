"""Broadcast tables (csrc/nbd_table.cuh): many small broadcasts in one launch with one flag
barrier - what a c10d coalescing block around the reference's per-parameter loop
(ref:README.md:122-123) turns into.  One GPU, emulated ranks; every check bit-exact.
Covered: mixed roots, misaligned and empty items, more items than one table holds, an item too
large for a table (falls back to the ordinary broadcast), back-to-back tables and collectives
without host synchronisation, the GPT-2 small state_dict of BASELINE config 3.

Flat tables (k_bcast_flat): a table over 256 KiB whose items share a root moves as ONE chunked
scatter pipeline over the packed layout.  Covered: ragged and misaligned items of every size
class, every root, many chunks per launch, nothing written outside the items, the whole GPT-2
small state_dict (149 tensors, 652 MB) in one launch, flat tables between other collectives.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rand(torch, n, seed):
    g = torch.Generator(device="cuda:0")
    g.manual_seed(seed)
    return torch.randint(0, 256, (max(n, 1),), generator=g, device="cuda:0", dtype=torch.uint8)[:n]


def _make(torch, world, sizes, roots, misalign=False):
    items, want = [], []
    for k, (n, root) in enumerate(zip(sizes, roots)):
        per_rank = []
        for r in range(world):
            back = torch.zeros(n + 16, dtype=torch.uint8, device="cuda:0")
            off = (k + r) % 5 if misalign else 0
            v = back[off:off + n]
            v.copy_(_rand(torch, n, 1000 * k + r))
            per_rank.append(v)
        want.append(per_rank[root].clone())
        items.append((per_rank, root))
    return items, want


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("misalign", [False, True])
def test_table_mixed_roots_and_sizes(world, misalign):
    import torch

    from nbdistributed_b200 import capi

    rng = np.random.default_rng(world)
    sizes = [0, 1, 15, 16, 17, 400, 40, 3072, 9216, 12288, 70001, 262144] + [int(x) for x in rng.integers(1, 50000, 30)]
    roots = [int(x) for x in rng.integers(0, world, len(sizes))]
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        items, want = _make(torch, world, sizes, roots, misalign)
        n0 = capi.launch_count()
        grp.broadcast_table(items)
        assert capi.launch_count() - n0 == 1, "42 small broadcasts must be one launch"
        assert grp.last_algo() == "bcast_table"
        torch.cuda.synchronize()
        grp.async_error()
        for k, (per_rank, root) in enumerate(items):
            for r in range(world):
                assert torch.equal(per_rank[r], want[k]), f"item {k} ({sizes[k]} B, root {root}) rank {r}"


def test_table_splits_and_oversize_items():
    import torch

    from nbdistributed_b200 import capi

    world = 4
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        # 300 items (> 256 per table), one of them larger than a staging half allows in a table
        sizes = [3072] * 150 + [(5 << 20) + 3] + [12288] * 149
        roots = [k % world for k in range(len(sizes))]
        items, want = _make(torch, world, sizes, roots)
        x = [torch.full((1 << 18,), float(r + 1), device="cuda:0") for r in range(world)]
        grp.all_reduce(x, x)
        n0 = capi.launch_count()
        grp.broadcast_table(items)
        launches = capi.launch_count() - n0
        grp.all_reduce(x, x)
        torch.cuda.synchronize()
        grp.async_error()
        assert launches <= 6, launches
        for k, (per_rank, root) in enumerate(items):
            for r in range(world):
                assert torch.equal(per_rank[r], want[k]), f"item {k} rank {r}"
        assert float(x[0][0]) == 10.0 * world


def test_table_gpt2_state_dict_shapes():
    """BASELINE config 3: the 98 small tensors (<= 12 KB) of GPT-2 small, root 0, in one launch."""
    import torch

    from nbdistributed_b200 import capi
    from tools.configs_bench import gpt2_small_shapes

    world = 4
    small = [int(np.prod(sh)) * 4 for sh in gpt2_small_shapes() if int(np.prod(sh)) * 4 <= (256 << 10)]
    assert len(small) == 98
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        items, want = _make(torch, world, small, [0] * len(small))
        n0 = capi.launch_count()
        grp.broadcast_table(items)
        assert capi.launch_count() - n0 == 1
        torch.cuda.synchronize()
        grp.async_error()
        for k, (per_rank, _) in enumerate(items):
            for r in range(world):
                assert torch.equal(per_rank[r], want[k]), f"tensor {k} rank {r}"


def _make_guarded(torch, world, sizes, root, seed=0):
    """Items inside zeroed backing buffers at odd offsets: returns (items, want, backs, offs)."""
    items, want, backs, offs = [], [], [], []
    for k, n in enumerate(sizes):
        per_rank, bk = [], []
        off = (3 * k + 1) % 7 if k % 3 == 0 else 0  # every third item misaligned
        for r in range(world):
            back = torch.zeros(n + 48, dtype=torch.uint8, device="cuda:0")
            v = back[16 + off:16 + off + n]
            v.copy_(_rand(torch, n, seed + 1000 * k + r))
            per_rank.append(v)
            bk.append(back)
        want.append(per_rank[root].clone())
        items.append((per_rank, root))
        backs.append(bk)
        offs.append(16 + off)
    return items, want, backs, offs


def _check_guarded(torch, world, sizes, items, want, backs, offs, what):
    for k, (per_rank, _) in enumerate(items):
        for r in range(world):
            assert torch.equal(per_rank[r], want[k]), f"{what}: item {k} ({sizes[k]} B) rank {r}"
            b, o, n = backs[k][r], offs[k], sizes[k]
            assert int(b[:o].sum()) == 0 and int(b[o + n:].sum()) == 0, f"{what}: item {k} rank {r} wrote outside the item"


@pytest.mark.parametrize("world", [2, 3, 8])
def test_flat_table_every_root_ragged_items(world):
    import torch

    from nbdistributed_b200 import capi

    rng = np.random.default_rng(10 + world)
    sizes = [1, 15, 16, 17, 3072, 70001, 262144, (1 << 20) + 5, 2359296, 9437184 + 3] + [int(x) for x in rng.integers(1, 400000, 20)]
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        for root in range(world):
            items, want, backs, offs = _make_guarded(torch, world, sizes, root, seed=root)
            n0 = capi.launch_count()
            grp.broadcast_table(items)
            assert capi.launch_count() - n0 == 1
            assert grp.last_algo() == "bcast_flat", grp.last_algo()
            torch.cuda.synchronize()
            grp.async_error()
            _check_guarded(torch, world, sizes, items, want, backs, offs, f"root {root}")


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("chunk", [7 * 16 * 3, 96 << 10])
def test_flat_table_many_chunks(world, chunk):
    """A tiny pipeline chunk: tens to hundreds of chunks per launch, both staging halves reused
    many times, items straddling chunk and shard boundaries everywhere."""
    import torch

    from nbdistributed_b200 import capi

    rng = np.random.default_rng(chunk)
    n_items, hi = (12, 3000) if chunk < 4096 else (40, 200000)
    sizes = [int(x) for x in rng.integers(1, hi, n_items)] + [300000 if chunk >= 4096 else 270000]
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        grp.set_tunable("chunk_bytes", chunk)
        items, want, backs, offs = _make_guarded(torch, world, sizes, world - 1)
        grp.broadcast_table(items)
        assert grp.last_algo() == "bcast_flat"
        torch.cuda.synchronize()
        grp.async_error()
        _check_guarded(torch, world, sizes, items, want, backs, offs, f"chunk {chunk}")


def test_flat_table_between_other_collectives():
    """No host synchronisation between flat tables, direct tables, all_reduces and ordinary
    broadcasts on one stream: the staging halves and epochs must line up across kernel kinds."""
    import torch

    from nbdistributed_b200 import capi

    world = 4
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        x = [torch.full((1 << 18,), float(r + 1), device="cuda:0") for r in range(world)]
        rounds = []
        for it in range(6):
            sizes = [3072, 12288, (1 << 20) + it, 50000 + it, 7]
            flat = _make_guarded(torch, world, sizes, it % world, seed=50 * it)
            small, small_want = _make(torch, world, [40, 3072, 9216], [(it + k) % world for k in range(3)])
            grp.all_reduce(x, x)
            grp.broadcast_table(flat[0])
            grp.broadcast_table(small)
            big = [_rand(torch, (2 << 20) + 9, 900 + 10 * it + r).clone() for r in range(world)]
            big_want = big[it % world].clone()
            grp.broadcast(big, root=it % world)
            rounds.append((sizes, flat, small, small_want, big, big_want))
        torch.cuda.synchronize()
        grp.async_error()
        for it, (sizes, flat, small, small_want, big, big_want) in enumerate(rounds):
            _check_guarded(torch, world, sizes, *flat, f"round {it}")
            for k, (per_rank, _) in enumerate(small):
                for r in range(world):
                    assert torch.equal(per_rank[r], small_want[k]), f"round {it} small item {k} rank {r}"
            for r in range(world):
                assert torch.equal(big[r], big_want), f"round {it} broadcast rank {r}"
        assert float(x[0][0]) == 10.0 * world ** 5


def test_flat_table_more_items_than_one_launch_holds():
    import torch

    from nbdistributed_b200 import capi

    world = 3
    sizes = [2048 + k for k in range(600)]
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        items, want, backs, offs = _make_guarded(torch, world, sizes, 1)
        n0 = capi.launch_count()
        grp.broadcast_table(items)
        assert capi.launch_count() - n0 == 3  # 256 + 256 + 88 items
        torch.cuda.synchronize()
        grp.async_error()
        _check_guarded(torch, world, sizes, items, want, backs, offs, "600 items")


def test_flat_table_whole_gpt2_state_dict_one_launch():
    """BASELINE config 3 as the c10d glue now issues it inside a coalescing block: all 149 tensors
    of GPT-2 small (652 MB, 2.4 KB to 154 MB each), root 0, ONE launch."""
    import torch

    from nbdistributed_b200 import capi
    from tools.configs_bench import gpt2_small_shapes

    world = 4
    sizes = [int(np.prod(sh)) * 4 for sh in gpt2_small_shapes()]
    assert len(sizes) == 149
    with capi.EmuGroup(world, device=0, stage_bytes=64 << 20) as grp:
        items = []
        for k, n in enumerate(sizes):
            per_rank = [_rand(torch, n, 7 * k + r) for r in range(world)]
            items.append((per_rank, 0))
        want = [per_rank[0].clone() for per_rank, _ in items]
        n0 = capi.launch_count()
        grp.broadcast_table(items)
        assert capi.launch_count() - n0 == 1
        assert grp.last_algo() == "bcast_flat"
        torch.cuda.synchronize()
        grp.async_error()
        for k, (per_rank, _) in enumerate(items):
            for r in range(world):
                assert torch.equal(per_rank[r], want[k]), f"tensor {k} rank {r}"
