"""Broadcast tables (csrc/nbd_table.cuh): many small broadcasts in one launch with one flag
barrier - what a c10d coalescing block around the reference's per-parameter loop
(ref:README.md:122-123) turns into.  One GPU, emulated ranks; every check bit-exact.
Covered: mixed roots, misaligned and empty items, more items than one table holds, an item too
large for a table (falls back to the ordinary broadcast), back-to-back tables and collectives
without host synchronisation, the GPT-2 small state_dict of BASELINE config 3.
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
        # 300 items (> 128 per table), one of them larger than a staging half allows in a table
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
    """BASELINE config 3: the 149 tensors of GPT-2 small, root 0 - the 98 small ones (<= 12 KB)
    as tables, the large ones as ordinary broadcasts (what the c10d glue does)."""
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
