#!/usr/bin/env python
"""Smallest run of the round-2 kernels for `compute-sanitizer --tool memcheck`: emulated ranks on one
GPU, tile pipeline (TMA rings), point-to-point ring, broadcast table, LL broadcast."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch

    from nbdistributed_b200 import capi

    world = 4
    with capi.EmuGroup(world, device=0, stage_bytes=8 << 20) as grp:
        grp.set_tunable("pipe", 1)
        x = [torch.full((300_003,), float(r + 1), device="cuda:0") for r in range(world)]
        grp.all_reduce(x, x)
        b = [torch.full((100_001,), r, device="cuda:0", dtype=torch.uint8) for r in range(world)]
        grp.broadcast(b, root=2)
        grp.set_tunable("pipe", 0)
        s = [torch.full((70_001,), r + 1, device="cuda:0", dtype=torch.uint8) for r in range(world)]
        d = [torch.zeros_like(t) for t in s]
        grp.p2p_batch([[("send", (r + 1) % world, s[r]), ("recv", (r - 1) % world, d[r])] for r in range(world)])
        items = [([torch.full((n,), r, device="cuda:0", dtype=torch.uint8) for r in range(world)], n % world) for n in (40, 3072, 12289)]
        grp.broadcast_table(items)
        small = [torch.full((1000,), r, device="cuda:0", dtype=torch.uint8) for r in range(world)]
        grp.broadcast(small, root=1)
        torch.cuda.synchronize()
        grp.async_error()
        assert float(x[0][0]) == 10.0 and int(b[0][0]) == 2 and int(d[0][0]) == world and int(small[3][5]) == 1
        assert all(int(per_rank[r][0]) == root for per_rank, root in items for r in range(world))
    print("sanitize_small ok")


if __name__ == "__main__":
    main()
