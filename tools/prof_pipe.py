#!/usr/bin/env python
"""One emulated-rank all_reduce through the tile-pipeline kernel (TMA rings + reducing warps),
for `ncu --set full -k regex:k_allreduce_pipe`: 4 ranks x 64 MiB fp32 on one GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch

    from nbdistributed_b200 import capi

    world, numel = 4, (64 << 20) // 4
    with capi.EmuGroup(world, device=0, stage_bytes=256 << 20) as grp:
        grp.set_tunable("pipe", 1)
        x = [torch.full((numel,), float(r + 1), device="cuda:0") for r in range(world)]
        y = [torch.empty_like(t) for t in x]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(4):
            if it == 1:
                e0.record()
            grp.all_reduce(x, y)
        e1.record()
        torch.cuda.synchronize()
        grp.async_error()
        assert grp.last_algo() == "ar_pipe" and float(y[0][0]) == 10.0 and float(y[3][-1]) == 10.0
        ms = e0.elapsed_time(e1) / 3
        print(f"ar_pipe emulated {world} x {numel * 4 >> 20} MiB: {ms:.3f} ms/launch, "
              f"{2 * world * numel * 4 / ms / 1e6:.0f} GB/s algorithmic (2*R*S)")


if __name__ == "__main__":
    main()
