#!/usr/bin/env python
"""1-GPU row: the emulated 8-rank 256 MiB all_reduce under the zero-copy kernel's variants
(fused loop / tiled read-then-push, tile sizes, grid sizes).  One JSON line per variant."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch

    from nbdistributed_b200 import capi

    R, S = 8, 256 << 20
    grp = capi.EmuGroup(R, device=0)
    x = [torch.randn(S // 4, device="cuda:0") for _ in range(R)]
    y = [torch.empty_like(t) for t in x]
    for mode, tile, blocks in ((0, 0, 0), (1, 0, 0), (2, 0, 0), (2, 1 << 20, 0), (1, 0, 18), (1, 0, 17), (1, 0, 16), (1, 0, 15), (1, 0, 14), (1, 0, 12), (1, 0, 10), (1, 0, 8), (2, 0, 16)):
        grp.set_tunable("zc_ar_mode", mode)
        grp.set_tunable("zc_tile_bytes", tile)
        grp.set_tunable("blocks", blocks)
        for _ in range(3):
            grp.all_reduce(x, y)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            grp.all_reduce(x, y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(json.dumps({"mode": mode, "tile": tile, "blocks": blocks, "ms": round(ms, 4), "algo": grp.last_algo(),
                          "GBps_2RS": round(2 * R * S / ms / 1e6, 1)}), flush=True)
    grp.close()


if __name__ == "__main__":
    main()
