#!/usr/bin/env python
"""Drive tools/p2p_probe.cu on 2 GPUs: which copy strategy sustains the most NVLink bandwidth?

    python tools/p2p_probe.py            # builds tools/_p2p_probe.so if needed, prints a table

One process, cuda:0 is the active GPU, cuda:1 the peer.  Bandwidth = bytes moved by GPU 0's
kernel / time (for mode "both": bytes in each direction)."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_p2p_probe.so")
SRC = os.path.join(HERE, "p2p_probe.cu")


def build():
    if os.path.exists(SO) and os.path.getmtime(SO) >= os.path.getmtime(SRC):
        return
    subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-shared",
                    "-Xcompiler", "-fPIC", "-o", SO, SRC], check=True)


def main():
    build()
    if "--build-only" in sys.argv:
        return
    import torch

    assert torch.cuda.device_count() >= 2, "needs 2 GPUs"
    lib = C.CDLL(SO)
    lib.probe_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                               C.c_void_p, C.c_void_p]
    lib.probe_tma_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p]
    assert lib.probe_enable_peer(0, 1) == 0 and lib.probe_enable_peer(1, 0) == 0
    torch.cuda.set_device(0)
    nbytes = 256 << 20
    loc_a = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device="cuda:0")
    loc_b = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
    rem_a = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device="cuda:1")
    rem_b = torch.empty(nbytes, dtype=torch.uint8, device="cuda:1")
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    stream = torch.cuda.current_stream().cuda_stream
    results = []

    def timed(fn, check, label, iters=5):
        rc = fn()
        torch.cuda.synchronize(0)
        if rc != 0:
            results.append({**label, "error": rc})
            print(json.dumps(results[-1]), flush=True)
            return
        ok = check()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(0)
        ms = e0.elapsed_time(e1) / iters
        results.append({**label, "us": round(ms * 1e3, 1), "GBps": round(nbytes / ms / 1e6, 1), "ok": bool(ok)})
        print(json.dumps(results[-1]), flush=True)

    def same(dst, src):
        torch.cuda.synchronize(1)
        return bool(torch.equal(dst.cpu()[:: 4099], src.cpu()[:: 4099]))

    for mode in ("pull", "push", "both"):
        for threads in (256, 512, 1024):
            for blocks in (148, 296, 592, 1184):
                for unroll in (1, 2, 4, 8):
                    if threads * blocks > 148 * 2048 * 4:
                        continue
                    if mode == "pull":
                        loc_b.zero_()
                        fn = lambda: lib.probe_copy(loc_b.data_ptr(), rem_a.data_ptr(), nbytes, blocks, threads, unroll, 0, None, None, stream)
                        chk = lambda: same(loc_b, rem_a)
                    elif mode == "push":
                        rem_b.zero_()
                        fn = lambda: lib.probe_copy(rem_b.data_ptr(), loc_a.data_ptr(), nbytes, blocks, threads, unroll, 0, None, None, stream)
                        chk = lambda: same(rem_b, loc_a)
                    else:
                        loc_b.zero_(); rem_b.zero_()
                        fn = lambda: lib.probe_copy(loc_b.data_ptr(), rem_a.data_ptr(), nbytes, blocks, threads, unroll, 1,
                                                    rem_b.data_ptr(), loc_a.data_ptr(), stream)
                        chk = lambda: same(loc_b, rem_a) and same(rem_b, loc_a)
                    timed(fn, chk, {"mode": mode, "threads": threads, "blocks": blocks, "unroll": unroll})
    for mode in ("tma_pull", "tma_push"):
        for blocks in (32, 74, 148, 296):
            for tile, stages in ((8192, 8), (16384, 8), (16384, 12), (32768, 6), (65536, 3)):
                if mode == "tma_pull":
                    loc_b.zero_()
                    fn = lambda: lib.probe_tma_copy(loc_b.data_ptr(), rem_a.data_ptr(), nbytes, blocks, tile, stages, stream)
                    chk = lambda: same(loc_b, rem_a)
                else:
                    rem_b.zero_()
                    fn = lambda: lib.probe_tma_copy(rem_b.data_ptr(), loc_a.data_ptr(), nbytes, blocks, tile, stages, stream)
                    chk = lambda: same(rem_b, loc_a)
                timed(fn, chk, {"mode": mode, "blocks": blocks, "tile": tile, "stages": stages})
    # reference points: cudaMemcpyPeer via torch
    for name, dst, src in (("memcpy_pull", loc_b, rem_a), ("memcpy_push", rem_b, loc_a)):
        timed(lambda: (dst.copy_(src, non_blocking=True), 0)[1], lambda: True, {"mode": name})


if __name__ == "__main__":
    main()
