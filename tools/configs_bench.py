#!/usr/bin/env python
"""BASELINE.json configs 3-5 on the b200 backend (run under torch.distributed.run).

  cfg3  rank-0 broadcast of a random-init GPT-2 small state_dict (149 tensors, 3 KB .. 154 MB)
  cfg4  DDP-style bf16 gradient all_reduce of a 1.236 G-parameter Llama-style model: the
        bucket list DDP would build (1 MiB first bucket, 25 MiB buckets, one 525 MB bucket for
        the tied embedding), issued asynchronously like the reducer does
  cfg5  ZeRO-style reduce_scatter + all_gather of a flat bf16 buffer of 2^32 elements (8 GiB)

Each is checked (bit-exact broadcast; exactly representable sums) and timed; NCCL on the same
GPUs is timed beside cfg4/cfg5 as a comparison point.  One JSON line per config on rank 0.
"""
from __future__ import annotations

import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def gpt2_small_shapes():
    """Parameter shapes of GPT2LMHeadModel(GPT2Config()) incl. the tied lm_head alias (149 entries)."""
    d, v, p, layers = 768, 50257, 1024, 12
    shapes = [(v, d), (p, d)]
    for _ in range(layers):
        shapes += [(d,), (d,), (d, 3 * d), (3 * d,), (d, d), (d,), (d,), (d,), (d, 4 * d), (4 * d,), (4 * d, d), (d,)]
    shapes += [(d,), (d,), (v, d)]
    return shapes


def main():
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    W.init_worker(rank, world, os.environ.get("MASTER_ADDR", "127.0.0.1"), os.environ.get("MASTER_PORT", "29500"),
                  gpu_id=local, backend="b200")
    dev = torch.device("cuda", local)
    cpu = dist.new_group(backend="gloo")
    nccl = dist.new_group(backend="nccl") if "--no-nccl" not in sys.argv else None
    be = B.get_backend()

    def emit(obj):
        if rank == 0:
            print(json.dumps(obj), flush=True)

    def timed(fn, iters, warm, group_for_barrier=cpu):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        dist.barrier(group=group_for_barrier)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=group_for_barrier)
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu)
        return float(t.item())

    # ---------------- cfg3: GPT-2 small state_dict broadcast ----------------
    shapes = gpt2_small_shapes()
    g = torch.Generator(device=dev)
    g.manual_seed(0)
    tensors = []
    for shp in shapes:
        if rank == 0:
            tensors.append(torch.randn(shp, generator=g, device=dev))
        else:
            tensors.append(torch.zeros(shp, device=dev))
    total = sum(t.numel() * 4 for t in tensors)

    def bcast_all():
        for t in tensors:
            dist.broadcast(t, src=0)

    bcast_all()
    torch.cuda.synchronize()
    chk = torch.stack([t.view(torch.int32).sum(dtype=torch.int64) for t in tensors]).cpu()
    allc = [torch.empty_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk, group=cpu)
    ok3 = all(torch.equal(a, allc[0]) for a in allc)
    ms3 = timed(bcast_all, 5, 1)
    small = [t for t in tensors if t.numel() * 4 <= 12 * 1024]
    ms_small = timed(lambda: [dist.broadcast(t, src=0) for t in small], 5, 1)
    # the same loop inside a c10d coalescing block: all 149 tensors share root 0 and become ONE
    # flat-table launch (csrc/nbd_table.cuh, k_bcast_flat)
    def bcast_coalesced():
        with dist._coalescing_manager(device=dev):
            for t in tensors:
                dist.broadcast(t, src=0)

    for t in tensors[1:]:
        if rank != 0:
            t.zero_()
    bcast_coalesced()
    torch.cuda.synchronize()
    chk2 = torch.stack([t.view(torch.int32).sum(dtype=torch.int64) for t in tensors]).cpu()
    ok3c = bool(torch.equal(chk2, allc[0]))
    ms3c = timed(bcast_coalesced, 5, 1)
    line3 = {"config": "cfg3 GPT-2 small state_dict broadcast", "ranks": world, "tensors": len(tensors), "bytes": total,
             "identical_on_all_ranks": ok3, "ms_total": round(ms3, 3), "effective_GBps": round(total / ms3 / 1e6, 1),
             "small_tensors": len(small), "us_per_small_broadcast": round(ms_small * 1e3 / len(small), 2),
             "coalesced_ms_total": round(ms3c, 3), "coalesced_identical": ok3c,
             "coalesced_effective_GBps": round(total / ms3c / 1e6, 1)}
    if nccl is not None:
        ms3n = timed(lambda: [dist.broadcast(t, src=0, group=nccl) for t in tensors], 5, 1)
        line3["nccl_ms_total"] = round(ms3n, 3)
    n0 = B.launch_count()
    bcast_coalesced()
    torch.cuda.synchronize()
    line3["coalesced_launches"] = B.launch_count() - n0
    line3["coalesced_algo"] = be.last_algo()
    # the same list as ONE call (no per-tensor dispatcher trip on the host)
    for t in tensors[1:]:
        if rank != 0:
            t.zero_()
    B.broadcast_tensors(tensors, src=0)
    torch.cuda.synchronize()
    chk3 = torch.stack([t.view(torch.int32).sum(dtype=torch.int64) for t in tensors]).cpu()
    line3["one_call_identical"] = bool(torch.equal(chk3, allc[0]))
    ms3o = timed(lambda: B.broadcast_tensors(tensors, src=0), 5, 1)
    line3["one_call_ms_total"] = round(ms3o, 3)
    line3["one_call_effective_GBps"] = round(total / ms3o / 1e6, 1)
    t0 = time.perf_counter()
    for _ in range(3):
        bcast_coalesced()
    host_us = (time.perf_counter() - t0) / 3 * 1e6
    torch.cuda.synchronize()
    line3["coalesced_host_issue_us"] = round(host_us, 1)
    if "--only-cfg3" in sys.argv:  # pipeline-chunk sweep of the flat table
        sweep = {}
        for mib in (16, 32, 128, 256):
            be.set_tunable("chunk_bytes", mib << 20)
            sweep[f"{mib}MiB"] = round(timed(lambda: B.broadcast_tensors(tensors, src=0), 5, 1), 3)
        be.set_tunable("chunk_bytes", 0)
        line3["one_call_ms_by_chunk"] = sweep
    emit(line3)
    del tensors
    if "--only-cfg3" in sys.argv:
        dist.barrier(group=cpu)
        W.shutdown_worker()
        return

    # ---------------- cfg4: DDP-style bf16 gradient buckets, 1.236 G params ----------------
    params = 1_235_814_400
    emb = 128256 * 2048
    sizes = [emb, (1 << 20) // 2]
    rest = params - sum(sizes)
    cap = (25 << 20) // 2
    while rest > 0:
        n = min(cap, rest)
        sizes.append(n)
        rest -= n
    buckets = [torch.full((n,), float(rank + 1), device=dev, dtype=torch.bfloat16) for n in sizes]

    def step(group=None):
        works = [dist.all_reduce(b, async_op=True, group=group) for b in reversed(buckets)]
        for w in works:
            w.wait()

    step()
    torch.cuda.synchronize()
    want = float(world * (world + 1) // 2)
    ok4 = all(bool((b[:: max(1, b.numel() // 1000)] == want).all()) for b in buckets)
    for b in buckets:
        b.fill_(0.001)
    ms4 = timed(step, 5, 2)
    nbytes4 = params * 2
    line4 = {"config": "cfg4 DDP bf16 gradient all_reduce, 1.236 G params", "ranks": world, "buckets": len(sizes),
             "bytes": nbytes4, "exact": ok4, "ms_per_step": round(ms4, 3),
             "busbw_per_rank": round(nbytes4 * 2 * (world - 1) / world / ms4 / 1e6, 1)}
    if nccl is not None:
        ms4n = timed(lambda: step(nccl), 5, 2)
        line4["nccl_ms_per_step"] = round(ms4n, 3)
        line4["nccl_busbw_per_rank"] = round(nbytes4 * 2 * (world - 1) / world / ms4n / 1e6, 1)
    emit(line4)
    del buckets
    torch.cuda.empty_cache()

    # ---------------- cfg5: ZeRO-style reduce_scatter + all_gather, 2^32 bf16 elements ------
    numel = 1 << 32
    if "--small" in sys.argv:
        numel = 1 << 28
    flat = torch.empty(numel, dtype=torch.bfloat16, device=dev)
    # exactly representable pattern: element i of rank r holds ((i mod 5) + r) -> sums are small integers
    pat = (torch.arange(1 << 20, device=dev) % 5 + rank).to(torch.bfloat16)
    flat.view(-1, 1 << 20)[:] = pat
    shard = torch.empty(numel // world, dtype=torch.bfloat16, device=dev)
    dist.reduce_scatter_tensor(shard, flat)
    torch.cuda.synchronize()
    base = (torch.arange(1 << 20, device=dev) % 5).to(torch.float32) * world + world * (world - 1) / 2
    exp = base.to(torch.bfloat16)
    ok_rs = bool((shard.view(-1, 1 << 20) == exp).all())
    dist.all_gather_into_tensor(flat, shard)
    torch.cuda.synchronize()
    ok_ag = bool((flat.view(-1, 1 << 20)[:: 97] == exp).all()) and bool((flat[-(1 << 20):] == exp).all())
    algo_rs = None
    ms_rs = timed(lambda: dist.reduce_scatter_tensor(shard, flat), 3, 2)
    algo_rs = be.last_algo()
    ms_ag = timed(lambda: dist.all_gather_into_tensor(flat, shard), 3, 2)
    algo_ag = be.last_algo()
    nb = numel * 2
    line5 = {"config": "cfg5 ZeRO reduce_scatter + all_gather, flat bf16", "ranks": world, "elements": numel, "bytes": nb,
             "reduce_scatter_exact": ok_rs, "all_gather_exact": ok_ag,
             "reduce_scatter_ms": round(ms_rs, 3), "reduce_scatter_busbw": round(nb * (world - 1) / world / ms_rs / 1e6, 1),
             "all_gather_ms": round(ms_ag, 3), "all_gather_busbw": round(nb * (world - 1) / world / ms_ag / 1e6, 1),
             "algo": [algo_rs, algo_ag]}
    if nccl is not None:
        ms_rsn = timed(lambda: dist.reduce_scatter_tensor(shard, flat, group=nccl), 3, 1)
        ms_agn = timed(lambda: dist.all_gather_into_tensor(flat, shard, group=nccl), 3, 1)
        line5["nccl_reduce_scatter_busbw"] = round(nb * (world - 1) / world / ms_rsn / 1e6, 1)
        line5["nccl_all_gather_busbw"] = round(nb * (world - 1) / world / ms_agn / 1e6, 1)
    emit(line5)
    del flat, shard
    dist.barrier(group=cpu)
    W.shutdown_worker()


if __name__ == "__main__":
    main()
