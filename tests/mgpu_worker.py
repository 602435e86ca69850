"""Per-rank body of the multi-GPU backend tests (spawned by test_backend_multigpu.py, one
process per GPU).  Every check compares the b200 backend (CUDA tensors) with the reference's
CPU path — gloo, reached through the same process group because the backend string is
"cpu:gloo,cuda:b200" — and with the numpy oracle, on identical seeded inputs."""
from __future__ import annotations

import os
import sys
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import collectives_oracle as O  # noqa: E402
from tests import helpers as H  # noqa: E402


def _gather_cpu(dist, torch, x_cpu):
    """All ranks' copies of a CPU tensor, through gloo."""
    out = [torch.empty_like(x_cpu) for _ in range(dist.get_world_size())]
    dist.all_gather(out, x_cpu)
    return out


def _np(t, dtype):
    return H.to_numpy(t, dtype)


def body(rank: int, world: int, port: int) -> None:
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    ns = W.init_worker(rank, world, "127.0.0.1", port, gpu_id=rank, backend="b200")
    assert ns["__backend__"] == "cpu:gloo,cuda:b200"
    assert ns["device"] == rank and ns["gpu_id"] == rank
    dev = torch.device("cuda", rank)
    be = B.get_backend()
    assert type(be).__name__ == "BackendB200", type(be)

    # ---- config 1: README example (ref:README.md:99-125) through the cell runner -------
    W.run_cell(ns, "torch.manual_seed(1234 + rank)\nx = torch.randn(100, 100)\nxc = x.clone()\nxg = x.cuda()\n"
                   "dist.all_reduce(xg)\ndist.all_reduce(xc)\n")
    xg, xc = ns["xg"], ns["xc"]
    allx = _gather_cpu(dist, torch, ns["x"])
    want = O.reduce_rank_order([_np(a, "f32") for a in allx], "f32")
    H.assert_same(_np(xg, "f32"), want, "f32", "README all_reduce vs oracle")
    if world == 2:
        assert torch.equal(xg.cpu(), xc), "README all_reduce: not bit-identical with gloo at N=2"
    else:
        tol = O.float_sum_tolerance([_np(a, "f32") for a in allx], "f32")
        assert (np.abs(_np(xg, "f32").astype(np.float64) - _np(xc, "f32")) <= tol).all()
    W.run_cell(ns, "torch.manual_seed(rank)\nmodel = torch.nn.Linear(10, 10).cuda()\n"
                   "ref = [p.detach().cpu().clone() for p in model.parameters()]\n"
                   "for param in model.parameters():\n    dist.broadcast(param.data, src=0)\n"
                   "for p in ref:\n    dist.broadcast(p, src=0)\n")
    for p, q in zip(ns["model"].parameters(), ns["ref"]):
        assert torch.equal(p.detach().cpu(), q), "broadcast of Linear params differs from gloo"

    # ---- all_reduce sweep: dtypes x sizes, bit-exact vs oracle (rank order) --------------
    for dtype in ("f32", "bf16", "f16", "f64", "i32", "i64", "u8", "i8"):
        for numel in (1, 127, 4096 + 5, (1 << 20) + 1, 3 << 20):
            x = O.make_inputs(world, numel, dtype)
            t = H.to_cuda(x[rank], dtype, dev)
            dist.all_reduce(t)
            H.assert_same(_np(t, dtype), O.reduce_rank_order(x, dtype), dtype,
                          f"all_reduce {dtype} n={numel} algo={be.last_algo()}")
    # integer all_reduce is bit-exact with gloo
    x = O.make_inputs(world, 70001, "i64")
    t = H.to_cuda(x[rank], "i64", dev)
    c = torch.from_numpy(x[rank].copy())
    dist.all_reduce(t)
    dist.all_reduce(c)
    assert torch.equal(t.cpu(), c)
    # other ops
    for dtype, op, dop in (("f32", "max", dist.ReduceOp.MAX), ("f32", "min", dist.ReduceOp.MIN),
                           ("i32", "bxor", dist.ReduceOp.BXOR), ("bf16", "avg", dist.ReduceOp.AVG),
                           ("f32", "prod", dist.ReduceOp.PRODUCT)):
        x = O.make_inputs(world, 33333, dtype)
        t = H.to_cuda(x[rank], dtype, dev)
        dist.all_reduce(t, op=dop)
        H.assert_same(_np(t, dtype), O.reduce_rank_order(x, dtype, op), dtype, f"all_reduce {op}")

    # ---- NVSwitch multicast (csrc/nbd_nvls.cu, nbd_pipe.cuh) ---------------------------------
    # set-up is all-or-nothing across ranks; broadcast over it is a plain copy (bit-exact);
    # in-switch float sums are opt-in, agree with the rank-order oracle within the north_star
    # tolerance, are identical on every rank, and exact where the sum is exact in any order
    nv_on, nv_why = be.nvls_status()
    flags = _gather_cpu(dist, torch, torch.tensor([int(nv_on)]))
    assert all(int(f) == int(flags[0]) for f in flags), f"ranks disagree on multicast: {[int(f) for f in flags]}"
    if rank == 0:
        print(f"[mgpu_worker] multicast: {nv_on} ({nv_why})", flush=True)
    if os.environ.get("NBD_TEST_REQUIRE_NVLS") == "1":
        assert nv_on, f"multicast expected on this box: {nv_why}"
    if nv_on:
        be.set_tunable("nvls_min_bytes", 1)
        be.set_tunable("nvls_bcast", 1)  # off by default: the scatter kernel is faster
        for nbytes in (40, (48 << 10) + 5, (1 << 20) + 17, (9 << 20) + 1):
            for root in sorted({0, world - 1}):
                x = O.make_inputs(world, nbytes, "u8", seed=70 + root)
                t = H.to_cuda(x[rank], "u8", dev)
                dist.broadcast(t, src=root)
                assert be.last_algo() == "bcast_nvls", be.last_algo()
                H.assert_same(_np(t, "u8"), x[root], "u8", f"multicast broadcast {nbytes}B root {root}")
        v = torch.zeros(70001 + 8, dtype=torch.uint8, device=dev)[3:3 + 70001]  # misaligned view
        x = O.make_inputs(world, 70001, "u8", seed=75)
        v.copy_(H.to_cuda(x[rank], "u8", dev))
        dist.broadcast(v, src=1 % world)
        H.assert_same(_np(v, "u8"), x[1 % world], "u8", "multicast broadcast into a misaligned view")
        be.set_tunable("nvls_bcast", 0)
        be.set_tunable("nvls_reduce", 2)  # 2: also between two ranks, where 1 leaves it to the peer-to-peer kernels
        for mode in (0, 2, 6):  # one / two reducing groups, 4 / 8 loads in flight
            be.set_tunable("pipe_mode", mode)
            xi = [np.random.default_rng(60 + r).integers(-1000, 1000, (2 << 20) + 11).astype(np.float32) for r in range(world)]
            t = H.to_cuda(xi[rank], "f32", dev)
            dist.all_reduce(t)
            assert be.last_algo() == "ar_nvls"
            H.assert_same(_np(t, "f32"), O.reduce_rank_order(xi, "f32"), "f32", f"in-switch sum, pipe_mode {mode}")
        be.set_tunable("pipe_mode", 0)
        for dtype in ("f32", "bf16", "f16"):
            for numel in (4099, (1 << 20) + 1, (5 << 20) + 3):
                x = O.make_inputs(world, numel, dtype, seed=80)
                t = H.to_cuda(x[rank], dtype, dev)
                dist.all_reduce(t)
                assert be.last_algo() == "ar_nvls", be.last_algo()
                got = _np(t, dtype)
                err = np.abs(O.to_f64(got, dtype) - O.to_f64(O.reduce_rank_order(x, dtype), dtype))
                tol = O.float_sum_tolerance(x, dtype)
                assert (err <= tol).all(), f"in-switch {dtype} sum n={numel}: max err/tol {np.max(err / tol)}"
                chk = torch.from_numpy(got.view(np.uint8).astype(np.int64)).sum().reshape(1)
                allchk = _gather_cpu(dist, torch, chk)
                assert all(int(a) == int(allchk[0]) for a in allchk), f"in-switch {dtype} sum differs between ranks"
        xi = [np.random.default_rng(90 + r).integers(-1000, 1000, (3 << 20) + 11).astype(np.float32) for r in range(world)]
        t = H.to_cuda(xi[rank], "f32", dev)
        dist.all_reduce(t)
        H.assert_same(_np(t, "f32"), O.reduce_rank_order(xi, "f32"), "f32", "in-switch sum of integer-valued fp32 (exact in any order)")
        x = O.make_inputs(world, (1 << 20) + 3, "f32", seed=81)
        t = H.to_cuda(x[rank], "f32", dev)
        dist.all_reduce(t, op=dist.ReduceOp.AVG)
        assert be.last_algo() == "ar_nvls"
        err = np.abs(_np(t, "f32").astype(np.float64) - O.to_f64(O.reduce_rank_order(x, "f32", "avg"), "f32"))
        assert (err <= O.float_sum_tolerance(x, "f32") / world + 1e-12).all(), "in-switch AVG"
        # what the switch cannot add keeps the rank-order kernels: bit-exact as ever
        for dtype, dop, op in (("i32", dist.ReduceOp.SUM, "sum"), ("f32", dist.ReduceOp.MAX, "max"), ("f64", dist.ReduceOp.SUM, "sum")):
            x = O.make_inputs(world, (1 << 19) + 7, dtype, seed=82)
            t = H.to_cuda(x[rank], dtype, dev)
            dist.all_reduce(t, op=dop)
            assert be.last_algo() != "ar_nvls"
            H.assert_same(_np(t, dtype), O.reduce_rank_order(x, dtype, op), dtype, f"{dtype} {op} with nvls_reduce on")
        # a misaligned in-place view through the in-switch path
        v = torch.zeros((1 << 18) + 9, dtype=torch.float32, device=dev)[1:1 + (1 << 18) + 5]
        x = [np.random.default_rng(95 + r).integers(-99, 99, (1 << 18) + 5).astype(np.float32) for r in range(world)]
        v.copy_(H.to_cuda(x[rank], "f32", dev))
        dist.all_reduce(v)
        H.assert_same(_np(v, "f32"), O.reduce_rank_order(x, "f32"), "f32", "in-switch sum on a storage_offset view")
        # symmetric memory from the multicast pool (allocated while in-switch reduction is on):
        # all_reduce needs no staging at all
        pool = B.empty_symmetric((3 << 20) + 64, torch.float32)
        for numel, off in (((3 << 20) + 64, 0), (1 << 20, 4096)):
            xi = [np.random.default_rng(120 + r).integers(-1000, 1000, numel).astype(np.float32) for r in range(world)]
            v = pool[off:off + numel]
            v.copy_(H.to_cuda(xi[rank], "f32", dev))
            dist.all_reduce(v)
            assert be.last_algo() == "ar_nvls_zerocopy", be.last_algo()
            H.assert_same(_np(v, "f32"), O.reduce_rank_order(xi, "f32"), "f32", f"zero-copy in-switch sum n={numel} off={off}")
        x = O.make_inputs(world, 1 << 20, "bf16", seed=121)
        pb = pool.view(torch.bfloat16)[: 1 << 20]
        pb.copy_(H.to_cuda(x[rank], "bf16", dev))
        dist.all_reduce(pb)
        assert be.last_algo() == "ar_nvls_zerocopy"
        err = np.abs(O.to_f64(_np(pb, "bf16"), "bf16") - O.to_f64(O.reduce_rank_order(x, "bf16"), "bf16"))
        assert (err <= O.float_sum_tolerance(x, "bf16")).all(), "zero-copy in-switch bf16 sum"
        # what the switch is not asked to do falls back to the staged kernels on the same memory
        vi = pool.view(torch.int32)[: 70001]
        xi = O.make_inputs(world, 70001, "i32", seed=122)
        vi.copy_(H.to_cuda(xi[rank], "i32", dev))
        dist.all_reduce(vi)
        H.assert_same(_np(vi, "i32"), O.reduce_rank_order(xi, "i32"), "i32", "int32 on multicast-pool memory")
        ragged = pool[7:7 + 100003]  # not a whole number of 16-byte units: staged in-switch pipeline
        xr = [np.random.default_rng(130 + r).integers(-50, 50, 100003).astype(np.float32) for r in range(world)]
        ragged.copy_(H.to_cuda(xr[rank], "f32", dev))
        dist.all_reduce(ragged)
        H.assert_same(_np(ragged, "f32"), O.reduce_rank_order(xr, "f32"), "f32", "ragged view of multicast-pool memory")
        del pool, pb, vi, ragged, v
        be.set_tunable("nvls_reduce", 0)
        be.set_tunable("nvls_min_bytes", 0)
    # ---- the tile pipeline over peer-to-peer transport on real ranks: bit-exact ----------------
    be.set_tunable("pipe", 1)
    for dtype, numel in (("f32", (1 << 20) + 1), ("bf16", (3 << 20) + 5), ("i64", 70001), ("u8", (1 << 20) + 3)):
        x = O.make_inputs(world, numel, dtype, seed=85)
        t = H.to_cuda(x[rank], dtype, dev)
        dist.all_reduce(t)
        assert be.last_algo() == "ar_pipe", be.last_algo()
        H.assert_same(_np(t, dtype), O.reduce_rank_order(x, dtype), dtype, f"pipe all_reduce {dtype} n={numel}")
    be.set_tunable("nvls_off", 1)
    for nbytes, root in (((48 << 10) + 5, 0), ((6 << 20) + 1, world - 1)):
        x = O.make_inputs(world, nbytes, "u8", seed=86)
        t = H.to_cuda(x[rank], "u8", dev)
        dist.broadcast(t, src=root)
        assert be.last_algo() == "bcast_pipe", be.last_algo()
        H.assert_same(_np(t, "u8"), x[root], "u8", f"pipe broadcast {nbytes}B root {root}")
    be.set_tunable("nvls_off", 0)
    be.set_tunable("pipe", 0)

    # ---- broadcast: every root, latency and bandwidth paths, bit-exact -------------------
    for nbytes in (40, 400, 12 * 1024 + 3, (1 << 20) + 17, (40 << 20) + 1):
        for root in sorted({0, world - 1}):
            x = O.make_inputs(world, nbytes, "u8", seed=7 + root)
            t = H.to_cuda(x[rank], "u8", dev)
            dist.broadcast(t, src=root)
            H.assert_same(_np(t, "u8"), x[root], "u8", f"broadcast {nbytes}B root {root} algo={be.last_algo()}")

    # ---- all_gather / reduce_scatter (tensor and list forms) -----------------------------
    for dtype, numel in (("f32", 25), ("bf16", 4099), ("i64", 1 << 16), ("bf16", (2 << 20) + 8)):
        x = O.make_inputs(world, numel, dtype)
        t = H.to_cuda(x[rank], dtype, dev)
        out = torch.empty(world * numel, dtype=t.dtype, device=dev)
        dist.all_gather_into_tensor(out, t)
        H.assert_same(_np(out, dtype), O.all_gather(x)[0], dtype, f"all_gather_into_tensor {dtype} {numel}")
        lst = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(lst, t)
        for r in range(world):
            H.assert_same(_np(lst[r], dtype), x[r], dtype, f"all_gather list slot {r}")
        xs = O.make_inputs(world, world * numel, dtype, seed=99)
        s = H.to_cuda(xs[rank], dtype, dev)
        o = torch.empty(numel, dtype=s.dtype, device=dev)
        dist.reduce_scatter_tensor(o, s)
        H.assert_same(_np(o, dtype), O.reduce_scatter(xs, dtype)[rank], dtype, f"reduce_scatter_tensor {dtype} {numel}")
        o2 = torch.empty(numel, dtype=s.dtype, device=dev)
        parts = [c.clone() for c in s.chunk(world)]  # separate allocations, as a caller's list would be
        torch.cuda.synchronize()
        n0 = B.launch_count()
        dist.reduce_scatter(o2, parts)
        torch.cuda.synchronize()
        assert B.launch_count() - n0 == 1 and be.last_algo() == "reduce_scatter_list", (B.launch_count() - n0, be.last_algo())
        assert torch.equal(o, o2), "reduce_scatter list form differs from tensor form"

    # ---- symmetric memory (Backend.allocate_tensor): zero-copy kernels over NVLink ----------
    assert be.supports_tensor_alloc(dev)
    be.set_tunable("zc_min_bytes", 256 << 10)  # default is 4 MiB; exercise the zero-copy kernels on smaller messages too
    n_sym = (6 << 20) + 64
    sym = B.empty_symmetric(n_sym, torch.float32)
    sym2 = B.empty_symmetric(n_sym, torch.bfloat16)
    assert sym.is_cuda and sym.numel() == n_sym and sym2.dtype == torch.bfloat16
    for numel, off in ((n_sym, 0), ((1 << 20) + 3, 16), (70001, 4 * 999)):
        x = O.make_inputs(world, numel, "f32", seed=21)
        v = sym[off:off + numel]
        v.copy_(H.to_cuda(x[rank], "f32", dev))
        dist.all_reduce(v)
        expect_zc = numel * 4 >= (256 << 10)
        assert (be.last_algo() == "ar_zerocopy") == expect_zc, be.last_algo()
        H.assert_same(_np(v, "f32"), O.reduce_rank_order(x, "f32"), "f32", f"zero-copy all_reduce n={numel} off={off}")
    x = O.make_inputs(world, 1 << 20, "bf16", seed=22)
    v = sym2[: 1 << 20]
    v.copy_(H.to_cuda(x[rank], "bf16", dev))
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
    assert be.last_algo() == "ar_zerocopy"
    H.assert_same(_np(v, "bf16"), O.reduce_rank_order(x, "bf16", "max"), "bf16", "zero-copy all_reduce max bf16")
    shard = (1 << 19) + 16
    xs = O.make_inputs(world, world * shard, "f32", seed=23)
    sym[: world * shard].copy_(H.to_cuda(xs[rank], "f32", dev))
    o = torch.empty(shard, dtype=torch.float32, device=dev)
    dist.reduce_scatter_tensor(o, sym[: world * shard])          # send registered
    assert be.last_algo() == "reduce_scatter_zerocopy", be.last_algo()
    H.assert_same(_np(o, "f32"), O.reduce_scatter(xs, "f32")[rank], "f32", "zero-copy reduce_scatter")
    full = torch.empty(world * shard, dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(full, sym[:shard])                # send registered -> pull
    assert be.last_algo() == "all_gather_zerocopy_pull", be.last_algo()
    want_ag = np.concatenate([a[:shard] for a in xs])
    H.assert_same(_np(full, "f32"), want_ag, "f32", "zero-copy all_gather (pull)")
    src = H.to_cuda(xs[rank][:shard], "f32", dev)
    dist.all_gather_into_tensor(sym[: world * shard], src)        # recv registered -> push
    assert be.last_algo() == "all_gather_zerocopy_push", be.last_algo()
    H.assert_same(_np(sym[: world * shard], "f32"), want_ag, "f32", "zero-copy all_gather (push)")
    # staged and zero-copy collectives interleave freely on one communicator
    y = H.to_cuda(xs[rank][:shard], "f32", dev)
    dist.all_reduce(y)
    dist.all_reduce(sym[:shard])
    dist.all_reduce(y)
    torch.cuda.synchronize()
    be.set_tunable("zc_min_bytes", 0)
    del sym, sym2

    # ---- auto-registration: ordinary tensors go zero-copy once every rank has imported them ---
    be.set_tunable("auto_min_bytes", 1 << 20)  # default is 64 MiB
    numel = 5 << 20  # 20 MiB fp32
    t = torch.empty(numel + 64, dtype=torch.float32, device=dev)[32:32 + numel]  # offset view of an allocation
    decisions = []
    for it in range(4):
        x = O.make_inputs(world, numel, "f32", seed=40 + it)
        t.copy_(H.to_cuda(x[rank], "f32", dev))
        dist.all_reduce(t)
        assert be.last_algo() == "ar_auto", be.last_algo()
        decisions.append(be.last_zc_decision())
        H.assert_same(_np(t, "f32"), O.reduce_rank_order(x, "f32"), "f32", f"auto all_reduce iteration {it} (zc={decisions[-1]})")
    # the first use of an allocation is staged (peers have not imported it yet) unless an earlier
    # collective already touched the same caching-allocator segment; afterwards it is zero-copy
    assert decisions[-1] == 1 and sorted(decisions) == decisions, decisions
    allowed = _gather_cpu(dist, torch, torch.tensor(decisions))
    assert all(torch.equal(a, allowed[0]) for a in allowed), "ranks took different paths"
    shard = 3 << 20
    xs = O.make_inputs(world, world * shard, "bf16", seed=50)
    big_in = H.to_cuda(xs[rank], "bf16", dev)
    o = torch.empty(shard, dtype=torch.bfloat16, device=dev)
    full = torch.empty(world * shard, dtype=torch.bfloat16, device=dev)
    rs_dec, ag_dec = [], []
    for it in range(3):
        dist.reduce_scatter_tensor(o, big_in)
        rs_dec.append(be.last_zc_decision())
        H.assert_same(_np(o, "bf16"), O.reduce_scatter(xs, "bf16")[rank], "bf16", f"auto reduce_scatter iteration {it}")
        dist.all_gather_into_tensor(full, o)
        ag_dec.append(be.last_zc_decision())
        H.assert_same(_np(full, "bf16"), np.concatenate(O.reduce_scatter(xs, "bf16")), "bf16", f"auto all_gather iteration {it}")
    assert rs_dec[-1] == 1 and ag_dec[-1] == 1, (rs_dec, ag_dec)
    be.set_tunable("auto_register_off", 1)
    dist.all_reduce(t)
    assert be.last_algo() in ("ar_oneshot", "ar_twoshot")
    be.set_tunable("auto_register_off", 0)
    be.set_tunable("auto_min_bytes", 0)
    del t, big_in, o, full

    # ---- host buffers: CPU tensors all-reduced through the GPUs, vs gloo on the same tensors ----
    be.set_tunable("host_chunk_bytes", 4 << 20)
    xh = O.make_inputs(world, (5 << 20) + 3, "i32", seed=60)
    h_in = torch.from_numpy(xh[rank].copy()).pin_memory()
    h_out = torch.zeros_like(h_in).pin_memory()
    B.all_reduce_host(h_in, out=h_out)
    ref = h_in.clone()
    dist.all_reduce(ref)  # CPU tensor -> gloo
    assert torch.equal(h_out, ref), "all_reduce_host differs from gloo (int32, bit-exact)"
    hf = torch.from_numpy(O.make_inputs(world, 300001, "f32", seed=61)[rank].copy()).pin_memory()
    B.all_reduce_host(hf)
    H.assert_same(hf.numpy(), O.reduce_rank_order(O.make_inputs(world, 300001, "f32", seed=61), "f32"), "f32", "all_reduce_host f32")
    be.set_tunable("host_chunk_bytes", 0)

    # ---- point to point and what is built on it, against gloo on the same data ----------------
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    for numel in (1, 4099, (3 << 20) + 5):
        a = H.to_cuda(O.make_inputs(world, numel, "f32", seed=200)[rank], "f32", dev)
        b = torch.empty_like(a)
        if rank % 2 == 0:  # blocking calls: pair them up so that they cannot deadlock
            dist.send(a, nxt)
            dist.recv(b, prv)
        else:
            dist.recv(b, prv)
            dist.send(a, nxt)
        H.assert_same(_np(b, "f32"), O.make_inputs(world, numel, "f32", seed=200)[prv], "f32", f"send/recv ring n={numel}")
        # the same exchange as one batch: order inside the batch does not matter
        b2 = torch.zeros_like(a)
        reqs = dist.batch_isend_irecv([dist.P2POp(dist.isend, a, nxt), dist.P2POp(dist.irecv, b2, prv)])
        for q in reqs:
            q.wait()
        assert torch.equal(b2, b), "batch_isend_irecv differs from send/recv"
    x = O.make_inputs(world, world * 70001, "i32", seed=201)
    t_in = H.to_cuda(x[rank], "i32", dev)
    t_out = torch.empty_like(t_in)
    dist.all_to_all_single(t_out, t_in)
    want = np.concatenate([x[s][rank * 70001:(rank + 1) * 70001] for s in range(world)])
    H.assert_same(_np(t_out, "i32"), want, "i32", "all_to_all_single")
    splits_in = [(rank + p) % 3 + 1 for p in range(world)]              # what I send to p
    splits_out = [(s + rank) % 3 + 1 for s in range(world)]             # what s sends to me
    ti = torch.arange(sum(splits_in) * 5, device=dev, dtype=torch.float32).reshape(-1, 5) + 1000 * rank
    to = torch.empty(sum(splits_out), 5, device=dev)
    dist.all_to_all_single(to, ti, output_split_sizes=splits_out, input_split_sizes=splits_in)
    ci, co = ti.cpu(), torch.empty(sum(splits_out), 5)
    dist.all_to_all_single(co, ci, output_split_sizes=splits_out, input_split_sizes=splits_in)  # gloo
    assert torch.equal(to.cpu(), co), "all_to_all_single with split sizes differs from gloo"
    outs = [torch.empty(1000 + 7 * s, device=dev) for s in range(world)]
    ins = [torch.full((1000 + 7 * rank,), float(100 * rank + p), device=dev) for p in range(world)]
    dist.all_to_all(outs, ins)
    for s_ in range(world):
        assert bool((outs[s_] == float(100 * s_ + rank)).all()), "all_to_all list form"
    for root in sorted({0, world - 1}):
        g_in = H.to_cuda(O.make_inputs(world, 33333, "bf16", seed=202)[rank], "bf16", dev)
        g_out = [torch.empty_like(g_in) for _ in range(world)] if rank == root else None
        dist.gather(g_in, g_out, dst=root)
        if rank == root:
            for s_ in range(world):
                H.assert_same(_np(g_out[s_], "bf16"), O.make_inputs(world, 33333, "bf16", seed=202)[s_], "bf16", f"gather slot {s_}")
        sc_out = torch.empty(4097, device=dev)
        sc_in = [torch.full((4097,), float(p + 1), device=dev) for p in range(world)] if rank == root else None
        dist.scatter(sc_out, sc_in, src=root)
        assert bool((sc_out == float(rank + 1)).all()), "scatter"
        xr = O.make_inputs(world, 50001, "f32", seed=203)
        tr = H.to_cuda(xr[rank], "f32", dev)
        dist.reduce(tr, dst=root)
        if rank == root:
            H.assert_same(_np(tr, "f32"), O.reduce_rank_order(xr, "f32"), "f32", "reduce at the root")
        else:
            H.assert_same(_np(tr, "f32"), xr[rank], "f32", "reduce must leave non-root tensors alone")
    # ---- coalescing: the reference's per-parameter broadcast loop as ONE table launch ----------
    torch.manual_seed(rank)
    params = [torch.randn(n, device=dev) for n in (768, 768, 2304, 10, 3072, 1, 768 * 4)] + [torch.randn(1 << 20, device=dev)]
    ref = [p.cpu().clone() for p in params]
    n0 = B.launch_count()
    with dist._coalescing_manager(device=dev):
        for p in params:
            dist.broadcast(p, src=0)
    torch.cuda.synchronize()
    launches = B.launch_count() - n0
    for q in ref:
        dist.broadcast(q, src=0)  # gloo
    for p, q in zip(params, ref):
        assert torch.equal(p.cpu(), q), "coalesced broadcast differs from gloo"
    assert launches == 1, f"7 small + 1 large broadcast from one root took {launches} launches"
    # the same list as one call, from the last rank, mixed dtypes
    torch.manual_seed(100 + rank)
    mixed = [torch.randn(n, device=dev).to(dt) for n, dt in ((5, torch.float64), (70001, torch.bfloat16), (3 << 20, torch.float32), (1, torch.float16))]
    mixed.append(torch.randint(0, 255, (12345,), device=dev, dtype=torch.uint8))
    ref = [p.cpu().clone() for p in mixed]
    B.broadcast_tensors(mixed, src=world - 1)
    torch.cuda.synchronize()
    assert B.get_backend().last_algo() == "bcast_flat", B.get_backend().last_algo()
    for q in ref:
        dist.broadcast(q, src=world - 1)  # gloo
    for p, q in zip(mixed, ref):
        assert torch.equal(p.cpu().view(torch.uint8), q.view(torch.uint8)), "broadcast_tensors differs from gloo"
    # ---- sub-groups: dist.new_group over a subset gets its own communicator ----------------------
    if world >= 4:
        evens = dist.new_group(ranks=list(range(0, world, 2)), backend="cpu:gloo,cuda:b200")
        if rank % 2 == 0:
            xs = O.make_inputs(world, 100003, "f32", seed=204)
            ts = H.to_cuda(xs[rank], "f32", dev)
            dist.all_reduce(ts, group=evens)
            H.assert_same(_np(ts, "f32"), O.reduce_rank_order([xs[r] for r in range(0, world, 2)], "f32"), "f32", "sub-group all_reduce")
            tb = H.to_cuda(xs[rank], "f32", dev)
            dist.broadcast(tb, src=2, group=evens)
            H.assert_same(_np(tb, "f32"), xs[2], "f32", "sub-group broadcast")
        dist.barrier()

    # ---- async ops, futures, barrier, non-contiguous rejection -----------------------------
    x = O.make_inputs(world, 1 << 20, "f32", seed=3)
    t = H.to_cuda(x[rank], "f32", dev)
    work = dist.all_reduce(t, async_op=True)
    work.wait()
    H.assert_same(_np(t, "f32"), O.reduce_rank_order(x, "f32"), "f32", "async all_reduce")
    t2 = H.to_cuda(x[rank], "f32", dev)
    fut = dist.all_reduce(t2, async_op=True).get_future()
    fut.wait()
    assert torch.equal(fut.value()[0], t), "future value differs"
    dist.barrier()
    dist.barrier(device_ids=[rank])
    try:
        dist.all_reduce(torch.zeros(8, 8, device=dev).t())
        raise AssertionError("non-contiguous tensor was accepted")
    except ValueError:
        pass
    try:
        dist.all_reduce(torch.zeros(8, device=dev, dtype=torch.int32), op=dist.ReduceOp.AVG)
        raise AssertionError("integer AVG was accepted")
    except Exception as e:  # noqa: BLE001
        assert "AVG" in str(e) or "unsupported" in str(e), e

    # ---- DDP: gradient all_reduce through getFuture (config 4 pattern, tiny model) -------
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 8)).to(dev)
    ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank])
    torch.manual_seed(100 + rank)
    inp = torch.randn(16, 64, device=dev)
    ddp(inp).square().mean().backward()
    # reference: local grads of an un-wrapped copy, averaged over ranks through gloo
    torch.manual_seed(0)
    plain = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 8)).to(dev)
    plain(inp).square().mean().backward()
    for p, q in zip(ddp.module.parameters(), plain.parameters()):
        g = q.grad.detach().cpu()
        dist.all_reduce(g)
        g /= world
        assert torch.allclose(p.grad.cpu(), g, rtol=1e-5, atol=1e-6), "DDP gradient mismatch"

    # ---- BASELINE-size message through properties: 256 MiB fp32, integer-valued ----------
    numel = 64 << 20
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    big = torch.randint(-1000, 1000, (numel,), generator=g, device=dev, dtype=torch.int32)
    ref = big.clone()
    dist.all_reduce(ref)  # int32: exact
    bigf = big.to(torch.float32)
    dist.all_reduce(bigf)
    assert torch.equal(bigf, ref.to(torch.float32)), "256 MiB all_reduce: fp32 sum of small ints must be exact"
    chk = bigf.view(torch.int32).sum(dtype=torch.int64).cpu()
    allchk = _gather_cpu(dist, torch, chk.reshape(1))
    assert all(int(a) == int(allchk[0]) for a in allchk), "ranks disagree bit-wise on a 256 MiB all_reduce"
    del big, ref, bigf
    torch.cuda.synchronize()

    # ---- LAST (it kills the communicator): one rank passes symmetric memory, the others an
    # ordinary tensor.  The ranks then run different kernels; both kinds must notice and report
    # NBD_ERR_MISMATCH instead of returning garbage or hanging.
    be.set_tunable("zc_min_bytes", 256 << 10)
    be.set_tunable("auto_register_off", 1)
    symm = B.empty_symmetric(1 << 20, torch.float32)
    plain = torch.ones(1 << 20, device=dev)
    symm.fill_(1.0)
    dist.all_reduce(symm if rank == 0 else plain)
    torch.cuda.synchronize()
    try:
        dist.all_reduce(plain)
        torch.cuda.synchronize()
        raise AssertionError("mixed registered/unregistered buffers went unnoticed")
    except RuntimeError as e:
        assert "different registered buffers" in str(e) or "MISMATCH" in str(e).upper() or "aborted" in str(e), e
    W.shutdown_worker()


def main(rank: int, world: int, port: int) -> None:
    import faulthandler

    faulthandler.enable(all_threads=True)
    try:
        body(rank, world, port)
    except Exception:
        traceback.print_exc()
        raise
