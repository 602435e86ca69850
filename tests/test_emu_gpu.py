"""Parity of the CUDA kernels against the oracle, on ONE GPU.

All R ranks are emulated by a single cooperative launch (nbd_emu_* entry points of the
C ABI): the device code, flag protocol, staging layout and chunk pipeline are exactly the
ones the multi-process backend runs, only the "peer" arenas live on the same GPU.
Bar: bit-exact against the oracle's rank-order restatement for every dtype (the kernels
implement that fixed order), and within the north_star tolerance against gloo's order.
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

    assert torch.cuda.is_available(), "these tests need a GPU (run with -m gpu on the B200 box)"
    return torch


@pytest.fixture(scope="module")
def groups(torch_cuda):
    """One emulated group per world size, small staging so that modest sizes already pipeline."""
    from nbdistributed_b200 import capi

    made = {}

    def get(world, stage_bytes=8 << 20):
        key = (world, stage_bytes)
        if key not in made:
            made[key] = capi.EmuGroup(world, device=0, stage_bytes=stage_bytes)
        return made[key]

    yield get
    for g in made.values():
        g.close()


class tuned:
    """Set tunables on a group for the duration of a with-block (0 restores the default)."""

    def __init__(self, grp, **kv):
        self.grp, self.kv = grp, kv

    def __enter__(self):
        for k, v in self.kv.items():
            self.grp.set_tunable(k, v)
        return self.grp

    def __exit__(self, *exc):
        for k in self.kv:
            self.grp.set_tunable(k, 0)


# staged = through the arena (what unregistered buffers of real ranks use);
# zerocopy = straight between the ranks' buffers (registered memory; always possible when emulated)
PATHS = {"staged": {"zero_copy_off": 1}, "zerocopy": {"zc_min_bytes": 1}}


def _run_all_reduce(torch, grp, x, dtype, op="sum", inplace=True):
    sends = [H.to_cuda(a, dtype) for a in x]
    recvs = sends if inplace else [torch.empty_like(s) for s in sends]
    grp.all_reduce(sends, recvs, op=OPCODE[op])
    torch.cuda.synchronize()
    grp.async_error()
    return [H.to_numpy(r, dtype) for r in recvs], [H.to_numpy(s, dtype) for s in sends]


@pytest.mark.parametrize("path", ["staged", "zerocopy"])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype", ["f32", "bf16", "f16", "f64", "i32", "i64", "u8", "i8"])
@pytest.mark.parametrize("numel", [1, 3, 127, 128, 1024, 4096 + 5, (1 << 18) + 1])
def test_all_reduce_sum_matches_oracle(torch_cuda, groups, world, dtype, numel, path):
    x = O.make_inputs(world, numel, dtype)
    with tuned(groups(world), **PATHS[path]) as grp:
        got, _ = _run_all_reduce(torch_cuda, grp, x, dtype)
        assert (grp.last_algo() == "ar_zerocopy") == (path == "zerocopy")
        got2, sends = _run_all_reduce(torch_cuda, grp, x, dtype, inplace=False)
    want = O.all_reduce(x, dtype, "sum", order="rank")
    for r in range(world):
        H.assert_same(got[r], want[r], dtype, f"rank {r}")
        H.assert_same(got2[r], want[r], dtype, f"out of place, rank {r}")
        H.assert_same(sends[r], x[r], dtype, f"send buffer of rank {r} was modified")


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_all_reduce_within_tolerance_of_gloo_order(torch_cuda, groups, world, dtype):
    numel = (1 << 16) + 13
    x = O.make_inputs(world, numel, dtype)
    got, _ = _run_all_reduce(torch_cuda, groups(world), x, dtype)
    ref = O.to_f64(O.reduce_gloo_order(x, dtype), dtype)
    tol = O.float_sum_tolerance(x, dtype)
    err = np.abs(O.to_f64(got[0], dtype) - ref)
    assert (err <= tol).all(), f"max err/tol {np.max(err / tol)}"
    if world == 2:  # a+b is commutative: bit-exact with the reference's gloo path
        H.assert_same(got[0], O.reduce_gloo_order(x, dtype), dtype)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("numel", [1 << 20, (3 << 20) + 7])
@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_all_reduce_two_shot_pipelined(torch_cuda, groups, world, numel, dtype):
    """Messages above the one-shot limit and above the 8 MiB staging: two-shot, several chunks."""
    grp = groups(world)
    grp.set_tunable("ar1_max_bytes", 1024)
    grp.set_tunable("zero_copy_off", 1)
    try:
        x = O.make_inputs(world, numel, dtype)
        for inplace in (True, False):
            got, sends = _run_all_reduce(torch_cuda, grp, x, dtype, inplace=inplace)
            assert grp.last_algo() == "ar_twoshot"
            want = O.all_reduce(x, dtype, "sum")
            for r in range(world):
                H.assert_same(got[r], want[r], dtype, f"rank {r} inplace={inplace}")
                if not inplace:
                    H.assert_same(sends[r], x[r], dtype, f"send buffer of rank {r} was modified")
    finally:
        grp.set_tunable("ar1_max_bytes", 0)
        grp.set_tunable("zero_copy_off", 0)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_all_reduce_one_shot_pipelined(torch_cuda, groups, world):
    grp = groups(world)
    grp.set_tunable("ar1_max_bytes", 1 << 40)
    grp.set_tunable("chunk_bytes", 64 << 10)
    grp.set_tunable("zero_copy_off", 1)
    try:
        x = O.make_inputs(world, (1 << 18) + 3, "f32")
        got, _ = _run_all_reduce(torch_cuda, grp, x, "f32")
        assert grp.last_algo() == "ar_oneshot"
        want = O.all_reduce(x, "f32", "sum")
        for r in range(world):
            H.assert_same(got[r], want[r], "f32", f"rank {r}")
    finally:
        grp.set_tunable("ar1_max_bytes", 0)
        grp.set_tunable("chunk_bytes", 0)
        grp.set_tunable("zero_copy_off", 0)


@pytest.mark.parametrize("dtype,op", [("f32", "prod"), ("f32", "min"), ("f32", "max"), ("f32", "avg"),
                                      ("bf16", "max"), ("bf16", "avg"), ("f64", "min"), ("f16", "prod"),
                                      ("i32", "prod"), ("i32", "min"), ("i32", "max"), ("i32", "band"),
                                      ("i32", "bor"), ("i32", "bxor"), ("i64", "max"), ("i64", "bxor"),
                                      ("u8", "max"), ("u8", "min"), ("i8", "min"), ("i8", "max"), ("u8", "bor"),
                                      ("i8", "prod")])
@pytest.mark.parametrize("algo", ["ll", "oneshot", "twoshot", "zerocopy"])
def test_all_reduce_other_ops(torch_cuda, groups, dtype, op, algo):
    world = 4
    knobs = {"ll": {"zero_copy_off": 1}, "oneshot": {"zero_copy_off": 1, "ll_max_bytes": 1},
             "twoshot": {"zero_copy_off": 1, "ar1_max_bytes": 1024, "ll_max_bytes": 1},
             "zerocopy": {"zc_min_bytes": 1}}[algo]
    with tuned(groups(world), **knobs) as grp:
        x = O.make_inputs(world, 2000 if algo == "ll" else 5000, dtype, scale=1.0)
        got, _ = _run_all_reduce(torch_cuda, grp, x, dtype, op=op)
        assert grp.last_algo() == "ar_" + algo
        want = O.all_reduce(x, dtype, op)
        for r in range(world):
            H.assert_same(got[r], want[r], dtype, f"{op} rank {r}")


def test_all_reduce_special_values(torch_cuda, groups):
    """inf/NaN propagation, denormals, and a cancellation vector that exposes the order."""
    world = 4
    base = np.array([1e8, 1.0, -1e8, np.inf, -np.inf, np.nan, 1e-45, -0.0, 3.0e38, 1.5], dtype=np.float32)
    x = [np.roll(base, r).copy() for r in range(world)]
    got, _ = _run_all_reduce(torch_cuda, groups(world), x, "f32")
    want = O.all_reduce(x, "f32", "sum")
    for r in range(world):
        H.assert_same(got[r], want[r], "f32", f"rank {r}")


def test_all_reduce_known_answers(torch_cuda, groups):
    """ones / rank constants: sum = N*x exactly in every dtype."""
    for world in (2, 8):
        for dtype in ("f32", "bf16", "i32", "u8"):
            n = 1000
            x = [np.full(n, r + 1, dtype=np.float32) for r in range(world)]
            if dtype == "bf16":
                x = [O.f32_to_bf16(a) for a in x]
            else:
                x = [a.astype(O.NP_DTYPE[dtype]) for a in x]
            got, _ = _run_all_reduce(torch_cuda, groups(world), x, dtype)
            total = world * (world + 1) // 2
            want = O.f32_to_bf16(np.full(n, total, np.float32)) if dtype == "bf16" else np.full(n, total, O.NP_DTYPE[dtype])
            H.assert_same(got[world - 1], want, dtype)


def test_all_reduce_misaligned_views(torch_cuda, groups):
    """storage_offset views: base pointer not 16-byte aligned, ragged tail."""
    torch = torch_cuda
    world = 4
    grp = groups(world)
    for zc, (dtype, off) in [(z, c) for z in (0, 1) for c in (("f32", 1), ("bf16", 3), ("u8", 5), ("f64", 1))]:
        grp.set_tunable("zc_min_bytes", 1 if zc else 0)
        grp.set_tunable("zero_copy_off", 0 if zc else 1)
        n = 3001
        x = O.make_inputs(world, n + off, dtype)
        full = [H.to_cuda(a, dtype) for a in x]
        views = [t[off:] for t in full]
        assert views[0].data_ptr() % 16 != 0
        grp.all_reduce(views, views)
        torch.cuda.synchronize()
        want = O.all_reduce([a[off:] for a in x], dtype, "sum")
        for r in range(world):
            H.assert_same(H.to_numpy(views[r], dtype), want[r], dtype, f"{dtype} rank {r}")
            H.assert_same(H.to_numpy(full[r][:off], dtype), x[r][:off], dtype, "bytes before the view were touched")
    grp.set_tunable("zc_min_bytes", 0)
    grp.set_tunable("zero_copy_off", 1)
    grp.set_tunable("ar1_max_bytes", 1024)
    try:
        x = O.make_inputs(world, 100003, "f32")
        full = [H.to_cuda(a, "f32") for a in x]
        views = [t[1:-1] for t in full]
        grp.all_reduce(views, views)
        torch.cuda.synchronize()
        want = O.all_reduce([a[1:-1] for a in x], "f32", "sum")
        for r in range(world):
            H.assert_same(H.to_numpy(views[r], "f32"), want[r], "f32", f"two-shot rank {r}")
            assert H.to_numpy(full[r], "f32")[0] == x[r][0] and H.to_numpy(full[r], "f32")[-1] == x[r][-1]
    finally:
        grp.set_tunable("ar1_max_bytes", 0)
        grp.set_tunable("zero_copy_off", 0)


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("nbytes", [1, 40, 400, 4096, 12 * 1024 + 3, (1 << 20) + 17, (5 << 20) + 1])
def test_broadcast_bit_exact(torch_cuda, groups, world, nbytes):
    torch = torch_cuda
    grp = groups(world)
    for root in sorted({0, world - 1, world // 2}):
        x = O.make_inputs(world, nbytes, "u8", seed=99 + root)
        bufs = [H.to_cuda(a, "u8") for a in x]
        grp.broadcast(bufs, root=root)
        torch.cuda.synchronize()
        grp.async_error()
        want = O.broadcast(x, root)
        for r in range(world):
            H.assert_same(H.to_numpy(bufs[r], "u8"), want[r], "u8", f"root {root} rank {r} algo {grp.last_algo()}")


@pytest.mark.parametrize("world", [3, 4, 8])
def test_broadcast_both_algorithms(torch_cuda, groups, world):
    torch = torch_cuda
    grp = groups(world)
    nbytes = (2 << 20) + 5
    for limit, algo in ((1 << 40, "bcast_direct"), (1, "bcast_scatter")):
        grp.set_tunable("bcast_direct_max_bytes", limit)
        try:
            x = O.make_inputs(world, nbytes, "u8", seed=5)
            bufs = [H.to_cuda(a, "u8") for a in x]
            grp.broadcast(bufs, root=1)
            torch.cuda.synchronize()
            assert grp.last_algo() == algo
            for r in range(world):
                H.assert_same(H.to_numpy(bufs[r], "u8"), x[1], "u8", f"{algo} rank {r}")
        finally:
            grp.set_tunable("bcast_direct_max_bytes", 0)


@pytest.mark.parametrize("path", ["staged", "zerocopy"])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype,numel", [("f32", 1), ("f32", 25), ("bf16", 4099), ("i64", 1 << 16), ("u8", (3 << 20) + 1)])
def test_all_gather_bit_exact(torch_cuda, groups, world, dtype, numel, path):
    torch = torch_cuda
    grp = groups(world)
    for k, v in PATHS[path].items():
        grp.set_tunable(k, v)
    x = O.make_inputs(world, numel, dtype)
    sends = [H.to_cuda(a, dtype) for a in x]
    recvs = [torch.zeros(world * numel, dtype=sends[0].dtype, device="cuda:0") for _ in range(world)]
    grp.all_gather(sends, recvs)
    torch.cuda.synchronize()
    grp.async_error()
    want = O.all_gather(x)
    for r in range(world):
        H.assert_same(H.to_numpy(recvs[r], dtype), want[r], dtype, f"rank {r}")
    # in place: send aliases its own slot of recv
    recvs2 = [torch.zeros(world * numel, dtype=sends[0].dtype, device="cuda:0") for _ in range(world)]
    for r in range(world):
        recvs2[r][r * numel:(r + 1) * numel].copy_(sends[r])
    grp.all_gather([recvs2[r][r * numel:(r + 1) * numel] for r in range(world)], recvs2)
    torch.cuda.synchronize()
    assert grp.last_algo() == ("all_gather_zerocopy_pull" if path == "zerocopy" else "all_gather_pull")
    for k in PATHS[path]:
        grp.set_tunable(k, 0)
    for r in range(world):
        H.assert_same(H.to_numpy(recvs2[r], dtype), want[r], dtype, f"in-place rank {r}")


@pytest.mark.parametrize("path", ["staged", "zerocopy"])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype,numel", [("f32", 1), ("f32", 1001), ("bf16", 4099), ("i32", 1 << 15), ("bf16", (1 << 20) + 8)])
def test_reduce_scatter_matches_oracle(torch_cuda, groups, world, dtype, numel, path):
    torch = torch_cuda
    grp = groups(world)
    for k, v in PATHS[path].items():
        grp.set_tunable(k, v)
    x = O.make_inputs(world, world * numel, dtype)
    sends = [H.to_cuda(a, dtype) for a in x]
    recvs = [torch.zeros(numel, dtype=sends[0].dtype, device="cuda:0") for _ in range(world)]
    grp.reduce_scatter(sends, recvs)
    torch.cuda.synchronize()
    grp.async_error()
    assert grp.last_algo() == ("reduce_scatter_zerocopy" if path == "zerocopy" else "reduce_scatter_push")
    for k in PATHS[path]:
        grp.set_tunable(k, 0)
    want = O.reduce_scatter(x, dtype, "sum")
    for r in range(world):
        H.assert_same(H.to_numpy(recvs[r], dtype), want[r], dtype, f"rank {r}")
        H.assert_same(H.to_numpy(sends[r], dtype), x[r], dtype, f"send of rank {r} modified")


def test_reduce_scatter_then_all_gather_equals_all_reduce(torch_cuda, groups):
    """ZeRO-style round trip (BASELINE config 5 shape, scaled): RS + AG == AR, bit for bit."""
    torch = torch_cuda
    world, shard = 8, (1 << 17) + 16
    grp = groups(world)
    x = O.make_inputs(world, world * shard, "bf16", scale=0.01)
    sends = [H.to_cuda(a, "bf16") for a in x]
    shards = [torch.empty(shard, dtype=torch.bfloat16, device="cuda:0") for _ in range(world)]
    full = [torch.empty(world * shard, dtype=torch.bfloat16, device="cuda:0") for _ in range(world)]
    grp.reduce_scatter(sends, shards)
    grp.all_gather(shards, full)
    ar = [s.clone() for s in sends]
    grp.set_tunable("ar1_max_bytes", 1024)
    grp.all_reduce(ar, ar)
    grp.set_tunable("ar1_max_bytes", 0)
    torch.cuda.synchronize()
    grp.async_error()
    want = O.all_reduce(x, "bf16", "sum")[0]
    for r in range(world):
        H.assert_same(H.to_numpy(full[r], "bf16"), want, "bf16", f"rs+ag rank {r}")
        H.assert_same(H.to_numpy(ar[r], "bf16"), want, "bf16", f"ar rank {r}")


def test_back_to_back_mixed_collectives_reuse_staging_safely(torch_cuda, groups):
    """Many different collectives in a row without host synchronisation: exercises the epoch
    counters and the alternation of staging halves across launches and grid sizes."""
    torch = torch_cuda
    world = 4
    grp = groups(world)
    rng = np.random.default_rng(7)
    expected, produced = [], []
    for it in range(40):
        kind = ("ar", "bc", "ag", "rs")[it % 4]
        n = int(rng.integers(1, 300000))
        if kind == "ar":
            x = O.make_inputs(world, n, "f32", seed=it)
            t = [H.to_cuda(a, "f32") for a in x]
            grp.all_reduce(t, t)
            expected.append(O.all_reduce(x, "f32"))
            produced.append((t, "f32"))
        elif kind == "bc":
            x = O.make_inputs(world, n, "u8", seed=it)
            t = [H.to_cuda(a, "u8") for a in x]
            grp.broadcast(t, root=it % world)
            expected.append(O.broadcast(x, it % world))
            produced.append((t, "u8"))
        elif kind == "ag":
            x = O.make_inputs(world, n, "i32", seed=it)
            s = [H.to_cuda(a, "i32") for a in x]
            t = [torch.empty(world * n, dtype=torch.int32, device="cuda:0") for _ in range(world)]
            grp.all_gather(s, t)
            expected.append(O.all_gather(x))
            produced.append((t, "i32"))
        else:
            x = O.make_inputs(world, world * n, "bf16", seed=it)
            s = [H.to_cuda(a, "bf16") for a in x]
            t = [torch.empty(n, dtype=torch.bfloat16, device="cuda:0") for _ in range(world)]
            grp.reduce_scatter(s, t)
            expected.append(O.reduce_scatter(x, "bf16"))
            produced.append((t, "bf16"))
        if it % 7 == 3:
            grp.barrier()
    torch.cuda.synchronize()
    grp.async_error()
    for i, (want, (t, dtype)) in enumerate(zip(expected, produced)):
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], dtype), want[r], dtype, f"op {i} rank {r}")


def test_large_all_reduce_properties(torch_cuda):
    """BASELINE-size message (256 MiB fp32 per rank, 4 ranks): checked through size-independent
    properties — all ranks bit-identical, linearity in exactly representable inputs, and a
    sampled comparison with the oracle."""
    torch = torch_cuda
    from nbdistributed_b200 import capi

    world, numel = 4, 64 << 20
    with capi.EmuGroup(world, device=0) as grp:
        g = torch.Generator(device="cuda:0")
        t = []
        for r in range(world):
            g.manual_seed(1234 + r)
            t.append(torch.randint(-1000, 1000, (numel,), generator=g, device="cuda:0", dtype=torch.int32).to(torch.float32))
        keep = [a.clone() for a in t]
        want = keep[0] + keep[1] + keep[2] + keep[3]  # small integers: exact in fp32 in any order
        for off, algo in ((1, "ar_twoshot"), (0, "ar_zerocopy")):
            grp.set_tunable("zero_copy_off", off)
            for r in range(world):
                t[r].copy_(keep[r])
            grp.all_reduce(t, t)
            torch.cuda.synchronize()
            grp.async_error()
            assert grp.last_algo() == algo
            for r in range(world):
                assert torch.equal(t[r], want), f"{algo} rank {r}"
        idx = np.random.default_rng(0).integers(0, numel, 4096)
        sample = [H.to_numpy(k[idx], "f32") for k in keep]
        H.assert_same(H.to_numpy(t[0][idx], "f32"), O.reduce_rank_order(sample, "f32"), "f32")


def test_invalid_arguments_are_rejected(torch_cuda, groups):
    from nbdistributed_b200 import capi

    torch = torch_cuda
    grp = groups(2)
    t = [torch.zeros(8, device="cuda:0") for _ in range(2)]
    with pytest.raises(capi.NbdError):
        grp.all_reduce(t, t, op=OPCODE["band"])  # bitwise on floats
    ti = [torch.zeros(8, device="cuda:0", dtype=torch.int32) for _ in range(2)]
    with pytest.raises(capi.NbdError):
        grp.all_reduce(ti, ti, op=OPCODE["avg"])  # gloo rejects AVG on ints too
    with pytest.raises(capi.NbdError):
        grp.broadcast(t, root=5)
    empty = [torch.zeros(0, device="cuda:0") for _ in range(2)]
    grp.all_reduce(empty, empty)  # empty tensors are a no-op, not an error
    torch.cuda.synchronize()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode,tile", [(1, 0), (2, 0), (2, 64 << 10), (2, 8192)])
def test_zero_copy_all_reduce_variants(torch_cuda, groups, world, mode, tile):
    """Fused and tiled (read-then-push) zero-copy all_reduce, in place and out of place."""
    with tuned(groups(world), zc_min_bytes=1, zc_ar_mode=mode, zc_tile_bytes=tile) as grp:
        for dtype, numel in (("f32", (1 << 20) + 7), ("bf16", 300001)):
            x = O.make_inputs(world, numel, dtype)
            want = O.all_reduce(x, dtype, "sum")
            for inplace in (True, False):
                got, sends = _run_all_reduce(torch_cuda, grp, x, dtype, inplace=inplace)
                assert grp.last_algo() == "ar_zerocopy"
                for r in range(world):
                    H.assert_same(got[r], want[r], dtype, f"mode {mode} tile {tile} rank {r} inplace={inplace}")
                    if not inplace:
                        H.assert_same(sends[r], x[r], dtype, "send buffer modified")


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("dtype,numel,chunk", [("f32", (3 << 20) + 5, 1 << 20), ("bf16", 100001, 64 << 10), ("i32", 777, 0)])
def test_all_reduce_host_buffers(torch_cuda, groups, world, dtype, numel, chunk):
    """nbd_emu_all_reduce_host: pinned host tensors in, pinned host tensors out, pipelined in chunks
    (several chunks, ragged last one, more chunks than bounce buffers)."""
    torch = torch_cuda
    with tuned(groups(world), host_chunk_bytes=chunk) as grp:
        x = O.make_inputs(world, numel, dtype)
        tdt = H.torch_dtype(dtype)
        hin = []
        for a in x:
            t = torch.empty(numel, dtype=tdt).pin_memory()
            t.copy_(H.to_cuda(a, dtype).cpu())
            hin.append(t)
        hout = [torch.zeros(numel, dtype=tdt).pin_memory() for _ in range(world)]
        for _ in range(2):  # twice: the bounce-buffer ring is reused across calls
            grp.all_reduce_host(hin, hout)
        torch.cuda.synchronize()
        grp.async_error()
        want = O.reduce_rank_order(x, dtype)
        for r in range(world):
            H.assert_same(H.to_numpy(hout[r], dtype), want, dtype, f"rank {r}")
            H.assert_same(H.to_numpy(hin[r], dtype), x[r], dtype, f"host input of rank {r} modified")
        grp.all_reduce_host(hin)  # in place
        torch.cuda.synchronize()
        H.assert_same(H.to_numpy(hin[0], dtype), want, dtype, "in place")


@pytest.mark.parametrize("world", [2, 3, 8])
def test_ll_all_reduce_back_to_back(torch_cuda, groups, world):
    """The flag-in-data path: many small all_reduces in a row without host synchronisation
    (buffer parity, sequence flags), mixed sizes and dtypes, in place and out of place."""
    torch = torch_cuda
    grp = groups(world)
    rng = np.random.default_rng(3)
    runs = []
    for it in range(60):
        dtype = ("f32", "bf16", "i64", "u8")[it % 4]
        n = int(rng.integers(1, (32 << 10) // O.NP_DTYPE[dtype]().itemsize))
        x = O.make_inputs(world, n, dtype, seed=it)
        t = [H.to_cuda(a, dtype) for a in x]
        out = t if it % 3 else [torch.empty_like(a) for a in t]
        grp.all_reduce(t, out)
        assert grp.last_algo() == "ar_ll"
        runs.append((x, out, dtype))
    torch.cuda.synchronize()
    grp.async_error()
    for i, (x, out, dtype) in enumerate(runs):
        want = O.reduce_rank_order(x, dtype)
        for r in range(world):
            H.assert_same(H.to_numpy(out[r], dtype), want, dtype, f"op {i} rank {r}")


@pytest.mark.parametrize("world", [2, 3, 8])
def test_ll_broadcast_and_all_reduce_interleaved(torch_cuda, groups, world):
    """The LL (flag-in-data) broadcast shares the two parity buffers with the LL all_reduce.  A
    root needs nothing from its peers, so it could lap them: 60 back-to-back LL operations with
    rotating roots and sizes, no host sync, exercise the ack back-pressure (kLLAckRow)."""
    torch = torch_cuda
    grp = groups(world)
    rng = np.random.default_rng(11)
    cases, outs = [], []
    for it in range(60):
        nbytes = int(rng.integers(1, 32 << 10))
        if it % 3 == 2:
            x = O.make_inputs(world, max(1, nbytes // 4), "f32", seed=300 + it)
            t = [H.to_cuda(a, "f32") for a in x]
            grp.all_reduce(t, t)
            assert grp.last_algo() == "ar_ll"
            cases.append(("ar", x, None))
        else:
            root = it % world
            x = O.make_inputs(world, nbytes, "u8", seed=300 + it)
            t = [H.to_cuda(a, "u8") for a in x]
            grp.broadcast(t, root=root)
            assert grp.last_algo() == "bcast_ll", grp.last_algo()
            cases.append(("bc", x, root))
        outs.append(t)
    torch.cuda.synchronize()
    grp.async_error()
    for (kind, x, root), t in zip(cases, outs):
        for r in range(world):
            if kind == "ar":
                H.assert_same(H.to_numpy(t[r], "f32"), O.reduce_rank_order(x, "f32"), "f32", f"LL all_reduce rank {r}")
            else:
                H.assert_same(H.to_numpy(t[r], "u8"), x[root], "u8", f"LL broadcast root {root} rank {r}")


@pytest.mark.parametrize("world", [2, 3, 8])
def test_ll_large_area_up_to_256kib(torch_cuda, world):
    """Arenas of 64 MiB and more carry an 8 MiB LL area (limit 256 KiB per operation; the planner's
    default threshold is lower, "ll_max_bytes" raises it).  Several units per thread: all packets
    are sent before any is awaited.  Back-to-back all_reduces and broadcasts of 1 B .. 256 KiB,
    in place and out of place, misaligned views, no host sync."""
    from nbdistributed_b200 import capi

    torch = torch_cuda
    rng = np.random.default_rng(17 + world)
    with capi.EmuGroup(world, device=0, stage_bytes=64 << 20) as grp:
        grp.set_tunable("ll_max_bytes", 256 << 10)
        grp.set_tunable("zero_copy_off", 1)  # emulated ranks would take 256 KiB zero-copy
        runs = []
        for it in range(40):
            nbytes = (256 << 10) if it == 0 else int(rng.integers(1, 256 << 10))
            if it % 4 == 3:
                root = it % world
                x = O.make_inputs(world, nbytes, "u8", seed=500 + it)
                back = [torch.zeros(nbytes + 32, dtype=torch.uint8, device="cuda:0") for _ in range(world)]
                off = it % 3
                t = [b[off:off + nbytes] for b in back]
                for v, a in zip(t, x):
                    v.copy_(H.to_cuda(a, "u8"))
                grp.broadcast(t, root=root)
                assert grp.last_algo() == "bcast_ll", grp.last_algo()
                runs.append(("bc", "u8", x, root, t))
            else:
                dtype = ("f32", "bf16", "i64")[it % 3]
                n = max(1, nbytes // O.NP_DTYPE[dtype]().itemsize)
                x = O.make_inputs(world, n, dtype, seed=500 + it)
                t = [H.to_cuda(a, dtype) for a in x]
                out = t if it % 2 else [torch.empty_like(a) for a in t]
                grp.all_reduce(t, out)
                assert grp.last_algo() == "ar_ll", grp.last_algo()
                runs.append(("ar", dtype, x, None, out))
        torch.cuda.synchronize()
        grp.async_error()
        for i, (kind, dtype, x, root, out) in enumerate(runs):
            want = O.reduce_rank_order(x, dtype) if kind == "ar" else x[root]
            for r in range(world):
                H.assert_same(H.to_numpy(out[r], dtype), want, dtype, f"op {i} ({kind}) rank {r}")
        # above the limit the planner falls back to the barrier kernels
        x = O.make_inputs(world, (256 << 10) // 4 + 4, "f32", seed=9)
        t = [H.to_cuda(a, "f32") for a in x]
        grp.all_reduce(t, t)
        assert grp.last_algo() != "ar_ll"
        torch.cuda.synchronize()
        for r in range(world):
            H.assert_same(H.to_numpy(t[r], "f32"), O.reduce_rank_order(x, "f32"), "f32", f"above the limit, rank {r}")


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("dtype,op", [("f32", "sum"), ("bf16", "sum"), ("i32", "sum"), ("f32", "max"), ("bf16", "avg"), ("u8", "bxor")])
@pytest.mark.parametrize("numel", [1, 4099, (1 << 19) + 5])
def test_reduce_scatter_list_form_matches_oracle(torch_cuda, groups, world, dtype, op, numel):
    """dist.reduce_scatter(out, [in_0 .. in_{N-1}]): N separate contributions per rank, taken
    where they lie (nbd_reduce_scatter_list) - same bits as the tensor form on their
    concatenation, contributions untouched; every other contribution sits at a misaligned
    address."""
    from nbdistributed_b200 import capi

    torch = torch_cuda
    opcode = {"sum": 0, "avg": 1, "max": 4, "bxor": 7}[op]
    grp = groups(world)
    x = O.make_inputs(world, world * numel, dtype, seed=21)
    itemsize = O.NP_DTYPE[dtype]().itemsize
    lists, keep = [], []
    for r in range(world):
        row = []
        for p in range(world):
            part = np.ascontiguousarray(x[r][p * numel:(p + 1) * numel])
            t = H.to_cuda(part, dtype)
            if (r + p) % 2 and itemsize < 16:  # a view one element into a larger buffer
                back = torch.empty(numel + 1, dtype=t.dtype, device="cuda:0")
                back[1:].copy_(t)
                keep.append(back)
                t = back[1:]
            row.append(t)
        lists.append(row)
    recvs = [torch.zeros(numel, dtype=lists[0][0].dtype, device="cuda:0") for _ in range(world)]
    n0 = capi.launch_count()
    grp.reduce_scatter_list(lists, recvs, op=opcode)
    assert capi.launch_count() - n0 == 1
    assert grp.last_algo() == "reduce_scatter_list"
    torch.cuda.synchronize()
    grp.async_error()
    want = O.reduce_scatter(x, dtype, op)
    for r in range(world):
        H.assert_same(H.to_numpy(recvs[r], dtype), want[r], dtype, f"rank {r}")
        for p in range(world):
            H.assert_same(H.to_numpy(lists[r][p], dtype), x[r][p * numel:(p + 1) * numel], dtype, f"contribution {r}->{p} modified")
