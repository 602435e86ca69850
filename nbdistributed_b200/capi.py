"""ctypes binding of the C ABI declared in include/nbdist_b200.h.

This is the same surface the c10d glue calls; tests and the bench use it
directly for the emulated-rank (single GPU) mode.  There is deliberately no
fallback: if the native library is missing or a call fails, it raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnbdist_b200.so")

NBD_MAX_RANKS = 8
NBD_HANDLE_BYTES = 256

# nbd_dtype
F32, BF16, F16, F64, I32, I64, U8, I8, BYTES = range(9)
# nbd_redop (c10d::ReduceOp numbering)
SUM, AVG, PROD, MIN, MAX, BAND, BOR, BXOR = range(8)

DTYPE_NAMES = {F32: "f32", BF16: "bf16", F16: "f16", F64: "f64", I32: "i32", I64: "i64", U8: "u8", I8: "i8",
               BYTES: "bytes"}
OP_NAMES = {SUM: "sum", AVG: "avg", PROD: "prod", MIN: "min", MAX: "max", BAND: "band", BOR: "bor", BXOR: "bxor"}

# every symbol include/nbdist_b200.h declares (checked by tests/test_capi_symbols.py)
SYMBOLS = [
    "nbd_version", "nbd_error_string", "nbd_last_error", "nbd_dtype_size",
    "nbd_comm_create", "nbd_comm_get_handle", "nbd_comm_connect", "nbd_comm_destroy", "nbd_comm_abort",
    "nbd_comm_async_error", "nbd_comm_rank", "nbd_comm_world_size", "nbd_comm_set_timeout_ms",
    "nbd_comm_set_tunable", "nbd_comm_get_tunable", "nbd_comm_nvls_status",
    "nbd_mem_alloc", "nbd_mem_connect", "nbd_mem_free",
    "nbd_emu_create", "nbd_emu_destroy",
    "nbd_all_reduce", "nbd_broadcast", "nbd_all_gather", "nbd_reduce_scatter", "nbd_barrier",
    "nbd_reduce_scatter_list", "nbd_emu_reduce_scatter_list",
    "nbd_emu_all_reduce", "nbd_emu_broadcast", "nbd_emu_all_gather", "nbd_emu_reduce_scatter", "nbd_emu_barrier",
    "nbd_all_reduce_host", "nbd_emu_all_reduce_host",
    "nbd_p2p_batch", "nbd_send", "nbd_recv", "nbd_emu_p2p_batch",
    "nbd_broadcast_table", "nbd_emu_broadcast_table",
    "nbd_local_all_reduce_noalloc",
    "nbd_comm_last_zc_decision", "nbd_plan_query",
    "nbd_launch_count", "nbd_last_algo",
]
NOALLOC_MAX_BYTES = 1 << 20


class P2POp(C.Structure):
    """nbd_p2p_op"""
    _fields_ = [("kind", C.c_int), ("peer", C.c_int), ("buf", C.c_void_p), ("bytes", C.c_size_t)]


class BcastItem(C.Structure):
    """nbd_bcast_item"""
    _fields_ = [("buf", C.c_void_p), ("bytes", C.c_size_t), ("root", C.c_int)]


class NbdError(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__(f"nbdist_b200: {what} (code {code})")
        self.code = code


_lib = None


def load() -> C.CDLL:
    """Load libnbdist_b200.so (built in-tree by nbdistributed_b200._build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). nbdistributed_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, sz, i, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64
    pp = C.POINTER(C.c_void_p)
    sig = {
        "nbd_version": (i, []),
        "nbd_error_string": (C.c_char_p, [i]),
        "nbd_last_error": (C.c_char_p, []),
        "nbd_dtype_size": (sz, [i]),
        "nbd_comm_create": (i, [pp, i, i, i, sz]),
        "nbd_comm_get_handle": (i, [vp, vp]),
        "nbd_comm_connect": (i, [vp, vp]),
        "nbd_comm_destroy": (i, [vp]),
        "nbd_comm_abort": (i, [vp]),
        "nbd_comm_async_error": (i, [vp]),
        "nbd_comm_rank": (i, [vp]),
        "nbd_comm_world_size": (i, [vp]),
        "nbd_comm_set_timeout_ms": (i, [vp, u64]),
        "nbd_comm_set_tunable": (i, [vp, C.c_char_p, u64]),
        "nbd_comm_get_tunable": (i, [vp, C.c_char_p, C.POINTER(u64)]),
        "nbd_comm_nvls_status": (i, [vp, C.c_char_p, sz]),
        "nbd_mem_alloc": (i, [vp, sz, pp, vp]),
        "nbd_mem_connect": (i, [vp, vp, vp]),
        "nbd_mem_free": (i, [vp, vp]),
        "nbd_emu_create": (i, [pp, i, i, sz]),
        "nbd_emu_destroy": (i, [vp]),
        "nbd_all_reduce": (i, [vp, vp, vp, sz, i, i, vp]),
        "nbd_broadcast": (i, [vp, vp, sz, i, i, vp]),
        "nbd_all_gather": (i, [vp, vp, vp, sz, i, vp]),
        "nbd_reduce_scatter": (i, [vp, vp, vp, sz, i, i, vp]),
        "nbd_reduce_scatter_list": (i, [vp, pp, vp, sz, i, i, vp]),
        "nbd_emu_reduce_scatter_list": (i, [vp, pp, pp, sz, i, i, vp]),
        "nbd_barrier": (i, [vp, vp]),
        "nbd_emu_all_reduce": (i, [vp, pp, pp, sz, i, i, vp]),
        "nbd_emu_broadcast": (i, [vp, pp, sz, i, i, vp]),
        "nbd_emu_all_gather": (i, [vp, pp, pp, sz, i, vp]),
        "nbd_emu_reduce_scatter": (i, [vp, pp, pp, sz, i, i, vp]),
        "nbd_emu_barrier": (i, [vp, vp]),
        "nbd_all_reduce_host": (i, [vp, vp, vp, sz, i, i, vp]),
        "nbd_emu_all_reduce_host": (i, [vp, pp, pp, sz, i, i, vp]),
        "nbd_broadcast_table": (i, [vp, vp, i, vp]),
        "nbd_emu_broadcast_table": (i, [vp, vp, i, vp]),
        "nbd_p2p_batch": (i, [vp, vp, i, vp]),
        "nbd_send": (i, [vp, vp, sz, i, vp]),
        "nbd_recv": (i, [vp, vp, sz, i, vp]),
        "nbd_emu_p2p_batch": (i, [vp, vp, vp, vp]),
        "nbd_local_all_reduce_noalloc": (i, [pp, pp, i, sz, i, i, vp]),
        "nbd_comm_last_zc_decision": (i, [vp]),
        "nbd_plan_query": (i, [i, sz, i, i, i, sz, C.POINTER(u64)]),
        "nbd_launch_count": (u64, []),
        "nbd_last_algo": (C.c_char_p, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        lib = load()
        detail = lib.nbd_last_error().decode() or lib.nbd_error_string(rc).decode()
        raise NbdError(rc, detail)


def launch_count() -> int:
    return int(load().nbd_launch_count())


def plan_query(world: int, kind: str, nbytes: int, stage_bytes: int = 0, num_sms: int = 148, emulated: bool = False) -> dict:
    """The planner's choice for a collective (GPU-free): algorithm code, chunking, grid."""
    lib = load()
    out = (C.c_uint64 * 6)()
    code = {"all_reduce": 0, "reduce_scatter": 1, "all_gather": 2, "broadcast": 3, "flat_table": 4, "pipe": 5}[kind]
    check(lib.nbd_plan_query(world, stage_bytes, num_sms, int(emulated), code, nbytes, out))
    return {"algo": int(out[0]), "chunk_bytes": int(out[1]), "nchunks": int(out[2]), "nblocks": int(out[3]),
            "half_bytes": int(out[4]), "ll_off": int(out[5])}


def torch_dtype_code(dtype) -> int:
    import torch

    table = {torch.float32: F32, torch.bfloat16: BF16, torch.float16: F16, torch.float64: F64,
             torch.int32: I32, torch.int64: I64, torch.uint8: U8, torch.int8: I8}
    if dtype not in table:
        raise NbdError(-6, f"dtype {dtype} is not supported by nbdist_b200 reductions")
    return table[dtype]


def _ptr_array(tensors: Sequence) -> C.Array:
    arr = (C.c_void_p * len(tensors))()
    for k, t in enumerate(tensors):
        arr[k] = t.data_ptr()
    return arr


def _stream_handle(stream) -> int:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def local_all_reduce_noalloc(sends, recvs=None, op: int = SUM, stream=None) -> None:
    """nbd_local_all_reduce_noalloc on torch CUDA tensors of the current device (index = rank).
    Uses the library's fixed scratch array: no communicator, no allocation; at most
    NOALLOC_MAX_BYTES per rank; calls must not overlap on the GPU."""
    lib = load()
    recvs = sends if recvs is None else recvs
    t = sends[0]
    check(lib.nbd_local_all_reduce_noalloc(_ptr_array(sends), _ptr_array(recvs), len(sends), t.numel(),
                                           torch_dtype_code(t.dtype), op, _stream_handle(stream)))


class EmuGroup:
    """R ranks emulated on one GPU through nbd_emu_* (one cooperative launch per collective).

    Tensors are torch CUDA tensors on ``device``; index in each list = rank.
    """

    def __init__(self, world_size: int, device: int = 0, stage_bytes: int = 0):
        self.lib = load()
        self.world_size = world_size
        self.device = device
        arr = (C.c_void_p * NBD_MAX_RANKS)()
        check(self.lib.nbd_emu_create(arr, world_size, device, stage_bytes))
        self.comms = [arr[k] for k in range(world_size)]
        self._h = self.comms[0]

    def close(self) -> None:
        if self._h is not None:
            check(self.lib.nbd_emu_destroy(self._h))
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_tunable(self, name: str, value: int) -> None:
        for c in self.comms:
            check(self.lib.nbd_comm_set_tunable(c, name.encode(), value))

    def last_algo(self) -> str:
        return self.lib.nbd_last_algo(self._h).decode()

    def set_timeout_ms(self, ms: int) -> None:
        """Flag-wait watchdog of every rank (nbd_comm_set_timeout_ms)."""
        for c in self.comms:
            check(self.lib.nbd_comm_set_timeout_ms(c, ms))

    def abort(self) -> None:
        """nbd_comm_abort on every rank: running kernels leave their flag waits."""
        for c in self.comms:
            check(self.lib.nbd_comm_abort(c))

    def async_error(self) -> None:
        for c in self.comms:
            check(self.lib.nbd_comm_async_error(c))

    def all_reduce(self, sends, recvs=None, op: int = SUM, stream=None) -> None:
        recvs = sends if recvs is None else recvs
        t = sends[0]
        check(self.lib.nbd_emu_all_reduce(self._h, _ptr_array(sends), _ptr_array(recvs), t.numel(),
                                          torch_dtype_code(t.dtype), op, _stream_handle(stream)))

    def all_reduce_host(self, host_sends, host_recvs=None, op: int = SUM, stream=None) -> None:
        """all_reduce of (pinned) CPU tensors: chunked H2D / collective / D2H pipeline in the library."""
        host_recvs = host_sends if host_recvs is None else host_recvs
        t = host_sends[0]
        assert not t.is_cuda, "all_reduce_host takes CPU tensors"
        check(self.lib.nbd_emu_all_reduce_host(self._h, _ptr_array(host_sends), _ptr_array(host_recvs), t.numel(),
                                               torch_dtype_code(t.dtype), op, _stream_handle(stream)))

    def broadcast(self, bufs, root: int = 0, stream=None) -> None:
        t = bufs[0]
        check(self.lib.nbd_emu_broadcast(self._h, _ptr_array(bufs), t.numel() * t.element_size(), BYTES, root,
                                         _stream_handle(stream)))

    def all_gather(self, sends, recvs, stream=None) -> None:
        t = sends[0]
        check(self.lib.nbd_emu_all_gather(self._h, _ptr_array(sends), _ptr_array(recvs),
                                          t.numel() * t.element_size(), BYTES, _stream_handle(stream)))

    def reduce_scatter(self, sends, recvs, op: int = SUM, stream=None) -> None:
        t = recvs[0]
        check(self.lib.nbd_emu_reduce_scatter(self._h, _ptr_array(sends), _ptr_array(recvs), t.numel(),
                                              torch_dtype_code(t.dtype), op, _stream_handle(stream)))

    def reduce_scatter_list(self, send_lists, recvs, op: int = SUM, stream=None) -> None:
        """send_lists[r][p] = rank r's contribution to rank p's result (separate tensors)."""
        t = recvs[0]
        flat = [x for row in send_lists for x in row]
        check(self.lib.nbd_emu_reduce_scatter_list(self._h, _ptr_array(flat), _ptr_array(recvs), t.numel(),
                                                   torch_dtype_code(t.dtype), op, _stream_handle(stream)))

    def barrier(self, stream=None) -> None:
        check(self.lib.nbd_emu_barrier(self._h, _stream_handle(stream)))

    def broadcast_table(self, items, stream=None) -> None:
        """items = list of (per-rank tensor list, root): many broadcasts in one launch
        (nbd_emu_broadcast_table)."""
        n = self.world_size
        rows, ptrs = [], (C.c_void_p * n)()
        for r in range(n):
            arr = (BcastItem * max(1, len(items)))()
            for k, (tensors, root) in enumerate(items):
                t = tensors[r]
                arr[k].buf = t.data_ptr()
                arr[k].bytes = t.numel() * t.element_size()
                arr[k].root = root
            rows.append(arr)
            ptrs[r] = C.cast(arr, C.c_void_p)
        check(self.lib.nbd_emu_broadcast_table(self._h, ptrs, len(items), _stream_handle(stream)))

    def p2p_batch(self, ops_per_rank, stream=None) -> None:
        """ops_per_rank[r] = list of ("send" | "recv", peer, tensor) for rank r; every rank's batch
        runs in the same cooperative launch (nbd_emu_p2p_batch)."""
        n = self.world_size
        arrays, ptrs, counts = [], (C.c_void_p * n)(), (C.c_int * n)()
        for r in range(n):
            ops = ops_per_rank[r]
            arr = (P2POp * max(1, len(ops)))()
            for k, (kind, peer, t) in enumerate(ops):
                arr[k].kind = 0 if kind == "send" else 1
                arr[k].peer = peer
                arr[k].buf = t.data_ptr()
                arr[k].bytes = t.numel() * t.element_size()
            arrays.append(arr)
            ptrs[r] = C.cast(arr, C.c_void_p)
            counts[r] = len(ops)
        check(self.lib.nbd_emu_p2p_batch(self._h, ptrs, counts, _stream_handle(stream)))
