"""The C-ABI library on a machine without a GPU: it loads, exports exactly what
include/nbdist_b200.h declares, and its GPU-free entry points behave."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "nbdist_b200.h")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nbd_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    from nbdistributed_b200 import capi

    assert _declared() == sorted(capi.SYMBOLS)


def test_library_exports_every_declared_symbol(native_lib):
    from nbdistributed_b200 import capi

    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (nbd_[a-z0-9_]+)", out))
    assert set(_declared()) <= exported, sorted(set(_declared()) - exported)
    # the statically linked CUDA runtime stays private
    assert not re.search(r" T cuda[A-Z]", out)
    for name in _declared():
        assert getattr(native_lib, name) is not None


def test_gpu_free_entry_points(native_lib):
    from nbdistributed_b200 import capi

    assert native_lib.nbd_version() == 100
    sizes = {capi.F32: 4, capi.BF16: 2, capi.F16: 2, capi.F64: 8, capi.I32: 4, capi.I64: 8, capi.U8: 1, capi.I8: 1,
             capi.BYTES: 1}
    for code, size in sizes.items():
        assert native_lib.nbd_dtype_size(code) == size
    assert native_lib.nbd_dtype_size(99) == 0
    assert native_lib.nbd_error_string(0) == b"ok"
    assert b"timed out" in native_lib.nbd_error_string(-4)
    assert native_lib.nbd_comm_rank(None) == -1
    assert native_lib.nbd_launch_count() >= 0


def test_invalid_arguments_fail_cleanly_without_a_gpu(native_lib):
    from nbdistributed_b200 import capi

    comm = C.c_void_p()
    rc = native_lib.nbd_comm_create(C.byref(comm), 0, 9, 0, 0)  # world_size > NBD_MAX_RANKS
    assert rc == -1 and b"world_size" in native_lib.nbd_last_error()
    rc = native_lib.nbd_comm_create(C.byref(comm), 3, 2, 0, 0)  # rank out of range
    assert rc == -1
    with pytest.raises(capi.NbdError):
        capi.check(native_lib.nbd_all_reduce(None, None, None, 4, capi.F32, capi.SUM, None))
    with pytest.raises(capi.NbdError):
        capi.check(native_lib.nbd_local_all_reduce_noalloc(None, None, 2, (2 << 20), capi.F32, capi.SUM, None))
    assert b"NBD_NOALLOC_MAX_BYTES" in native_lib.nbd_last_error()


def test_no_cpu_fallback_when_cuda_is_absent(native_lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("this check is for the GPU-less container")
    from nbdistributed_b200 import capi

    comm = C.c_void_p()
    rc = native_lib.nbd_comm_create(C.byref(comm), 0, 1, 0, 0)
    assert rc != 0, "creating a communicator without a GPU must fail, not fall back"
    with pytest.raises(capi.NbdError):
        capi.EmuGroup(2)
