"""Shared helpers for the parity tests: numpy <-> torch CUDA conversion (bf16 travels as
uint16 bit patterns on the numpy side) and comparison against the oracle."""
from __future__ import annotations

import numpy as np

from oracle import collectives_oracle as O

TORCH_NAME = {"f32": "float32", "bf16": "bfloat16", "f16": "float16", "f64": "float64",
              "i32": "int32", "i64": "int64", "u8": "uint8", "i8": "int8"}


def torch_dtype(name: str):
    import torch

    return getattr(torch, TORCH_NAME[name])


def to_cuda(a: np.ndarray, dtype: str, device="cuda:0"):
    import torch

    a = np.ascontiguousarray(a)
    if dtype == "bf16":
        return torch.from_numpy(a.view(np.int16).copy()).to(device).view(torch.bfloat16)
    return torch.from_numpy(a.copy()).to(device)


def to_numpy(t, dtype: str) -> np.ndarray:
    import torch

    t = t.detach().cpu().contiguous()
    if dtype == "bf16":
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


def assert_bits_equal(got: np.ndarray, want: np.ndarray, what: str = "") -> None:
    """Bit-exact comparison, except that any-NaN == any-NaN for floating dtypes."""
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, f"{what}: shape {got.shape} != {want.shape}"
    g = got.reshape(-1).view(np.uint8)
    w = want.reshape(-1).view(np.uint8)
    if np.array_equal(g, w):
        return
    if got.dtype.kind == "f":
        both_nan = np.isnan(got) & np.isnan(want)
        ok = both_nan | (got.view(f"u{got.dtype.itemsize}") == want.view(f"u{want.dtype.itemsize}"))
        if ok.all():
            return
        bad = np.flatnonzero(~ok.reshape(-1))
    else:
        bad = np.flatnonzero(got.reshape(-1) != want.reshape(-1))
    i = int(bad[0])
    raise AssertionError(f"{what}: {bad.size} of {got.size} elements differ; first at {i}: "
                         f"got {got.reshape(-1)[i]!r} want {want.reshape(-1)[i]!r}")


def assert_bf16_bits_equal(got_u16: np.ndarray, want_u16: np.ndarray, what: str = "") -> None:
    gf, wf = O.bf16_to_f32(got_u16), O.bf16_to_f32(want_u16)
    ok = (got_u16 == want_u16) | (np.isnan(gf) & np.isnan(wf))
    if not ok.all():
        bad = np.flatnonzero(~ok.reshape(-1))
        i = int(bad[0])
        raise AssertionError(f"{what}: {bad.size} of {got_u16.size} bf16 elements differ; first at {i}: "
                             f"got {gf.reshape(-1)[i]!r} want {wf.reshape(-1)[i]!r}")


def assert_same(got: np.ndarray, want: np.ndarray, dtype: str, what: str = "") -> None:
    if dtype == "bf16":
        assert_bf16_bits_equal(got, want, what)
    else:
        assert_bits_equal(got, want, what)
