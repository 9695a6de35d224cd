"""The allocation-free sum-of-squares entry point, exercised through the C ABI only.

``detb200_sumsq_noalloc`` exists for users of ``libdetb200.so`` that cannot supply scratch
memory: a trainer whose HBM is wholly owned by a pool/caching allocator, a bare C program with no
allocator, or a hot path that must not take ``cudaMalloc``'s device-wide synchronisation.  Its
scratch is a ``__device__`` array inside the library, shared by all calls on a GPU and never
cleaned between them, so the things worth checking are: it agrees with the ordinary entry point
(which gets fresh caller-provided scratch), it still agrees on a second call that finds the first
call's leftovers in the array, and it refuses inputs above its stated limit.

This file is named so that pytest collects it last: a failure here cannot hide other results in a
run that stops at the first failure.  It must pass alone in a fresh process and after every other
test (the library may or may not be loaded already).
"""
import ctypes
import pathlib

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = pathlib.Path(__file__).resolve().parent.parent / "determined_b200" / "_native" / "libdetb200.so"


def _load() -> ctypes.CDLL:
    lib = ctypes.CDLL(str(LIB))
    p, i64, i = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    lib.detb200_last_error.restype = ctypes.c_char_p
    lib.detb200_sumsq_scratch_bytes.restype = i64
    lib.detb200_sumsq.restype = i
    lib.detb200_sumsq.argtypes = [p, i64, p, p, p, i, p]
    lib.detb200_sumsq_noalloc_max_elems.restype = i64
    lib.detb200_sumsq_noalloc.restype = i
    lib.detb200_sumsq_noalloc.argtypes = [p, i64, p, p, i, p]
    return lib


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _reference(lib: ctypes.CDLL, x: torch.Tensor, accumulate_into: float = None):
    """The existing entry point, with scratch the caller allocated (torch) and zeroed."""
    scratch = torch.zeros(int(lib.detb200_sumsq_scratch_bytes()), dtype=torch.uint8, device=x.device)
    out = torch.zeros(1, dtype=torch.float32, device=x.device)
    flag = torch.zeros(1, dtype=torch.int32, device=x.device)
    if accumulate_into is not None:
        out.fill_(accumulate_into)
    rc = lib.detb200_sumsq(x.data_ptr(), x.numel(), scratch.data_ptr(), out.data_ptr(), flag.data_ptr(),
                           int(accumulate_into is not None), _stream())
    assert rc == 0, lib.detb200_last_error()
    torch.cuda.synchronize()
    return float(out.item()), int(flag.item())


def _noalloc(lib: ctypes.CDLL, x: torch.Tensor, accumulate_into: float = None):
    # inputs and outputs still belong to the caller; only the scratch is the library's
    out = torch.zeros(1, dtype=torch.float32, device=x.device)
    flag = torch.zeros(1, dtype=torch.int32, device=x.device)
    if accumulate_into is not None:
        out.fill_(accumulate_into)
    rc = lib.detb200_sumsq_noalloc(x.data_ptr(), x.numel(), out.data_ptr(), flag.data_ptr(),
                                   int(accumulate_into is not None), _stream())
    assert rc == 0, lib.detb200_last_error()
    torch.cuda.synchronize()
    return float(out.item()), int(flag.item())


def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row():
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(11)
    # first call: big input -> the widest grid, every partial slot of the built-in array is written
    x1 = torch.randn((1 << 23) + 5, device=dev)
    want1, _ = _reference(lib, x1)
    got1, bad1 = _noalloc(lib, x1)
    torch.testing.assert_close(torch.tensor(got1), torch.tensor(want1), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(torch.tensor(got1, dtype=torch.float64), x1.double().pow(2).sum().cpu(), rtol=1e-5, atol=1e-6)
    assert bad1 == 0
    # second call right after: small input -> a narrow grid that must not pick up the first call's
    # leftover partials or a stale ticket from the shared array
    x2 = torch.randn(4099, device=dev) * 3.0
    want2, _ = _reference(lib, x2)
    got2, bad2 = _noalloc(lib, x2)
    torch.testing.assert_close(torch.tensor(got2), torch.tensor(want2), rtol=1e-5, atol=1e-6)
    assert bad2 == 0
    # accumulate mode and the non-finite flag behave like the existing entry point too
    want3, _ = _reference(lib, x2, accumulate_into=7.5)
    got3, _ = _noalloc(lib, x2, accumulate_into=7.5)
    torch.testing.assert_close(torch.tensor(got3), torch.tensor(want3), rtol=1e-5, atol=1e-6)
    x2[17] = float("inf")
    _, want_flag = _reference(lib, x2)
    _, got_flag = _noalloc(lib, x2)
    assert want_flag == got_flag == 1


def test_noalloc_refuses_sizes_beyond_its_limit():
    lib = _load()
    torch.cuda.set_device(0)
    x = torch.zeros(16, device="cuda")
    out = torch.zeros(1, device="cuda")
    limit = int(lib.detb200_sumsq_noalloc_max_elems())
    assert limit == 1 << 30
    # rejected on the host before anything is launched, so a tiny buffer is safe to pass
    rc = lib.detb200_sumsq_noalloc(x.data_ptr(), limit + 1, out.data_ptr(), None, 0, _stream())
    assert rc == -2, rc
    msg = lib.detb200_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs a second GPU")
def test_noalloc_keeps_one_scratch_address_per_gpu():
    """The library caches the array's device address per GPU.  Use it on cuda:0, then on cuda:1, then on cuda:0 again,
    and compare each with the ordinary entry point on the same device; finally hand it a stream of the other device,
    which it must refuse on the host instead of launching with the wrong GPU's scratch."""
    lib = _load()
    torch.manual_seed(5)
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        torch.cuda.set_device(dev)
        x = torch.randn((1 << 20) + 3, device=dev) * (d + 1)
        want, _ = _reference(lib, x)
        got, bad = _noalloc(lib, x)
        torch.testing.assert_close(torch.tensor(got), torch.tensor(want), rtol=1e-5, atol=1e-6)
        assert bad == 0
    torch.cuda.set_device(0)
    foreign = torch.cuda.Stream(device=1)
    x = torch.zeros(16, device="cuda:0")
    out = torch.zeros(1, device="cuda:0")
    rc = lib.detb200_sumsq_noalloc(x.data_ptr(), 16, out.data_ptr(), None, 0, foreign.cuda_stream)
    assert rc == -1, rc
    assert "current" in lib.detb200_last_error().decode()
