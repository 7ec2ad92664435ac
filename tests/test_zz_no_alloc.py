"""The allocation-free statistics entry point (``tdl_tensor_stats_noalloc``), exercised the way its
intended users reach it: straight through the C ABI with ctypes, no Python wrapper in between.

Those users are host programs that link ``libtdl_kernels.so`` without PyTorch (no caching allocator to
borrow scratch from) or jobs that have committed all device memory up front and refuse driver
allocations afterwards.  The test checks that the built-in scratch gives the same 12 numbers as the
regular entry point fed with caller-owned scratch, on two calls in a row (the second call sees the
first call's leftovers in the shared array), and that an oversize request is refused.

The file name sorts last on purpose: if this fails under ``-x``, everything else has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1 << 30          # TDL_STATS_NOALLOC_MAX_ELEMS in include/tdl_kernels.h


def _load():
    from trustworthy_dl.ops import _lib
    lib = _lib.lib()
    lib.tdl_tensor_stats_noalloc.restype = ctypes.c_int
    lib.tdl_tensor_stats.restype = ctypes.c_int
    lib.tdl_stats_scratch_bytes.restype = ctypes.c_size_t
    lib.tdl_last_error.restype = ctypes.c_char_p
    return lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _reference_call(lib, x, code):
    """Existing entry point with scratch we allocate ourselves."""
    nbytes = lib.tdl_stats_scratch_bytes(ctypes.c_int64(x.numel()))
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device=x.device)
    out = torch.empty(12, dtype=torch.float64, device=x.device)
    rc = lib.tdl_tensor_stats(ctypes.c_void_p(x.data_ptr()), ctypes.c_int(code), ctypes.c_int64(x.numel()),
                              ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(scratch.data_ptr()),
                              ctypes.c_size_t(nbytes), _stream())
    assert rc == 0, lib.tdl_last_error()
    return out


def _noalloc_call(lib, x, code, out):
    return lib.tdl_tensor_stats_noalloc(ctypes.c_void_p(x.data_ptr()), ctypes.c_int(code), ctypes.c_int64(x.numel()),
                                        ctypes.c_void_p(out.data_ptr()), _stream())


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    g = torch.Generator().manual_seed(2024)
    # two different tensors (dtype, size, distribution) so leftovers from call 1 would corrupt call 2
    inputs = [((torch.randn(1_000_003, generator=g) * 2.0 + 0.5).to(dev), 0),
              ((torch.randn(777_216, generator=g).abs() * 0.3 - 1.0).bfloat16().to(dev), 1)]
    outs = [torch.empty(12, dtype=torch.float64, device=dev) for _ in inputs]   # outputs belong to the caller
    for (x, code), out in zip(inputs, outs):
        rc = _noalloc_call(lib, x, code, out)
        assert rc == 0, lib.tdl_last_error()
    torch.cuda.synchronize()
    for (x, code), out in zip(inputs, outs):
        ref = _reference_call(lib, x, code)
        torch.cuda.synchronize()
        for i in range(12):
            assert out[i].item() == pytest.approx(ref[i].item(), rel=1e-5, abs=1e-5), i
    # an oracle that shares no code with either entry point: the plain PyTorch/numpy definition of the 12 statistics
    from trustworthy_dl.ops.reference_impl import tensor_statistics as oracle
    for (x, code), out in zip(inputs, outs):
        want = oracle(x.float().cpu())
        tol = 1e-4 if code == 0 else 2e-3
        for i in range(12):
            assert out[i].item() == pytest.approx(want[i].item(), rel=tol, abs=tol), i
    # and once more on the first tensor: a third use of the same array still agrees
    again = torch.empty(12, dtype=torch.float64, device=dev)
    assert _noalloc_call(lib, inputs[0][0], inputs[0][1], again) == 0
    torch.cuda.synchronize()
    assert torch.allclose(again, outs[0], rtol=1e-6, atol=1e-9)


def test_noalloc_rejects_oversize_and_bad_arguments():
    lib = _load()
    dev = torch.device("cuda", 0)
    x = torch.zeros(16, device=dev)
    out = torch.empty(12, dtype=torch.float64, device=dev)
    rc = lib.tdl_tensor_stats_noalloc(ctypes.c_void_p(x.data_ptr()), ctypes.c_int(0), ctypes.c_int64(MAX_ELEMS + 1),
                                      ctypes.c_void_p(out.data_ptr()), _stream())
    assert rc != 0
    msg = lib.tdl_last_error().decode()
    assert "exceed" in msg and str(MAX_ELEMS) in msg
    rc = lib.tdl_tensor_stats_noalloc(ctypes.c_void_p(x.data_ptr()), ctypes.c_int(0), ctypes.c_int64(0),
                                      ctypes.c_void_p(out.data_ptr()), _stream())
    assert rc != 0
    torch.cuda.synchronize()
