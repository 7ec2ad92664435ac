"""ctypes binding to the in-tree sm_100a kernel library (``trustworthy_dl/_C/libtdl_kernels.so``).

The library is a plain C-ABI shared object built by ``nvcc`` from ``csrc/*.cu`` (see
``__graft_entry__.build`` / ``setup.py``); declarations live in ``include/tdl_kernels.h``.  Python
passes raw device pointers (``tensor.data_ptr()``) and the current CUDA stream, so the kernels can
be captured into CUDA graphs and never allocate.

On a machine with a GPU a missing library is a hard error: ops must not silently fall back to
eager PyTorch (the fallbacks in ``reference_impl`` are for CPU tensors only).
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

import torch

_LIB_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_C")
LIB_PATH = os.path.join(_LIB_DIR, "libtdl_kernels.so")
_lib: Optional[ctypes.CDLL] = None
_lock = threading.Lock()
launch_count = 0          # number of kernel-library calls issued (bench.py reports it)

DTYPE_CODES = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


class KernelLibraryMissing(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise KernelLibraryMissing(
                        f"{LIB_PATH} not found - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(nvcc -gencode arch=compute_100a,code=sm_100a). CUDA tensors never fall back to eager ops.")
                _lib = ctypes.CDLL(LIB_PATH)
                _lib.tdl_last_error.restype = ctypes.c_char_p
                _lib.tdl_stats_scratch_bytes.restype = ctypes.c_size_t
                _lib.tdl_exchange_pad_bytes.restype = ctypes.c_size_t
    return _lib


def stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def i64(v) -> ctypes.c_int64:
    return ctypes.c_int64(int(v))


def i32(v) -> ctypes.c_int:
    return ctypes.c_int(int(v))


def f32(v) -> ctypes.c_float:
    return ctypes.c_float(float(v))


def f64(v) -> ctypes.c_double:
    return ctypes.c_double(float(v))


def dtype_code(t: torch.Tensor) -> int:
    try:
        return DTYPE_CODES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype} for the sm_100a kernels") from None


def call(name: str, *args) -> None:
    """Invoke ``int name(...)``; non-zero return raises with the library's error string."""
    global launch_count
    fn = getattr(lib(), name)
    rc = fn(*args)
    launch_count += 1
    if rc != 0:
        msg = lib().tdl_last_error()
        raise RuntimeError(f"{name} failed (code {rc}): {msg.decode() if msg else 'unknown'}")


_scratch = {}


def scratch(device: torch.device, nbytes: int, tag: str = "default") -> torch.Tensor:
    """Per-(device, tag) reusable scratch buffer (uint8), grown geometrically, zero-initialised."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), tag)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf
