"""ctypes loaders for the two native libraries.

`cuda_lib()`  -> libpsb200_cuda.so  (sm_100a kernels, C ABI declared in include/psb200.h)
`host_lib()`  -> libpsb200_host.so  (C++ host runtime: van, executor, parsers, filters, ...)

On a machine with a GPU a missing CUDA library is a hard error: there is no silent PyTorch
fallback for the kernels.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

_LIBDIR = Path(__file__).resolve().parent / "lib"
_cuda = None
_host = None


class NativeError(RuntimeError):
    pass


def _load(name: str) -> ctypes.CDLL:
    path = _LIBDIR / name
    if not path.exists():
        # build lazily when a compiler is around (dev box); never on import
        if os.environ.get("PSB200_NO_AUTOBUILD") != "1":
            from . import build as _build
            if "cuda" in name:
                _build.build_cuda()
            else:
                _build.build_host()
    if not path.exists():
        raise NativeError(
            f"{path} is missing: run `python -m parameter_server_b200.build` "
            "(the kernels have no PyTorch fallback)")
    # RTLD_LOCAL: the libraries carry their own (static) C++ runtime, which must not interpose
    # with the libstdc++ that torch brings into the process
    return ctypes.CDLL(str(path), mode=ctypes.RTLD_LOCAL)


def cuda_lib() -> ctypes.CDLL:
    global _cuda
    if _cuda is None:
        lib = _load("libpsb200_cuda.so")
        lib.psb_last_error.restype = ctypes.c_char_p
        lib.psb_darlin_evaluate_scratch_bytes.restype = ctypes.c_int64
        _cuda = lib
    return _cuda


def host_lib() -> ctypes.CDLL:
    global _host
    if _host is None:
        lib = _load("libpsb200_host.so")
        lib.psh_last_error.restype = ctypes.c_char_p
        _host = lib
    return _host


def check(rc: int) -> None:
    """Raise NativeError with the library's message when a C-ABI call failed."""
    if rc != 0:
        msg = cuda_lib().psb_last_error()
        raise NativeError(f"psb200 CUDA call failed (rc={rc}): {msg.decode() if msg else '?'}")


def hcheck(rc: int) -> None:
    if rc != 0:
        msg = host_lib().psh_last_error()
        raise NativeError(f"psb200 host call failed (rc={rc}): {msg.decode() if msg else '?'}")


def ptr(t) -> ctypes.c_void_p:
    """Device/host pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def current_stream() -> ctypes.c_void_p:
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
