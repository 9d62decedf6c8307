"""ctypes binding to the in-tree native libraries.

``lib()`` returns ``libharp_b200.so`` (CUDA kernels, C ABI declared in
``include/harp_b200.h``); ``host()`` returns ``libharp_host.so`` (C++ host
runtime).  On a GPU box a missing CUDA library is a hard error — ops never fall
back silently to eager PyTorch when a GPU is present.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_int64, c_longlong, c_size_t, c_void_p
from pathlib import Path

_PKG = Path(__file__).resolve().parent
_CUDA_LIB = None
_HOST_LIB = None


class NativeError(RuntimeError):
    pass


def cuda_lib_path() -> Path:
    # HB_CUDA_LIB points at an alternative build of the same library (A/B runs of compile-time knobs)
    alt = os.environ.get("HB_CUDA_LIB")
    return Path(alt) if alt else _PKG / "libharp_b200.so"


def host_lib_path() -> Path:
    return _PKG / "libharp_host.so"


def _declare_cuda(lib: ctypes.CDLL) -> None:
    P = c_void_p
    lib.hb_last_error.restype = c_char_p
    lib.hb_last_error.argtypes = []
    lib.hb_version.restype = c_int
    lib.hb_device_sm_count.restype = c_int
    lib.hb_l2_flush.argtypes = [P, c_size_t, P]
    lib.hb_kmeans_workspace_bytes.restype = c_size_t
    lib.hb_kmeans_workspace_bytes.argtypes = [c_int, c_int]
    lib.hb_kmeans_min_workspace_bytes.restype = c_size_t
    lib.hb_kmeans_min_workspace_bytes.argtypes = [c_int, c_int]
    lib.hb_kmeans_simt_workspace_bytes.restype = c_size_t
    lib.hb_kmeans_simt_workspace_bytes.argtypes = [c_longlong, c_int, c_int]
    lib.hb_kmeans_tc_supported.argtypes = [c_int, c_int]
    lib.hb_kmeans_step.argtypes = [P, c_longlong, c_int, P, c_int, P, P, P, P, P, c_size_t, c_int, P]
    lib.hb_kmeans_finalize.argtypes = [P, P, c_int, c_int, P, c_int, P]
    # optional symbols are declared lazily by the modules that use them


def lib() -> ctypes.CDLL:
    """The CUDA library; raises NativeError if it has not been built."""
    global _CUDA_LIB
    if _CUDA_LIB is None:
        path = cuda_lib_path()
        if not path.exists():
            raise NativeError(
                f"{path} is missing: run `python -m harp_b200.build` (or __graft_entry__.build()) first"
            )
        _CUDA_LIB = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        _declare_cuda(_CUDA_LIB)
    return _CUDA_LIB


def host() -> ctypes.CDLL:
    global _HOST_LIB
    if _HOST_LIB is None:
        path = host_lib_path()
        if not path.exists():
            raise NativeError(f"{path} is missing: run `python -m harp_b200.build` first")
        _HOST_LIB = ctypes.CDLL(str(path))
    return _HOST_LIB


def have_cuda_lib() -> bool:
    return cuda_lib_path().exists()


def have_host_lib() -> bool:
    return host_lib_path().exists()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().hb_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def ptr(t) -> c_void_p:
    """Device/host pointer of a torch tensor (None → NULL)."""
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def stream_ptr(stream=None) -> c_void_p:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return c_void_p(s.cuda_stream)


# how many of OUR kernels have been launched (bench.py reports it as gpu_launches)
_LAUNCHES = 0


def count_launch(n: int = 1) -> None:
    global _LAUNCHES
    _LAUNCHES += n


def launches() -> int:
    return _LAUNCHES


def reset_launches() -> None:
    global _LAUNCHES
    _LAUNCHES = 0
