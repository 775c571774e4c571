"""ctypes binding of include/psla_frontend.h (the C ABI of libpsla_frontend.so).

There is deliberately no fallback: if the shared library is missing or cannot be loaded the import
of the product path fails with an error that says how to build it.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

from .build import LIB

PSLA_OK = 0
STATUS_NAMES = {0: "PSLA_OK", -1: "PSLA_E_INVALID", -2: "PSLA_E_UNSUPPORTED", -3: "PSLA_E_SHORT_CLIP",
                -4: "PSLA_E_MASK", -5: "PSLA_E_WORKSPACE", -6: "PSLA_E_CUDA", -7: "PSLA_E_NOMEM", -8: "PSLA_E_STATIC_LIMIT"}

# every symbol include/psla_frontend.h declares: name -> (restype, argtypes).  The host metadata arrays of the forward entry
# points (int64 / int32 / double pointers in the header) are bound as void*: the caller passes the numpy buffers' addresses as
# plain integers, which costs 1 us per array instead of the 2.7 us of `ndarray.ctypes.data_as(POINTER(...))` -- six arrays
# per call next to a 48 us kernel.
SYMBOLS = {
    "psla_frontend_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_int, c_float, c_float, c_float, c_float, c_float]),
    "psla_frontend_destroy": (None, [c_void_p]),
    "psla_frontend_set_tables": (c_int, [c_void_p, c_void_p, c_void_p]),
    "psla_frontend_get_tables": (c_int, [c_void_p, c_void_p, c_void_p]),
    "psla_frontend_workspace_bytes": (c_size_t, [c_void_p, c_int]),
    "psla_frontend_forward": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_int,
                                      c_void_p, c_void_p, c_size_t, c_void_p]),
    "psla_frontend_forward_host": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_int,
                                           c_void_p, c_void_p]),
    "psla_frontend_device_out": (c_void_p, [c_void_p]),
    "psla_frontend_forward_static": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                             c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_int, c_void_p, c_void_p]),
    "psla_frontend_forward_pcm16": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                            c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_int,
                                            c_void_p, c_void_p, c_size_t, c_void_p]),
    "psla_frontend_forward_host_pcm16": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_void_p,
                                                 c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float,
                                                 c_int, c_void_p, c_void_p]),
    "psla_frontend_noise_roll": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, POINTER(c_float), POINTER(c_int32), c_int, c_int,
                                         ctypes.c_uint64, c_void_p]),
    "psla_frontend_num_frames": (c_int, [c_void_p, c_int64]),
    "psla_frontend_launch_count": (c_int64, [c_void_p]),
    "psla_frontend_get_geometry": (c_int, [c_void_p, POINTER(c_int), POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "psla_frontend_set_generic_path": (c_int, [c_void_p, c_int]),
    "psla_frontend_set_fast_log": (c_int, [c_void_p, c_int]),
    "psla_frontend_set_exact_clip_mean": (c_int, [c_void_p, c_int]),
    "psla_frontend_set_output_layout": (c_int, [c_void_p, c_int]),
    "psla_frontend_set_timing": (c_int, [c_void_p, c_int]),
    "psla_frontend_get_timing": (c_int, [c_void_p, POINTER(c_double), POINTER(c_double), POINTER(c_int64)]),
    "psla_frontend_set_trace": (c_int, [c_void_p, c_void_p, c_size_t]),
    "psla_frontend_trace_records": (c_int, [c_void_p]),
    "psla_frontend_trace_words": (c_int, []),
    "psla_frontend_set_cost_model": (c_int, [c_void_p, c_int, c_int, c_int]),
    "psla_frontend_set_meta_stream": (c_int, [c_void_p, c_int]),
    "psla_frontend_set_launch_overlap": (c_int, [c_void_p, c_int]),
    "psla_frontend_set_param_meta": (c_int, [c_void_p, c_int]),
    "psla_strerror": (c_char_p, [c_int]),
    "psla_last_cuda_error": (c_int, [c_void_p]),
    "psla_version": (c_char_p, []),
}

_lib = None


class PslaError(RuntimeError):
    def __init__(self, status: int, where: str, cuda: int = 0):
        self.status = status
        msg = f"{where}: {STATUS_NAMES.get(status, status)} ({strerror(status)})"
        if cuda:
            msg += f" [cudaError {cuda}]"
        super().__init__(msg)


class ShortClipError(PslaError, AssertionError):
    """The reference raises AssertionError (torchaudio kaldi.py:142) for clips shorter than one frame."""


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            raise ImportError(
                f"{LIB} is missing: the CUDA frontend has not been built (python -m psla_b200.build, or "
                "__graft_entry__.build()).  psla_b200 has no CPU fallback.")
        lib = ctypes.CDLL(LIB)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def strerror(status: int) -> str:
    try:
        return load().psla_strerror(int(status)).decode()
    except Exception:
        return "?"


def check(status: int, where: str, handle=None):
    if status == PSLA_OK:
        return
    cuda = load().psla_last_cuda_error(handle) if handle else 0
    if status == -3:
        raise ShortClipError(status, where, cuda)
    raise PslaError(status, where, cuda)
