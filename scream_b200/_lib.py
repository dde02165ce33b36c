"""ctypes binding of libscream_render.so (include/scream_render.h).

There is NO fallback: if the shared object is missing it is built with nvcc (scream_b200/csrc/build.py);
if that fails the import raises.  Nothing here touches oracle/.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from .csrc import build as _build

ABI_VERSION = 9

F_RAW = 1
F_DENSE = 2
F_ZRANGE_IN = 4
F_NOSTREAM = 8
F_THINSEED = 16
F_DUAL = 32
F_HALF = 64
F_DUAL_NOWIN1 = 128
NOALLOC_WS_BYTES = 65536

_lock = threading.Lock()
_lib = None

_P = C.c_void_p
_SIGNATURES = {
    "scream_abi_version": (C.c_int, []),
    "scream_last_error": (C.c_char_p, []),
    "scream_render_workspace_bytes": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int64, C.POINTER(C.c_size_t)]),
    "scream_render_forward": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, C.c_int, C.c_int, C.c_int,
                                        C.c_double, C.c_uint, _P, _P, _P, _P, C.c_size_t, _P]),
    "scream_render_forward_noalloc": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, C.c_int, C.c_int, C.c_int,
                                                C.c_double, C.c_uint, _P, _P, _P, _P]),
    "scream_render_backward": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, C.c_int, C.c_int, C.c_int,
                                         C.c_double, C.c_uint, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "scream_render_forward_pair": (C.c_int, [_P, C.c_int64, _P, C.c_int64, _P, _P, C.c_int, C.c_int, C.c_double, C.c_uint,
                                             _P, _P, _P, _P]),
    "scream_render_backward_pair": (C.c_int, [_P, C.c_int64, _P, C.c_int64, _P, _P, C.c_int, C.c_int, C.c_double, C.c_uint,
                                              _P, _P, _P, _P, _P, _P]),
    "scream_render_forward_sharded": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, C.c_int, C.c_int, C.c_int, C.c_double, C.c_uint,
                                                C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, C.c_size_t,
                                                C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), _P]),
    "scream_pack_keys": (C.c_int, [_P, _P, C.c_int64, C.c_int64, C.c_int, C.c_int, _P, _P]),
    "scream_unpack_keys": (C.c_int, [_P, C.c_int, C.c_int, _P, _P, _P, _P]),
    "scream_render_transform": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, _P, C.c_int, C.c_int,
                                          _P, _P, _P, _P, C.c_size_t, _P]),
    "scream_half_rotate": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, C.c_int, C.c_int, _P, _P, _P]),
    "scream_half_rotate_backward": (C.c_int, [_P, _P, C.c_int64, C.c_int64, _P, C.c_int, C.c_int, _P, _P]),
    "scream_nearest_target": (C.c_int, [_P, C.c_int64, _P, C.c_int64, C.c_float, _P, _P, _P, C.c_size_t, _P]),
    "scream_rigid_transform": (C.c_int, [_P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_int, C.c_int64, C.c_int64,
                                         _P, _P, C.c_size_t, _P]),
    "scream_disc_stem_forward": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "scream_disc_stem_backward": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P]),
    "scream_depth_to_rgb8": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "scream_render_forward_host": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_double, C.c_uint,
                                             _P, _P, _P, C.c_int]),
}
EXPORTS = tuple(_SIGNATURES)


class ScreamError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIB


def load():
    """Load (building first if needed) libscream_render.so; raises if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            path = _build.LIB if os.path.exists(_build.LIB) and not _can_rebuild() else _build.build()
            lib = C.CDLL(path)
            for name, (res, args) in _SIGNATURES.items():
                fn = getattr(lib, name)      # AttributeError if the symbol is missing: fail loudly
                fn.restype, fn.argtypes = res, args
            got = lib.scream_abi_version()
            if got != ABI_VERSION:
                raise ScreamError(f"libscream_render.so has ABI {got}, python expects {ABI_VERSION}; rebuild")
            _lib = lib
    return _lib


def _can_rebuild() -> bool:
    """True when nvcc is around (build container): then stale libraries are rebuilt from source."""
    try:
        _build.find_nvcc()
        return True
    except RuntimeError:
        return False


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().scream_last_error().decode("utf-8", "replace")
        raise ScreamError(f"{what} failed (code {rc}): {msg}")


def workspace_bytes(B: int, V: int, w: int, total_pts: int) -> int:
    out = C.c_size_t(0)
    check(load().scream_render_workspace_bytes(B, V, w, total_pts, C.byref(out)), "scream_render_workspace_bytes")
    return int(out.value)
