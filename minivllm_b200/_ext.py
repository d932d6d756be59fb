"""ctypes binding of libmvl_b200.so (the C ABI declared in include/mvl_abi.h).

The library is built in-tree (``minivllm_b200/csrc/libmvl_b200.so``) by ``__graft_entry__.build()``
or ``make -C minivllm_b200/csrc``.  There is no fallback: if the library is missing, ``lib()``
raises, and every operator in ``minivllm_b200.layers.attention`` goes through ``lib()``.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmvl_b200.so")
ABI_VERSION = 1

# mvl_dtype
F32, F16, BF16 = 0, 1, 2
# flags of mvl_paged_decode
DECODE_SKIP_COMBINE, DECODE_ONLY_COMBINE, DECODE_FUSED_COMBINE, DECODE_LOAD_TMA, DECODE_CUDA_CORES = 1, 2, 4, 8, 16
DECODE_SHARE_PREFIX = 32
DECODE_SHARE_NARROW = 64
DECODE_SPLIT_GRID, DECODE_STREAM, DECODE_OVERLAP_PREV, DECODE_GROUP_MMA = 128, 256, 512, 1024
# flags of mvl_flash_prefill
PREFILL_MMA_SYNC, PREFILL_ONE_TILE, PREFILL_PAIR = 1, 2, 4

_vp, _i, _i64, _sz, _f, _u = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float, C.c_uint

# name -> (restype, argtypes); one entry per declaration in include/mvl_abi.h
PROTOTYPES = {
    "mvl_abi_version": (_i, []),
    "mvl_strerror": (C.c_char_p, [_i]),
    "mvl_last_cuda_error": (C.c_char_p, []),
    "mvl_store_kvcache": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i64, _i, _i, _i64, _i64, _i64, _i, _vp]),
    "mvl_paged_decode_default_chunk": (_i, [_i64, _i, _i, _i, _i64, _i, _i]),
    "mvl_paged_decode_workspace_bytes": (_sz, [_i64, _i, _i, _i, _i64, _i, _i, _i]),
    "mvl_paged_decode": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i64, _i64, _i64, _i, _i, _i, _i,
                              _i64, _f, _i, _vp, _sz, _i, _u, _vp]),
    "mvl_paged_decode_peers": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _i64, _i64, _i64, _i,
                                    _i, _i, _i, _i64, _f, _i, _vp, _sz, _i, _u, _vp]),
    "mvl_flash_prefill": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp, _i, _i, _i, _i, _i, _f, _i, _u,
                               _vp]),
    "mvl_flash_prefill_paged": (_i, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _i, _i, _i, _i, _i, _i,
                                     _i, _i64, _f, _i, _u, _vp]),
    "mvl_qk_norm_rope_store": (_i, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _f, _vp, _i64, _vp, _i,
                                    _vp, _vp, _vp, _i, _i64, _i64, _i, _i, _i, _i, _vp]),
    "mvl_decode_step": (_i, [_vp] * 9 + [_i, _i64, _i64, _i64, _i64, _i64, _i, _i, _i, _i, _i64, _f, _i,
                              _vp, _sz, _i, _u, _vp]),
    "mvl_decode_step_peers": (_i, [_vp, _i, _i, _i] + [_vp] * 8 + [_i, _i64, _i64, _i64, _i64, _i64, _i,
                                   _i, _i, _i, _i64, _f, _i, _vp, _sz, _i, _u, _vp]),
    "mvl_decode_step_peers_sync": (_i, [_vp, _vp, _i, _i, _i, _i] + [_vp] * 8 + [_i, _i64, _i64, _i64, _i64, _i64,
                                        _i, _i, _i, _i, _i64, _f, _i, _vp, _sz, _i, _u, _vp]),
    "mvl_allgather_heads": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _i, _vp, _vp]),
    "mvl_debug_set_times": (None, [_vp]),
    "mvl_peer_wait": (_i, [_vp, _i, _i64, _vp]),
    "mvl_decode_step_host": (_i, [_vp] * 16 + [_i, _i64, _i64, _i, _i, _i, _i, _i64, _f, _i, _vp,
                                               _sz, _i, _u, _vp]),
    "mvl_decode_step_host_workspace_bytes": (_sz, [_i64, _i, _i, _i, _i64, _i, _i, _i, _i]),
    "mvl_decode_step_host_overlapped": (_i, [_vp] * 16 + [_i, _i64, _i64, _i, _i, _i, _i, _i64, _f, _i, _vp,
                                                          _sz, _i, _u, _i, _vp]),
    "mvl_paged_decode_static": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i64, _i64, _i64, _i, _i, _i,
                                     _i, _i64, _f, _i, _i, _vp]),
    "mvl_static_scratch_bytes": (_sz, []),
    "mvl_static_last_error": (C.c_char_p, []),
}

_lib = None
_lock = threading.Lock()


class MvlError(RuntimeError):
    """A libmvl_b200 call returned a negative mvl_status."""

    def __init__(self, status: int, message: str):
        super().__init__(f"libmvl_b200: status {status}: {message}")
        self.status = status


def lib() -> C.CDLL:
    """Load (once) and return the library with prototypes set.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise ImportError(
                    f"{LIB_PATH} is missing: build it with `make -C minivllm_b200/csrc` or "
                    "`python -c 'import __graft_entry__ as g; g.build()'`. "
                    "minivllm_b200 has no CPU or Triton fallback."
                )
            handle = C.CDLL(LIB_PATH)
            for name, (res, args) in PROTOTYPES.items():
                fn = getattr(handle, name)  # AttributeError if the ABI lost a symbol
                fn.restype = res
                fn.argtypes = args
            got = handle.mvl_abi_version()
            if got != ABI_VERSION:
                raise ImportError(f"libmvl_b200 ABI version {got}, binding expects {ABI_VERSION}")
            _lib = handle
    return _lib


def check(status: int) -> None:
    """Raise ValueError (bad arguments) or MvlError (CUDA / workspace) for a negative status."""
    if status >= 0:
        return
    handle = lib()
    msg = handle.mvl_strerror(status).decode()
    if status == -6:
        msg += ": " + handle.mvl_last_cuda_error().decode()
        raise MvlError(status, msg)
    if status in (-1, -2, -3, -4, -7):
        raise ValueError(f"libmvl_b200: status {status}: {msg}")
    raise MvlError(status, msg)
