"""ctypes wrapper of oracle/paged_oracle.c (TEST INFRASTRUCTURE ONLY; see that file's header)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpaged_oracle.so")
_CODES = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
_lib = None


def build() -> str:
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.oracle_paged_decode.restype = C.c_int
        _lib.oracle_paged_decode.argtypes = [C.c_void_p] * 6 + [C.c_int64, C.c_int64, C.c_int, C.c_int,
                                                               C.c_int, C.c_int, C.c_double, C.c_int]
        _lib.oracle_store_kvcache.restype = None
        _lib.oracle_store_kvcache.argtypes = [C.c_void_p] * 5 + [C.c_int64, C.c_int64]
    return _lib


def paged_decode(q, k_cache, v_cache, block_tables, context_lens, scale, num_heads, num_kv_heads,
                 head_dim, block_size) -> torch.Tensor:
    """Exact (double-accumulated) decode attention on CPU tensors; returns float32 (B, H, D)."""
    q, k_cache, v_cache = (t.detach().cpu().contiguous() for t in (q, k_cache, v_cache))
    bt = block_tables.detach().cpu().to(torch.int32).contiguous()
    cl = context_lens.detach().cpu().to(torch.int64).contiguous()
    out = torch.empty(q.shape, dtype=torch.float32)
    rc = lib().oracle_paged_decode(out.data_ptr(), q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                   bt.data_ptr(), cl.data_ptr(), q.shape[0], bt.shape[1], num_heads,
                                   num_kv_heads, head_dim, block_size, float(scale), _CODES[q.dtype])
    if rc:
        raise MemoryError("oracle_paged_decode: allocation failed")
    return out


def store_kvcache(key, value, k_cache, v_cache, slot_mapping) -> None:
    """Byte-exact scatter into CPU caches (in place)."""
    assert not k_cache.is_cuda and k_cache.is_contiguous() and v_cache.is_contiguous()
    key, value = key.detach().cpu().contiguous(), value.detach().cpu().contiguous()
    slots = slot_mapping.detach().cpu().to(torch.int64).contiguous()
    row_bytes = key.shape[1] * key.shape[2] * key.element_size()
    lib().oracle_store_kvcache(key.data_ptr(), value.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                               slots.data_ptr(), key.shape[0], row_bytes)
