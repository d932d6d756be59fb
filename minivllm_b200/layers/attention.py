"""Host-side mirror of the reference attention operator interface, backed by libmvl_b200.

Same names, positional order and return conventions as reference
src/myvllm/layers/attention.py: ``store_kvcache`` (:67-108), ``paged_attention_decode``
(:418-473) and the ``Attention`` module (:476-557).  Where the reference launches Triton
programs, these functions pass raw device pointers to the C ABI (include/mvl_abi.h) on torch's
current stream.  Nothing here computes on the CPU and nothing falls back: non-CUDA tensors and a
missing library are errors.

``flash_attention_prefill`` (:212-280) is here too for 16-bit inputs, so
``Attention.forward`` covers both branches of the reference.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import _ext
from ..utils.context import get_context

_DTYPES = {torch.float32: _ext.F32, torch.float16: _ext.F16, torch.bfloat16: _ext.BF16}

# ABI flags used when a caller does not pass any: chunk combine fused into the split kernel.
DEFAULT_DECODE_FLAGS = _ext.DECODE_FUSED_COMBINE

# One scratch tensor per (device, stream) for split-KV partials, grown on demand and reused, so
# steady-state calls allocate nothing (a CUDA-graph capture sees a fixed pointer).
_workspaces: dict = {}


def _dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise ValueError(f"unsupported dtype {t.dtype}; expected float32, float16 or bfloat16") from None


def _need_cuda(**tensors: torch.Tensor) -> torch.device:
    dev = None
    for name, t in tensors.items():
        if not isinstance(t, torch.Tensor):
            raise TypeError(f"{name} must be a torch.Tensor")
        if not t.is_cuda:
            raise ValueError(f"{name} must be a CUDA tensor (minivllm_b200 has no CPU path)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise ValueError(f"{name} is on {t.device}, expected {dev}")
    return dev


def _index_flag(name: str, t: torch.Tensor) -> int:
    if t.dtype == torch.int64:
        return 1
    if t.dtype == torch.int32:
        return 0
    raise ValueError(f"{name} must be int32 or int64, got {t.dtype}")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream_ptr(dev: torch.device) -> int:
    """cudaStream_t of torch's current stream on ``dev``.  The raw getter avoids building a
    torch.cuda.Stream object per call (~3 us each, twice per operator call, out of ~23 us of host time
    for a decode step at a tiny shape)."""
    if _raw_stream is not None:
        return _raw_stream(dev.index if dev.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(dev).cuda_stream


class _on_device:
    """``with torch.cuda.device(dev)`` only when ``dev`` is not already current (saves ~10 us/call)."""

    __slots__ = ("dev", "ctx")

    def __init__(self, dev: torch.device):
        self.dev, self.ctx = dev, None

    def __enter__(self):
        if torch.cuda.current_device() != self.dev.index:
            self.ctx = torch.cuda.device(self.dev)
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)


# (batch, H, KVH, D, max_blocks, BS, dtype code, chunk, device) -> workspace bytes; the plan is a
# pure function of these, so it is asked of the library once per shape.
_plan_cache: dict = {}


def workspace_for(dev: torch.device, nbytes: int) -> Optional[torch.Tensor]:
    """Scratch for the split-KV partials on torch's current stream of ``dev`` (None if nbytes == 0)."""
    if nbytes == 0:
        return None
    key = (dev.index, _stream_ptr(dev))
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        # zero-filled: its first 256 KiB are the arrival counters of the fused combine, which every
        # call leaves at zero (include/mvl_abi.h, MVL_DECODE_FUSED_COMBINE)
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    return ws


def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
                  v_cache: torch.Tensor, slot_mapping: torch.Tensor, block_size: int) -> None:
    """Scatter ``key``/``value`` rows into the paged caches at ``slot_mapping`` (-1 = skip).

    key, value: (num_tokens, num_kv_heads, head_dim); k_cache, v_cache:
    (num_blocks, block_size, num_kv_heads, head_dim) contiguous; slot_mapping: (num_tokens,).
    Mutates the caches in place, returns None (reference attention.py:67-108).  Views whose last
    two dims are dense (e.g. the k/v slices of a fused qkv projection) are consumed without a copy.
    """
    dev = _need_cuda(key=key, value=value, k_cache=k_cache, v_cache=v_cache, slot_mapping=slot_mapping)
    if key.dim() != 3 or value.shape != key.shape:
        raise ValueError("key and value must both be (num_tokens, num_kv_heads, head_dim)")
    num_tokens, num_kv_heads, head_dim = key.shape
    assert k_cache.shape == v_cache.shape, "K and V cache shapes must match"
    assert slot_mapping.numel() == num_tokens, "Slot mapping size must match number of tokens"
    if k_cache.dim() != 4 or k_cache.shape[2] != num_kv_heads or k_cache.shape[3] != head_dim:
        raise ValueError("k_cache must be (num_blocks, block_size, num_kv_heads, head_dim)")
    if k_cache.shape[1] != block_size:
        raise ValueError(f"block_size {block_size} != k_cache.shape[1] {k_cache.shape[1]}")
    if not (k_cache.is_contiguous() and v_cache.is_contiguous()):
        raise ValueError("k_cache and v_cache must be contiguous")
    if key.dtype != k_cache.dtype or value.dtype != v_cache.dtype or key.dtype != value.dtype:
        raise ValueError("key, value and the caches must share one dtype")

    def rows(x: torch.Tensor) -> torch.Tensor:
        dense = x.stride(2) == 1 and x.stride(1) == head_dim and (
            num_tokens <= 1 or x.stride(0) >= num_kv_heads * head_dim)
        return x if dense else x.contiguous()

    key, value = rows(key), rows(value)
    slot_mapping = slot_mapping.contiguous()
    dense_stride = num_kv_heads * head_dim
    with _on_device(dev):
        _ext.check(_ext.lib().mvl_store_kvcache(
            key.data_ptr(), value.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
            slot_mapping.data_ptr(), _index_flag("slot_mapping", slot_mapping), num_tokens,
            num_kv_heads, head_dim,
            key.stride(0) if num_tokens > 1 else dense_stride,
            value.stride(0) if num_tokens > 1 else dense_stride,
            k_cache.shape[0] * k_cache.shape[1], _dtype_code(key), _stream_ptr(dev)))


def paged_attention_decode(query: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                           block_tables: torch.Tensor, context_lens: torch.Tensor, scale: float,
                           num_heads: int, num_kv_heads: int, head_dim: int, block_size: int,
                           *, out: Optional[torch.Tensor] = None, chunk_tokens: int = 0,
                           flags: Optional[int] = None, _store=None) -> torch.Tensor:
    """Decode attention of one query per sequence over the paged KV cache.

    query: (batch, num_heads, head_dim); caches as in ``store_kvcache``; block_tables:
    (batch, max_num_blocks) int32, padded with -1; context_lens: (batch,) int32/int64.
    Returns a new (batch, num_heads, head_dim) tensor in query's dtype (reference
    attention.py:418-473).  ``out``, ``chunk_tokens`` and ``flags`` are keyword-only additions
    (preallocated output, split size override, ABI flags); defaults reproduce the reference call
    and merge the chunks inside the split kernel (MVL_DECODE_FUSED_COMBINE).
    """
    if flags is None:
        flags = DEFAULT_DECODE_FLAGS
    dev = _need_cuda(query=query, k_cache=k_cache, v_cache=v_cache, block_tables=block_tables,
                     context_lens=context_lens)
    if query.dim() != 3 or query.shape[1] != num_heads or query.shape[2] != head_dim:
        raise ValueError("query must be (batch, num_heads, head_dim)")
    if k_cache.dim() != 4 or k_cache.shape != v_cache.shape:
        raise ValueError("k_cache and v_cache must be (num_blocks, block_size, num_kv_heads, head_dim)")
    if tuple(k_cache.shape[1:]) != (block_size, num_kv_heads, head_dim):
        raise ValueError(
            f"cache shape {tuple(k_cache.shape)} does not match block_size={block_size}, "
            f"num_kv_heads={num_kv_heads}, head_dim={head_dim}")
    if not (k_cache.is_contiguous() and v_cache.is_contiguous()):
        raise ValueError("k_cache and v_cache must be contiguous")
    if query.dtype != k_cache.dtype or k_cache.dtype != v_cache.dtype:
        raise ValueError("query and the caches must share one dtype")
    if block_tables.dim() != 2 or block_tables.dtype != torch.int32:
        raise ValueError("block_tables must be a 2-D int32 tensor")
    batch = query.shape[0]
    if block_tables.shape[0] != batch or context_lens.numel() != batch:
        raise ValueError("block_tables / context_lens batch size does not match query")
    query = query.contiguous()
    context_lens = context_lens.contiguous()
    if block_tables.stride(1) != 1:
        block_tables = block_tables.contiguous()
    if out is None:
        out = torch.empty_like(query)
    elif out.shape != query.shape or out.dtype != query.dtype or not out.is_contiguous():
        raise ValueError("out must be a contiguous tensor shaped and typed like query")
    max_num_blocks = block_tables.shape[1]
    if batch == 0 or max_num_blocks == 0:
        return out.zero_() if max_num_blocks == 0 else out
    code = _dtype_code(query)
    handle = _ext.lib()
    with _on_device(dev):
        key = (batch, num_heads, num_kv_heads, head_dim, max_num_blocks, block_size, code, chunk_tokens,
               dev.index)
        nbytes = _plan_cache.get(key)
        if nbytes is None:
            nbytes = _plan_cache[key] = handle.mvl_paged_decode_workspace_bytes(
                batch, num_heads, num_kv_heads, head_dim, max_num_blocks, block_size, code, chunk_tokens)
        ws = workspace_for(dev, nbytes)
        wsp, wsn = (ws.data_ptr(), ws.numel()) if ws is not None else (None, 0)
        if _store is None:
            _ext.check(handle.mvl_paged_decode(
                out.data_ptr(), query.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                block_tables.data_ptr(), context_lens.data_ptr(),
                _index_flag("context_lens", context_lens), batch, max_num_blocks,
                block_tables.stride(0), num_heads, num_kv_heads, head_dim, block_size,
                k_cache.shape[0], float(scale), code, wsp, wsn, chunk_tokens, flags, _stream_ptr(dev)))
        else:
            key, value, slots = _store
            _ext.check(handle.mvl_decode_step(
                out.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(), k_cache.data_ptr(),
                v_cache.data_ptr(), slots.data_ptr(), block_tables.data_ptr(), context_lens.data_ptr(),
                _index_flag("context_lens", context_lens), batch, max_num_blocks,
                block_tables.stride(0), key.stride(0), value.stride(0), num_heads, num_kv_heads,
                head_dim, block_size, k_cache.shape[0], float(scale), code, wsp, wsn, chunk_tokens,
                flags, _stream_ptr(dev)))
    return out


# the reference's benchmarks import the decode operator under this name (benchmark_decoding.py:107-118,
# benchmark_models.py:118); same signature, same kernel
paged_attention_decode_triton = paged_attention_decode


def flash_attention_prefill(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens: torch.Tensor,
                            scale: float, num_heads: int, num_kv_heads: int, head_dim: int,
                            *, max_seqlen: Optional[int] = None, flags: int = 0,
                            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Causal attention over packed variable-length sequences (prefill), reference signature
    (attention.py:212-221).  q (total_tokens, num_heads, head_dim); k, v (total_tokens,
    num_kv_heads, head_dim); cu_seqlens (num_seqs + 1) cumulative lengths.  Returns a new
    (total_tokens, num_heads, head_dim) tensor.  ``max_seqlen`` (keyword-only) sizes the grid
    without reading ``cu_seqlens`` back; when omitted it is read from the device, as the
    reference does (attention.py:263-264).  ``flags``: ``_ext.PREFILL_MMA_SYNC`` / ``PREFILL_ONE_TILE`` /
    ``PREFILL_PAIR`` force one kernel (default: by max_seqlen, see include/mvl_abi.h).  bf16 / fp16 on the
    tensor cores; fp32 (the reference's default dtype) on a CUDA-core kernel."""
    dev = _need_cuda(q=q, k=k, v=v, cu_seqlens=cu_seqlens)
    if q.dim() != 3 or k.dim() != 3 or v.shape != k.shape or q.shape[0] != k.shape[0]:
        raise ValueError("q must be (T, H, D) and k, v (T, KVH, D)")
    total, H, D = q.shape
    if (H, D) != (num_heads, head_dim) or k.shape[1:] != (num_kv_heads, head_dim):
        raise ValueError("tensor shapes do not match num_heads / num_kv_heads / head_dim")
    if q.dtype not in (torch.bfloat16, torch.float16, torch.float32) or k.dtype != q.dtype or v.dtype != q.dtype:
        raise NotImplementedError("prefill attention takes bf16 / fp16 (tensor cores) or fp32 (CUDA cores) inputs")

    def rows(x, heads):
        dense = x.stride(2) == 1 and x.stride(1) == D and (total <= 1 or x.stride(0) >= heads * D)
        x = x if dense else x.contiguous()
        return x, (x.stride(0) if total > 1 else heads * D)

    (q, qs), (k, ks), (v, vs) = rows(q, H), rows(k, num_kv_heads), rows(v, num_kv_heads)
    cu = cu_seqlens.to(torch.int32).contiguous()
    num_seqs = cu.numel() - 1
    if out is None:
        out = torch.empty((total, H, D), dtype=q.dtype, device=dev)
    elif out.shape != (total, H, D) or out.dtype != q.dtype or out.device != dev or not out.is_contiguous():
        raise ValueError("out must be a contiguous (total_tokens, num_heads, head_dim) tensor like q")
    if total == 0 or num_seqs <= 0:
        return out
    if max_seqlen is None:
        max_seqlen = int((cu[1:] - cu[:-1]).max())
    with _on_device(dev):
        _ext.check(_ext.lib().mvl_flash_prefill(
            out.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), qs, ks, vs, total, cu.data_ptr(),
            num_seqs, int(max_seqlen), H, num_kv_heads, D, float(scale), _dtype_code(q), int(flags),
            _stream_ptr(dev)))
    return out


def flash_attention_prefill_paged(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                                  cu_seqlens_q: torch.Tensor, block_tables: torch.Tensor,
                                  context_lens: torch.Tensor, scale: float, num_heads: int, num_kv_heads: int,
                                  head_dim: int, block_size: int, *, max_seqlen_q: Optional[int] = None,
                                  flags: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Prefill attention over the paged cache: the ``n_s`` new queries of sequence ``s`` (packed in
    ``q`` and delimited by ``cu_seqlens_q``) attend causally to all ``context_lens[s]`` keys of the
    sequence — a cached prefix plus the new tokens, whose K/V must already be in the cache.  This is
    what the reference's prefill branch is handed when a prefix was cached (``block_tables`` set,
    ``cu_seqlens_k > cu_seqlens_q``, model_runner.py:209-215) and does not do (attention.py:514-525,
    SURVEY.md A-8); it also serves chunked prefill.  bf16 / fp16 (``scale > 0``) or fp32.  Returns
    (total_q_tokens, num_heads, head_dim)."""
    dev = _need_cuda(q=q, k_cache=k_cache, v_cache=v_cache, cu_seqlens_q=cu_seqlens_q,
                     block_tables=block_tables, context_lens=context_lens)
    if q.dim() != 3 or q.shape[1:] != (num_heads, head_dim):
        raise ValueError("q must be (total_q_tokens, num_heads, head_dim)")
    if k_cache.dim() != 4 or k_cache.shape != v_cache.shape or k_cache.shape[1:] != (
            block_size, num_kv_heads, head_dim) or not (k_cache.is_contiguous() and v_cache.is_contiguous()):
        raise ValueError("caches must be contiguous (num_blocks, block_size, num_kv_heads, head_dim)")
    if q.dtype not in (torch.bfloat16, torch.float16, torch.float32) or k_cache.dtype != q.dtype or \
            v_cache.dtype != q.dtype:
        raise NotImplementedError("prefill attention takes bf16 / fp16 (tensor cores) or fp32 (CUDA cores) inputs")
    if q.dtype != torch.float32 and not scale > 0:
        raise ValueError("paged prefill on the tensor cores needs scale > 0")
    total, H, D = q.shape
    dense = q.stride(2) == 1 and q.stride(1) == D and (total <= 1 or q.stride(0) >= H * D)
    q = q if dense else q.contiguous()
    qs = q.stride(0) if total > 1 else H * D
    cu = cu_seqlens_q.to(torch.int32).contiguous()
    num_seqs = cu.numel() - 1
    if block_tables.dim() != 2 or block_tables.shape[0] != num_seqs or context_lens.numel() != num_seqs:
        raise ValueError("block_tables (num_seqs, max_blocks) and context_lens (num_seqs,) must match cu_seqlens_q")
    bt = block_tables if block_tables.dtype == torch.int32 and block_tables.stride(1) == 1 else \
        block_tables.to(torch.int32).contiguous()
    ctx = context_lens.contiguous()
    if out is None:
        out = torch.empty((total, H, D), dtype=q.dtype, device=dev)
    elif out.shape != (total, H, D) or out.dtype != q.dtype or out.device != dev or not out.is_contiguous():
        raise ValueError("out must be a contiguous (total_tokens, num_heads, head_dim) tensor like q")
    if total == 0 or num_seqs <= 0:
        return out
    if max_seqlen_q is None:
        max_seqlen_q = int((cu[1:] - cu[:-1]).max())
    with _on_device(dev):
        _ext.check(_ext.lib().mvl_flash_prefill_paged(
            out.data_ptr(), q.data_ptr(), qs, total, cu.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
            bt.data_ptr(), bt.stride(0), ctx.data_ptr(), _index_flag("context_lens", ctx), num_seqs,
            int(max_seqlen_q), H, num_kv_heads, D, block_size, k_cache.shape[0], float(scale), _dtype_code(q),
            int(flags), _stream_ptr(dev)))
    return out


def decode_step(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
                v_cache: torch.Tensor, slot_mapping: torch.Tensor, block_tables: torch.Tensor,
                context_lens: torch.Tensor, scale: float, num_heads: int, num_kv_heads: int,
                head_dim: int, block_size: int, *, out: Optional[torch.Tensor] = None,
                chunk_tokens: int = 0, flags: Optional[int] = None) -> torch.Tensor:
    """``store_kvcache`` then ``paged_attention_decode`` of one decode step in ONE kernel launch
    (``mvl_decode_step``).  ``key``/``value`` are the step's new rows (batch, num_kv_heads,
    head_dim); ``slot_mapping[b]`` is the slot of token ``context_lens[b] - 1`` (or -1), as the
    reference's ``prepare_decode`` produces (model_runner.py:238-239).  Same result and same cache
    contents as the two separate calls."""
    _need_cuda(key=key, value=value, slot_mapping=slot_mapping)
    batch = query.shape[0]
    if key.shape != (batch, num_kv_heads, head_dim) or value.shape != key.shape:
        raise ValueError("key/value must be (batch, num_kv_heads, head_dim) for a decode step")
    if key.dtype != k_cache.dtype or value.dtype != k_cache.dtype:
        raise ValueError("key, value and the caches must share one dtype")
    if slot_mapping.numel() != batch or slot_mapping.dtype != context_lens.dtype:
        raise ValueError("slot_mapping must be (batch,) with context_lens' integer dtype")

    def rows(x: torch.Tensor) -> torch.Tensor:
        dense = x.stride(2) == 1 and x.stride(1) == head_dim and (
            batch <= 1 or x.stride(0) >= num_kv_heads * head_dim)
        x = x if dense else x.contiguous()
        return x if batch > 1 else x.reshape(1, num_kv_heads, head_dim).contiguous()

    return paged_attention_decode(query, k_cache, v_cache, block_tables, context_lens, scale, num_heads,
                                  num_kv_heads, head_dim, block_size, out=out,
                                  chunk_tokens=chunk_tokens, flags=flags,
                                  _store=(rows(key), rows(value), slot_mapping.contiguous()))


class Attention(nn.Module):
    """Paged attention layer with the reference's constructor and forward signature.

    ``k_cache`` / ``v_cache`` start as empty placeholders and are assigned per-layer views of the
    engine's cache allocation by the model runner, which finds this module by those two
    attribute names (reference attention.py:477-491, model_runner.py:164-168).
    """

    def __init__(self, num_heads: int, head_dim: int, scale: float = 1.0,
                 num_kv_heads: Optional[int] = None, block_size: int = 16):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.scale = scale
        self.num_kv_heads = num_kv_heads if num_kv_heads is not None else num_heads
        self.block_size = block_size
        self.k_cache = self.v_cache = torch.tensor([])
        # mvl_paged_decode flags of the decode step (None = DEFAULT_DECODE_FLAGS).  An engine that knows
        # what runs in front of this layer on the stream may add DECODE_OVERLAP_PREV (include/mvl_abi.h).
        self.decode_flags: Optional[int] = None

    def _effective_block_size(self) -> int:
        # The reference builds Attention with block_size 16 while the engine's cache uses 256
        # (qwen3.py:56-61 vs model_runner.py:19,162).  The attached cache
        # is the ground truth for the page size.
        if self.k_cache.dim() == 4:
            return self.k_cache.shape[1]
        return self.block_size

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, *,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``out`` (keyword-only, optional): a preallocated (tokens, num_heads, head_dim) result
        buffer, for callers that run under CUDA-graph capture with static buffers (decode) or that
        collect the layers' results in one tensor (prefill)."""
        context = get_context()
        k_cache, v_cache = self.k_cache, self.v_cache
        block_size = self._effective_block_size()

        has_cache = k_cache.numel() > 0 and v_cache.numel() > 0 and context.slot_mapping is not None
        # The reference's own caller can hand over mixed dtypes: with a 16-bit model its rotary
        # embedding multiplies by an fp32 table, so q and k arrive in fp32 next to a 16-bit v and
        # 16-bit caches (rotary_embedding.py:15-23; qwen3.py:108-110).  Its Triton kernels cast on
        # store / load; here the operands are brought to the cache's dtype once and the result goes
        # back to the caller's q dtype (the reference returns empty_like(query), attention.py:450).
        want_dtype = q.dtype
        if k_cache.numel() > 0 and (q.dtype != k_cache.dtype or k.dtype != k_cache.dtype or v.dtype != k_cache.dtype):
            q, k, v = q.to(k_cache.dtype), k.to(k_cache.dtype), v.to(k_cache.dtype)
            if out is not None and out.dtype != k_cache.dtype:
                raise ValueError("out= must have the cache's dtype when the operands are cast to it")
            o = self.forward(q, k, v, out=out)
            return o.to(want_dtype)
        if has_cache and k.dim() == 4:  # (B, N, KVH, D) -> (B*N, KVH, D)
            k = k.reshape(-1, k.shape[-2], k.shape[-1])
            v = v.reshape(-1, v.shape[-2], v.shape[-1])
        scale = self.scale / (self.head_dim ** 0.5)

        if context.is_prefill:
            # reference attention.py:514-531: store, then causal attention over the new tokens
            if has_cache:
                store_kvcache(k, v, k_cache, v_cache, context.slot_mapping, block_size)
            if q.dim() != 3:
                raise NotImplementedError("batched (4-D) prefill: the reference calls an undefined "
                                          "flash_attention_batched here (attention.py:528)")
            if context.cu_seqlens_q is None:
                raise ValueError("cu_seqlens_q must be provided for varlen attention")
            if context.block_tables is not None and has_cache:
                # a prefix of these prompts is already cached (model_runner.py:209-215): attend over
                # the pages — cached prefix + the rows just stored — not over the new tokens only
                if context.context_lens is not None:
                    ctx_lens = context.context_lens
                elif context.cu_seqlens_k is not None:
                    ctx_lens = (context.cu_seqlens_k[1:] - context.cu_seqlens_k[:-1]).to(torch.int32)
                else:
                    raise ValueError("prefill over cached pages needs cu_seqlens_k (or context_lens)")
                o = flash_attention_prefill_paged(q, k_cache, v_cache, context.cu_seqlens_q,
                                                  context.block_tables, ctx_lens, scale, self.num_heads,
                                                  self.num_kv_heads, self.head_dim, block_size,
                                                  max_seqlen_q=context.max_seqlen_q or None, out=out)
                return o.reshape(o.shape[0], self.num_heads * self.head_dim)
            o = flash_attention_prefill(q, k, v, context.cu_seqlens_q, scale, self.num_heads,
                                        self.num_kv_heads, self.head_dim,
                                        max_seqlen=context.max_seqlen_q or None, out=out)
            return o.reshape(o.shape[0], self.num_heads * self.head_dim)
        if q.dim() == 4:  # (B, 1, H, D) -> (B, H, D)
            q = q.squeeze(1)
        one_token_each = (has_cache and k.shape[0] == q.shape[0]
                          and context.slot_mapping.dtype == context.context_lens.dtype)
        if one_token_each:
            # the reference's decode step: one new token per sequence -> store fused into decode
            o = decode_step(q, k, v, k_cache, v_cache, context.slot_mapping, context.block_tables,
                            context.context_lens, scale, self.num_heads, self.num_kv_heads,
                            self.head_dim, block_size, out=out, flags=self.decode_flags)
        else:
            if has_cache:
                store_kvcache(k, v, k_cache, v_cache, context.slot_mapping, block_size)
            o = paged_attention_decode(q, k_cache, v_cache, context.block_tables, context.context_lens,
                                       scale, self.num_heads, self.num_kv_heads, self.head_dim,
                                       block_size, out=out)
        return o.reshape(o.shape[0], self.num_heads * self.head_dim)


class HostDecodeStep:
    """One layer's decode-attention step driven from HOST (pinned) buffers.

    Wraps ``mvl_decode_step_host``: per call it copies the step's inputs (q, the new token's k/v,
    slot_mapping, context_lens, block_tables) host->device into staging buffers owned by this
    object, scatters k/v into the resident paged cache, runs the split-KV decode and copies the
    (batch, num_heads, head_dim) result device->host.  This is the path an engine takes when its
    model runner builds the step on the CPU (reference model_runner.py:231-256 then
    attention.py:493-557).  The KV cache itself stays resident in HBM, as in the reference engine.
    All buffers are allocated here, once; ``__call__`` allocates nothing.
    """

    def __init__(self, k_cache: torch.Tensor, v_cache: torch.Tensor, batch: int, max_num_blocks: int,
                 num_heads: int, scale: float, index_dtype: torch.dtype = torch.int64,
                 chunk_tokens: int = 0, flags: int = 0, pieces: Optional[int] = None,
                 direct_out: bool = True, direct_q: Optional[bool] = None, direct_kv: bool = True):
        """``direct_out`` / ``direct_q`` (serial form only; default on): the kernels write the result rows
        straight into the pinned host buffer / read q straight from it over PCIe as they go, instead of
        a separate copy before and after (C3 on B200: 0.856 -> 0.766 ms per step; same bits).
        ``direct_q`` unset: on, except with ``DECODE_SHARE_PREFIX`` — there every work item of the group
        kernel and every chunk CTA behind it would start with a PCIe round trip for its q rows (C5 on two
        boxes: 0.388 / 0.580 ms with q read across PCIe, 0.378 / 0.451 ms with q in the step's copy).
        ``direct_kv``: the new token's k / v rows likewise (read once, by the fused store).
        ``pieces``: how many runs of consecutive sequences the step is cut into so that the copies
        of one run travel while another computes (``mvl_decode_step_host_overlapped``); 1 = the strictly
        serial ``mvl_decode_step_host`` (the default: measured on B200 the smaller launches lose what the overlap gains)."""
        dev = _need_cuda(k_cache=k_cache, v_cache=v_cache)
        if k_cache.dim() != 4 or k_cache.shape != v_cache.shape or not (
                k_cache.is_contiguous() and v_cache.is_contiguous()):
            raise ValueError("caches must be contiguous (num_blocks, block_size, num_kv_heads, head_dim)")
        self.dev, self.k_cache, self.v_cache = dev, k_cache, v_cache
        self.num_blocks, self.block_size, self.num_kv_heads, self.head_dim = k_cache.shape
        self.batch, self.max_num_blocks, self.num_heads = batch, max_num_blocks, num_heads
        self.scale, self.chunk_tokens, self.flags = float(scale), chunk_tokens, flags
        self.idx_flag = 1 if index_dtype == torch.int64 else 0
        dt = k_cache.dtype
        self.code = _DTYPES[dt]

        # The six input arrays live back to back in ONE pinned arena and ONE device arena, so the
        # library moves a step's inputs with a single host-to-device copy.
        es = k_cache.element_size()
        isz = 8 if index_dtype == torch.int64 else 4
        sizes = [batch * num_heads * self.head_dim * es, batch * self.num_kv_heads * self.head_dim * es,
                 batch * self.num_kv_heads * self.head_dim * es, batch * isz, batch * isz,
                 batch * max_num_blocks * 4]
        total = sum(sizes)
        self._h_arena = torch.empty(total, dtype=torch.uint8, pin_memory=True)
        self._d_arena = torch.empty(total, dtype=torch.uint8, device=dev)

        def views(arena):
            out, at = [], 0
            shapes = [((batch, num_heads, self.head_dim), dt), ((batch, self.num_kv_heads, self.head_dim), dt),
                      ((batch, self.num_kv_heads, self.head_dim), dt), ((batch,), index_dtype),
                      ((batch,), index_dtype), ((batch, max_num_blocks), torch.int32)]
            for n, (shape, dtype) in zip(sizes, shapes):
                out.append(arena[at:at + n].view(dtype).view(shape))
                at += n
            return out

        self.q_h, self.k_h, self.v_h, self.slot_h, self.ctx_h, self.bt_h = views(self._h_arena)
        self.q_d, self.k_d, self.v_d, self.slot_d, self.ctx_d, self.bt_d = views(self._d_arena)
        self.out_h = torch.empty((batch, num_heads, self.head_dim), dtype=dt, pin_memory=True)
        self.out_d = torch.empty((batch, num_heads, self.head_dim), dtype=dt, device=dev)
        if direct_q is None:
            direct_q = not (flags & _ext.DECODE_SHARE_PREFIX)
        self.direct_out, self.direct_q, self.direct_kv = bool(direct_out), bool(direct_q), bool(direct_kv)
        self.pieces = 1 if pieces is None else max(1, min(int(pieces), 8, batch))
        with torch.cuda.device(dev):
            if self.pieces > 1:
                nbytes = _ext.lib().mvl_decode_step_host_workspace_bytes(
                    batch, num_heads, self.num_kv_heads, self.head_dim, max_num_blocks, self.block_size,
                    self.code, chunk_tokens, self.pieces)
            else:
                nbytes = _ext.lib().mvl_paged_decode_workspace_bytes(
                    batch, num_heads, self.num_kv_heads, self.head_dim, max_num_blocks, self.block_size,
                    self.code, chunk_tokens)
        self.ws = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=dev)  # counters start at zero
        self.ws_bytes = nbytes
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in
                             (self.q_h, self.k_h, self.v_h, self.slot_h, self.ctx_h, self.bt_h))
        self.d2h_bytes = self.out_h.numel() * self.out_h.element_size()

    def __call__(self, synchronize: bool = True) -> torch.Tensor:
        """Run one step on the host buffers ``q_h, k_h, v_h, slot_h, ctx_h, bt_h``; returns ``out_h``
        (valid once the stream is synchronised, which ``synchronize=True`` does)."""
        with torch.cuda.device(self.dev):
            stream = _stream_ptr(self.dev)
            direct = self.pieces == 1
            q_d = self.q_h.data_ptr() if (direct and self.direct_q) else self.q_d.data_ptr()
            out_d = self.out_h.data_ptr() if (direct and self.direct_out) else self.out_d.data_ptr()
            kv = direct and self.direct_kv
            k_d = self.k_h.data_ptr() if kv else self.k_d.data_ptr()
            v_d = self.v_h.data_ptr() if kv else self.v_d.data_ptr()
            args = (self.q_h.data_ptr(), self.k_h.data_ptr(), self.v_h.data_ptr(), self.slot_h.data_ptr(),
                    self.ctx_h.data_ptr(), self.bt_h.data_ptr(), self.out_h.data_ptr(),
                    q_d, k_d, v_d, self.slot_d.data_ptr(),
                    self.ctx_d.data_ptr(), self.bt_d.data_ptr(), out_d,
                    self.k_cache.data_ptr(), self.v_cache.data_ptr(), self.idx_flag, self.batch,
                    self.max_num_blocks, self.num_heads, self.num_kv_heads, self.head_dim,
                    self.block_size, self.num_blocks, self.scale, self.code, self.ws.data_ptr(),
                    self.ws_bytes, self.chunk_tokens, self.flags)
            if self.pieces > 1:
                _ext.check(_ext.lib().mvl_decode_step_host_overlapped(*args, self.pieces, stream))
            else:
                _ext.check(_ext.lib().mvl_decode_step_host(*args, stream))
            if synchronize:
                torch.cuda.current_stream(self.dev).synchronize()
        return self.out_h
