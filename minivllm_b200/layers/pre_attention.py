"""q/k RMSNorm + rotary embedding + KV-cache write in one kernel.

``fused_qk_norm_rope_store`` replaces what the reference runs between the QKV projection and the
attention call of every layer (reference models/qwen3.py:104-110): ``LayerNorm.rms_forward`` on q
and on k (layers/layernorm.py:13-15), ``RotaryEmbedding.forward`` (layers/rotary_embedding.py:74-83)
and ``store_kvcache`` (layers/attention.py:498-510).  ``rope_cache`` builds the reference's
``cos_sin_cache`` (rotary_embedding.py:52-68).  ``Qwen3AttentionCore`` strings the fused
pre-attention pass and the paged decode attention together: the part of
``Qwen3Attention.forward`` (qwen3.py:86-111) that lies between its two GEMMs, for decode steps.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import _ext
from ..utils.context import get_context
from .attention import (Attention, _dtype_code, _index_flag, _need_cuda, _on_device, _stream_ptr,
                        paged_attention_decode)


def rope_cache(base: float, rotary_dim: int, max_position: int, device=None) -> torch.Tensor:
    """(max_position, rotary_dim) fp32 table, row p = [cos(p * inv_freq) | sin(p * inv_freq)].

    Same quantities as the reference's ``cos_sin_cache`` (rotary_embedding.py:61-71): fp32
    ``inv_freq``, fp32 angles ``p * inv_freq`` (one IEEE product each, the same on every host).  The
    reference then takes cos / sin with the host's fp32 vector math, whose last bits differ from CPU
    to CPU; here they are evaluated in float64 on those fp32 angles and rounded once, so the table is
    the correctly rounded value of the reference's expression and does not depend on the host.
    """
    exponent = torch.arange(0, rotary_dim, 2, dtype=torch.float64) / rotary_dim
    inv_freq = (1.0 / (torch.tensor(float(base), dtype=torch.float64) ** exponent)).float()
    freqs = (torch.arange(max_position, dtype=torch.float32)[:, None] * inv_freq[None, :]).double()
    return torch.cat([freqs.cos(), freqs.sin()], dim=-1).float().to(device)


def fused_qk_norm_rope_store(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, positions: torch.Tensor,
                             gamma: Optional[torch.Tensor], eps: float,
                             cos_sin_cache: Optional[torch.Tensor], k_cache: Optional[torch.Tensor],
                             v_cache: Optional[torch.Tensor], slot_mapping: Optional[torch.Tensor],
                             *, q_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q (N, H, D), k/v (N, KVH, D) (dense head dims, any token stride) -> normalised + rotated q
    (N, H, D); k (normalised + rotated) and v are written to the paged caches at ``slot_mapping``."""
    dev = _need_cuda(q=q, k=k, v=v, positions=positions)
    if q.dim() != 3 or k.dim() != 3 or v.shape != k.shape or q.shape[0] != k.shape[0]:
        raise ValueError("q must be (N, H, D) and k, v (N, KVH, D)")
    n, H, D = q.shape
    KVH = k.shape[1]
    if k.shape[2] != D or not (q.dtype == k.dtype == v.dtype):
        raise ValueError("q, k, v must share head_dim and dtype")

    def rows(x, heads):
        dense = x.stride(2) == 1 and x.stride(1) == D and (n <= 1 or x.stride(0) >= heads * D)
        x = x if dense else x.contiguous()
        return x, (x.stride(0) if n > 1 else heads * D)

    (q, qs), (k, ks), (v, vs) = rows(q, H), rows(k, KVH), rows(v, KVH)
    if q_out is None:
        q_out = torch.empty((n, H, D), dtype=q.dtype, device=dev)
    elif q_out.shape != (n, H, D) or q_out.dtype != q.dtype or not q_out.is_contiguous():
        raise ValueError("q_out must be a contiguous (N, H, D) tensor of q's dtype")
    if gamma is not None:
        gamma = gamma.to(device=dev, dtype=torch.float32).contiguous()
        if gamma.numel() != D:
            raise ValueError("gamma must have head_dim elements")
    max_pos = 0
    if cos_sin_cache is not None:
        if cos_sin_cache.dtype != torch.float32 or cos_sin_cache.dim() != 2 or cos_sin_cache.shape[1] != D \
                or not cos_sin_cache.is_contiguous() or cos_sin_cache.device != dev:
            raise ValueError("cos_sin_cache must be a contiguous fp32 (max_position, head_dim) CUDA tensor")
        max_pos = cos_sin_cache.shape[0]
        if positions.numel() != n:
            raise ValueError("positions must have one entry per token")
    positions = positions.contiguous()
    num_slots = 0
    if slot_mapping is not None:
        _need_cuda(k_cache=k_cache, v_cache=v_cache, slot_mapping=slot_mapping)
        if k_cache.shape != v_cache.shape or tuple(k_cache.shape[2:]) != (KVH, D) or k_cache.dtype != k.dtype \
                or not (k_cache.is_contiguous() and v_cache.is_contiguous()):
            raise ValueError("caches must be contiguous (num_blocks, block_size, KVH, D) of k's dtype")
        if slot_mapping.numel() != n:
            raise ValueError("Slot mapping size must match number of tokens")
        slot_mapping = slot_mapping.contiguous()
        num_slots = k_cache.shape[0] * k_cache.shape[1]
    with _on_device(dev):
        _ext.check(_ext.lib().mvl_qk_norm_rope_store(
            q_out.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), qs, ks, vs,
            gamma.data_ptr() if gamma is not None else None, float(eps),
            cos_sin_cache.data_ptr() if cos_sin_cache is not None else None, max_pos,
            positions.data_ptr(), _index_flag("positions", positions),
            k_cache.data_ptr() if slot_mapping is not None else None,
            v_cache.data_ptr() if slot_mapping is not None else None,
            slot_mapping.data_ptr() if slot_mapping is not None else None,
            _index_flag("slot_mapping", slot_mapping) if slot_mapping is not None else 0,
            num_slots, n, H, KVH, D, _dtype_code(q), _stream_ptr(dev)))
    return q_out


class Qwen3AttentionCore(nn.Module):
    """What ``Qwen3Attention.forward`` does between ``qkv_projection`` and ``output_projection``
    (reference models/qwen3.py:86-111) for a decode step: split the fused QKV activation, RMSNorm q
    and k (skipped when ``qkv_bias``), rotate both, write k/v to the cache, attend.  Two launches
    (pre-attention pass, decode) instead of the reference's seven."""

    def __init__(self, num_heads: int, num_kv_heads: int, head_dim: int, scale: float = 1.0,
                 qkv_bias: bool = False, base: float = 10000, max_position: int = 16384,
                 eps: float = 1e-5, block_size: int = 16):
        super().__init__()
        self.num_heads, self.num_kv_heads, self.head_dim = num_heads, num_kv_heads, head_dim
        self.q_size, self.kv_size = num_heads * head_dim, num_kv_heads * head_dim
        self.qkv_bias, self.eps = qkv_bias, eps
        self.register_buffer("gamma", torch.ones(head_dim))
        # the rotation table stays fp32 whatever dtype the module is cast to (it is fp32 in the
        # reference too, rotary_embedding.py:60-68), so it is not a buffer; one copy per device
        self._rope_args = (base, head_dim, max_position)
        self._rope_tables: dict = {}
        self.attention = Attention(num_heads, head_dim, scale, num_kv_heads, block_size)

    def forward(self, qkv: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        """qkv (tokens, (H + 2*KVH) * D) -> (tokens, H * D)."""
        ctx = get_context()
        if ctx.is_prefill:
            raise NotImplementedError("prefill attention is outside the B200 decode path (§8(f) N3)")
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        q = q.view(-1, self.num_heads, self.head_dim)
        k = k.view(-1, self.num_kv_heads, self.head_dim)
        v = v.view(-1, self.num_kv_heads, self.head_dim)
        att = self.attention
        has_cache = att.k_cache.numel() > 0 and ctx.slot_mapping is not None
        cache = self._rope_tables.get(qkv.device)
        if cache is None:
            cache = self._rope_tables[qkv.device] = rope_cache(*self._rope_args, device=qkv.device)
        q = fused_qk_norm_rope_store(
            q, k, v, positions, None if self.qkv_bias else self.gamma, self.eps, cache,
            att.k_cache if has_cache else None, att.v_cache if has_cache else None,
            ctx.slot_mapping if has_cache else None)
        block_size = att._effective_block_size()
        o = paged_attention_decode(q, att.k_cache, att.v_cache, ctx.block_tables, ctx.context_lens,
                                   att.scale / (self.head_dim ** 0.5), self.num_heads,
                                   self.num_kv_heads, self.head_dim, block_size)
        return o.reshape(o.shape[0], self.num_heads * self.head_dim)
