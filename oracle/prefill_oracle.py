"""CPU oracle for prefill attention.  TEST INFRASTRUCTURE ONLY.

Restates reference benchmark_prefilling.py:9-52 (``pytorch_standard_attention``): per sequence
delimited by cu_seqlens, expand KV heads to query heads (:33-36), scores = q.k^T * scale (:39),
mask the strict upper triangle with -inf (:42-43), softmax, probabilities @ v (:45-46) — evaluated
in float64.  Pinned against the reference function itself: tests/golden/make_golden_prefill.py ->
tests/golden/prefill_*.npz (tests/test_oracle_golden.py)."""
from __future__ import annotations

import torch


def causal_varlen_attention(q, k, v, cu_seqlens, scale, num_heads, num_kv_heads, head_dim):
    """q (T, H, D), k/v (T, KVH, D) -> float64 (T, H, D)."""
    out = torch.zeros(q.shape, dtype=torch.float64)
    bounds = [int(x) for x in cu_seqlens.tolist()]
    group = num_heads // num_kv_heads
    for s, e in zip(bounds[:-1], bounds[1:]):
        n = e - s
        if n <= 0:
            continue
        qs = q[s:e].double().permute(1, 0, 2)                       # (H, n, D)
        ks = k[s:e].double().permute(1, 0, 2).repeat_interleave(group, dim=0)
        vs = v[s:e].double().permute(1, 0, 2).repeat_interleave(group, dim=0)
        scores = qs @ ks.transpose(1, 2) * scale
        future = torch.ones(n, n, dtype=torch.bool).triu(1)
        scores = scores.masked_fill(future, float("-inf"))
        out[s:e] = (torch.softmax(scores, dim=-1) @ vs).permute(1, 0, 2)
    return out


def causal_paged_attention(q, k_cache, v_cache, cu_seqlens_q, block_tables, context_lens, scale, num_heads,
                           num_kv_heads, head_dim, block_size):
    """Prefill over cached pages: sequence s has context_lens[s] keys in the paged cache (gathered
    through its block-table row, as decode_attention_* do), its n_s new queries sit at positions
    ctx_s - n_s .. ctx_s - 1 and see keys 0 .. their own position.  float64 (T_q, H, D).

    Parity anchor: the reference has no working implementation of this case (its prefill ignores the
    pages), so this function is pinned through ``causal_varlen_attention`` (itself
    pinned to the reference's pytorch_standard_attention): the rows it returns must equal the LAST n_s
    rows of the dense causal attention over the sequence's full ctx_s tokens
    (tests/test_oracle_golden.py::test_paged_prefill_oracle_equals_dense_tail)."""
    out = torch.zeros(q.shape, dtype=torch.float64)
    bounds = [int(x) for x in cu_seqlens_q.tolist()]
    group = num_heads // num_kv_heads
    flat_k = k_cache.reshape(-1, num_kv_heads, head_dim)
    flat_v = v_cache.reshape(-1, num_kv_heads, head_dim)
    for s, (b, e) in enumerate(zip(bounds[:-1], bounds[1:])):
        n, ctx = e - b, int(context_lens[s])
        if n <= 0:
            continue
        pos = torch.arange(ctx)
        slots = block_tables[s, pos // block_size].long() * block_size + pos % block_size
        ks = flat_k[slots].double().permute(1, 0, 2).repeat_interleave(group, dim=0)   # (H, ctx, D)
        vs = flat_v[slots].double().permute(1, 0, 2).repeat_interleave(group, dim=0)
        qs = q[b:e].double().permute(1, 0, 2)                                           # (H, n, D)
        scores = qs @ ks.transpose(1, 2) * scale
        qpos = torch.arange(ctx - n, ctx)
        scores = scores.masked_fill(pos[None, :] > qpos[:, None], float("-inf"))
        out[b:e] = (torch.softmax(scores, dim=-1) @ vs).permute(1, 0, 2)
    return out
