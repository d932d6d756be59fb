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
