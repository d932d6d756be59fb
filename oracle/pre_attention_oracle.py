"""CPU oracle for the pre-attention pass (q/k RMSNorm + RoPE + KV store).  TEST INFRASTRUCTURE ONLY.

Restates, in float64 torch on CPU:
  rms_norm          reference src/myvllm/layers/layernorm.py:13-15:
                    variance = mean(x^2, -1) + eps;  x / sqrt(variance) * gamma
  rope_table        reference src/myvllm/layers/rotary_embedding.py:52-68:
                    inv_freq = 1 / base^(arange(0, d, 2) / d);  freqs = outer(positions, inv_freq);
                    table = cat([cos(freqs), sin(freqs)], -1)
  rotate            reference rotary_embedding.py:4-23 (varlen branch): x1, x2 = halves of the head
                    dim; out = cat([x1*cos - x2*sin, x1*sin + x2*cos], -1), cos/sin broadcast over heads
  pre_attention     reference src/myvllm/models/qwen3.py:104-108 followed by the cache write of
                    layers/attention.py:27-64 (slot -1 skipped)
Pinned against the reference's own LayerNorm / RotaryEmbedding modules run on CPU:
tests/golden/make_golden_pre.py -> tests/golden/pre_*.npz, checked by tests/test_oracle_golden.py.
"""
from __future__ import annotations

import torch


def rms_norm(x: torch.Tensor, gamma: torch.Tensor, eps: float) -> torch.Tensor:
    x = x.double()
    variance = x.pow(2).mean(dim=-1, keepdim=True) + eps
    return x / variance.sqrt() * gamma.double()


def rope_table(base: float, rotary_dim: int, max_position: int) -> torch.Tensor:
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2).float() / rotary_dim))
    freqs = torch.einsum("i,j->ij", torch.arange(max_position).float(), inv_freq)
    return torch.cat([freqs.cos(), freqs.sin()], dim=-1)  # fp32, as the reference builds it


def rotate(x: torch.Tensor, table_rows: torch.Tensor) -> torch.Tensor:
    """x (N, heads, D); table_rows (N, D) = table[positions]."""
    cos, sin = table_rows.double().chunk(2, dim=-1)
    cos, sin = cos.unsqueeze(1), sin.unsqueeze(1)
    x1, x2 = x.double().chunk(2, dim=-1)
    return torch.cat([x1 * cos - x2 * sin, x1 * sin + x2 * cos], dim=-1)


def pre_attention(q, k, v, positions, gamma, eps, table, k_cache, v_cache, slot_mapping):
    """Returns q' (float64) and writes k' (rounded to the cache dtype) and v into the caches."""
    qn = rms_norm(q, gamma, eps) if gamma is not None else q.double()
    kn = rms_norm(k, gamma, eps) if gamma is not None else k.double()
    if table is not None:
        rows = table[positions.long()]
        qn, kn = rotate(qn, rows), rotate(kn, rows)
    if slot_mapping is not None:
        keep = slot_mapping != -1
        kf = k_cache.view(-1, k_cache.shape[-2], k_cache.shape[-1])
        vf = v_cache.view(-1, v_cache.shape[-2], v_cache.shape[-1])
        kf[slot_mapping[keep].long()] = kn[keep].to(k_cache.dtype)
        vf[slot_mapping[keep].long()] = v[keep]
    return qn, kn
