"""CPU oracle for the paged-KV decode-attention path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this module; the product (``minivllm_b200``) never does.

Parity pinning: the reference ships no tests or golden vectors for this path, so
``decode_attention_padded`` was pinned against the reference's own function
``decode_torch_optimized`` (reference benchmark_decoding.py:136-189) imported and run on CPU in
the build container; ``tests/golden/make_golden.py`` is the committed script that did it and
``tests/golden/*.npz`` are the vectors it wrote (``tests/test_oracle_golden.py`` re-checks them).

Functions
  store_kvcache_oracle     restates the index math of the Triton store kernel, reference
                           src/myvllm/layers/attention.py:27-64 (no PyTorch version exists there).
  decode_attention_padded  restates reference benchmark_decoding.py:136-189: gather each
                           sequence's pages, zero-pad to the batch maximum, expand KV heads to query
                           heads, scores = q.K^T * scale, mask past context_len with -inf, softmax,
                           probabilities @ V.  Vectorised torch on CPU in the inputs' dtype — this
                           is the "port" timed as the CPU baseline.
  decode_attention_exact   independent per-sequence float64 numpy evaluation of the same
                           definition (semantics of reference attention.py:303-415: GQA map
                           h // (H // KVH), token j at cache[table[j // BS], j % BS]).
"""
from __future__ import annotations

import numpy as np
import torch


def store_kvcache_oracle(key, value, k_cache, v_cache, slot_mapping):
    """cache.view(-1, KVH, D)[s] = x[t] for every token t whose slot s is not -1.

    Follows reference attention.py:27-64: a token with slot -1 is skipped (:29-30); the flat
    element offset is slot * KVH * D + head * D + d (:33-34, :52-56), independent of block size.
    Works on numpy arrays or torch tensors; mutates the caches in place.
    """
    slots = np.asarray(slot_mapping.cpu() if isinstance(slot_mapping, torch.Tensor) else slot_mapping)
    keep = np.nonzero(slots != -1)[0]
    if isinstance(k_cache, torch.Tensor):
        kf = k_cache.view(-1, k_cache.shape[-2], k_cache.shape[-1])
        vf = v_cache.view(-1, v_cache.shape[-2], v_cache.shape[-1])
        idx = torch.as_tensor(slots[keep], dtype=torch.long)
        src = torch.as_tensor(keep, dtype=torch.long)
        kf[idx] = key[src]
        vf[idx] = value[src]
    else:
        kf = k_cache.reshape(-1, k_cache.shape[-2], k_cache.shape[-1])
        vf = v_cache.reshape(-1, v_cache.shape[-2], v_cache.shape[-1])
        kf[slots[keep]] = key[keep]
        vf[slots[keep]] = value[keep]


def _slot_table(block_tables: torch.Tensor, context_lens: torch.Tensor, block_size: int, width: int):
    """(B, width) flat slot index of token j of each sequence, and the (B, width) validity mask."""
    pos = torch.arange(width)
    valid = pos[None, :] < context_lens[:, None].to(torch.long)
    blk = torch.gather(block_tables.to(torch.long), 1,
                       (pos // block_size).clamp_(max=block_tables.shape[1] - 1)[None, :]
                       .expand(block_tables.shape[0], -1))
    slots = blk * block_size + (pos % block_size)[None, :]
    return torch.where(valid, slots, torch.zeros_like(slots)), valid


def decode_attention_padded(q, k_cache, v_cache, block_tables, context_lens, scale, num_heads,
                            num_kv_heads, head_dim, block_size):
    """Vectorised-PyTorch paged decode attention (restates benchmark_decoding.py:136-189).

    q (B, H, D); caches (NB, BS, KVH, D); block_tables (B, max_blocks) int32 padded with -1;
    context_lens (B,).  Returns (B, H, D) in q's dtype.  Every context_len must be >= 1 (the
    reference returns NaN rows for 0).
    """
    batch = q.shape[0]
    width = int(context_lens.max())
    slots, valid = _slot_table(block_tables, context_lens, block_size, width)
    flat_k = k_cache.reshape(-1, num_kv_heads, head_dim)
    flat_v = v_cache.reshape(-1, num_kv_heads, head_dim)
    # gather + zero padding (reference :154-168)
    keys = flat_k[slots.reshape(-1)].reshape(batch, width, num_kv_heads, head_dim)
    vals = flat_v[slots.reshape(-1)].reshape(batch, width, num_kv_heads, head_dim)
    keep = valid[:, :, None, None]
    keys = torch.where(keep, keys, torch.zeros((), dtype=keys.dtype))
    vals = torch.where(keep, vals, torch.zeros((), dtype=vals.dtype))
    # GQA expansion (reference :170-174)
    group = num_heads // num_kv_heads
    if group != 1:
        keys = keys.repeat_interleave(group, dim=2)
        vals = vals.repeat_interleave(group, dim=2)
    keys = keys.permute(0, 2, 1, 3)  # (B, H, N, D)
    vals = vals.permute(0, 2, 1, 3)
    scores = torch.matmul(q[:, :, None, :], keys.transpose(-1, -2)) * scale  # (B, H, 1, N)
    scores = scores.masked_fill(~valid[:, None, None, :], float("-inf"))
    probs = torch.softmax(scores, dim=-1)
    return torch.matmul(probs, vals)[:, :, 0, :]


def decode_attention_exact(q, k_cache, v_cache, block_tables, context_lens, scale, num_heads,
                           num_kv_heads, head_dim, block_size):
    """float64 per-sequence evaluation; returns float64 numpy (B, H, D).  context_len 0 -> zeros."""
    def to_np(x):
        if isinstance(x, torch.Tensor):
            x = x.detach().cpu()
            if x.dtype in (torch.bfloat16, torch.float16):
                x = x.float()
            return x.numpy()
        return np.asarray(x)

    qn, kn, vn = (to_np(t).astype(np.float64) for t in (q, k_cache, v_cache))
    bt, cl = to_np(block_tables), to_np(context_lens)
    flat_k = kn.reshape(-1, num_kv_heads, head_dim)
    flat_v = vn.reshape(-1, num_kv_heads, head_dim)
    group = num_heads // num_kv_heads
    out = np.zeros((qn.shape[0], num_heads, head_dim))
    for b in range(qn.shape[0]):
        n = int(cl[b])
        if n <= 0:
            continue
        j = np.arange(n)
        slots = bt[b, j // block_size].astype(np.int64) * block_size + j % block_size
        kb, vb = flat_k[slots], flat_v[slots]  # (n, KVH, D)
        for h in range(num_heads):
            s = kb[:, h // group, :] @ qn[b, h] * scale
            p = np.exp(s - s.max())
            out[b, h] = (p / p.sum()) @ vb[:, h // group, :]
    return out
