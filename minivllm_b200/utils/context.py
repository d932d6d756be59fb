"""Per-process forward context: what the engine's model runner tells the attention layer.

Mirrors the interface of reference src/myvllm/utils/context.py:5-27 (same field names, same
positional order in ``set_context``, same replace-wholesale semantics) so that code written
against ``myvllm.utils`` works unchanged.  The decode path reads ``slot_mapping``,
``context_lens`` and ``block_tables`` (produced by reference model_runner.py:231-256).
"""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch

_FIELDS = (
    ("is_prefill", bool, False),
    ("cu_seqlens_q", Optional[torch.Tensor], None),
    ("cu_seqlens_k", Optional[torch.Tensor], None),
    ("max_seqlen_q", int, 0),
    ("max_seqlen_k", int, 0),
    ("slot_mapping", Optional[torch.Tensor], None),   # (num_tokens,) int64/int32, -1 = no slot
    ("context_lens", Optional[torch.Tensor], None),   # (batch,) int64/int32
    ("block_tables", Optional[torch.Tensor], None),   # (batch, max_num_blocks) int32, pad -1
)

Context = dataclasses.make_dataclass(
    "Context", [(n, t, dataclasses.field(default=d)) for n, t, d in _FIELDS]
)
Context.__doc__ = "Decode/prefill metadata of the current forward pass (module-global, per process)."

_current = Context()


def get_context() -> "Context":
    return _current


def set_context(is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
                slot_mapping=None, context_lens=None, block_tables=None) -> None:
    """Replace the current context (every field, not a merge), as the reference does."""
    global _current
    _current = Context(is_prefill, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                       slot_mapping, context_lens, block_tables)


def reset_context() -> None:
    global _current
    _current = Context()
