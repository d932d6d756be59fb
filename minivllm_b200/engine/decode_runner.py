"""Engine-shaped decode loop for the attention path.

What the reference's ModelRunner does around the hot path, restricted to that path and repaired
where the reference cannot run:

* one KV-cache allocation ``(2, L, NB, BS, KVH, D)`` sliced per layer and attached to every module
  that has ``k_cache``/``v_cache`` (reference engine/model_runner.py:140-168);
* ``prepare_decode``: per-sequence ``context_lens = len(seq)``, ``slot = table[-1]*BS + pos``,
  block tables padded with -1 (reference :231-256) — here vectorised with numpy into pinned
  staging arrays and copied with ONE H2D per array, and with the slot of the LAST token computed as
  ``(len-1) % BS`` (the reference's ``last_block_num_tokens - 1`` is -1 at exact multiples, A-4);
* CUDA-graph capture per batch-size bucket and replay with copy-in (reference :267-290, :317-369):
  static metadata buffers, padded rows get ``context_len = 0`` and ``slot = -1`` (the reference
  leaves stale slots, A-5), block-table width fixed at ``ceil(max_model_length / BS)``.

The model around attention (projections, norms, RoPE, MLP, sampler) is out of scope: ``run`` takes
the per-layer ``q, k, v`` the projections would have produced.
"""
from __future__ import annotations

from typing import Dict, List, NamedTuple, Optional, Sequence

import numpy as np
import torch

from .. import _ext
from ..layers.attention import Attention
from ..utils.context import reset_context, set_context


class SeqState:
    """The two facts about a sequence the decode path needs (reference engine/sequence.py:
    ``len(seq)`` and ``seq.block_table``)."""

    __slots__ = ("num_tokens", "block_table")

    def __init__(self, num_tokens: int = 0):
        self.num_tokens = num_tokens
        self.block_table: List[int] = []

    def __len__(self) -> int:
        return self.num_tokens


class BlockPool:
    """Free-list page allocator: the part of reference engine/block_manager.py:99-139 the decode
    loop exercises (allocate for a prompt, append a page when a token crosses a page boundary,
    free).  No prefix hashing here (row N2)."""

    def __init__(self, num_blocks: int, block_size: int, seed: Optional[int] = None):
        order = np.arange(num_blocks)
        if seed is not None:
            np.random.RandomState(seed).shuffle(order)
        self.free: List[int] = order[::-1].tolist()
        self.block_size = block_size

    def allocate(self, seq: SeqState, num_tokens: int) -> None:
        need = (num_tokens + self.block_size - 1) // self.block_size
        if need > len(self.free):
            raise MemoryError("KV-cache pool exhausted")
        seq.block_table = [self.free.pop() for _ in range(need)]
        seq.num_tokens = num_tokens

    def append_token(self, seq: SeqState) -> None:
        """Account for one generated token; takes a new page when the token starts one."""
        if seq.num_tokens % self.block_size == 0:
            if not self.free:
                raise MemoryError("KV-cache pool exhausted")
            seq.block_table.append(self.free.pop())
        seq.num_tokens += 1

    def deallocate(self, seq: SeqState) -> None:
        self.free.extend(reversed(seq.block_table))
        seq.block_table = []
        seq.num_tokens = 0


class PrefillMeta(NamedTuple):
    cu_seqlens_q: np.ndarray   # (n+1) int32 over the NEW tokens of each prompt
    cu_seqlens_k: np.ndarray   # (n+1) int32 over all tokens (cached prefix + new)
    slot_mapping: np.ndarray   # (sum new) int64
    max_seqlen_q: int
    max_seqlen_k: int
    any_cached: bool


def prefill_metadata(seqs: Sequence["SeqState"], block_size: int) -> PrefillMeta:
    """Metadata of a prefill batch — what reference model_runner.py:186-213 builds with per-sequence,
    per-block Python loops, here a repeat + arange: the tokens of a prompt that are not already cached
    (``seq.num_cached_tokens``, set by the block manager, a multiple of the block size) are the
    queries; token ``j`` of a prompt lands in ``table[j // BS] * BS + j % BS``."""
    if not seqs:
        raise ValueError("empty prefill batch")
    BS = block_size
    n = np.fromiter((s.num_tokens for s in seqs), dtype=np.int64, count=len(seqs))
    c = np.fromiter((getattr(s, "num_cached_tokens", 0) for s in seqs), dtype=np.int64, count=len(seqs))
    if (n <= 0).any() or (c < 0).any() or (c >= n).any() or (c % BS).any():
        raise ValueError("prefill needs 0 <= num_cached_tokens < num_tokens, cached in whole blocks")
    n_pages = (n + BS - 1) // BS
    for s, need in zip(seqs, n_pages):
        if len(s.block_table) < need:
            raise ValueError("a sequence's block table is shorter than its prompt")
    nq = n - c
    cu_q = np.zeros(len(seqs) + 1, dtype=np.int32)
    cu_k = np.zeros(len(seqs) + 1, dtype=np.int32)
    np.cumsum(nq, out=cu_q[1:])
    np.cumsum(n, out=cu_k[1:])
    pos = np.arange(int(cu_q[-1]), dtype=np.int64) - np.repeat(cu_q[:-1].astype(np.int64) - c, nq)
    pages = np.concatenate([np.asarray(s.block_table[:need], dtype=np.int64)
                            for s, need in zip(seqs, n_pages)])
    first_page = np.zeros(len(seqs), dtype=np.int64)
    np.cumsum(n_pages[:-1], out=first_page[1:])
    slots = pages[np.repeat(first_page, nq) + pos // BS] * BS + pos % BS
    return PrefillMeta(cu_q, cu_k, slots, int(nq.max()), int(n.max()), bool(c.any()))


class DecodeRunner:
    def __init__(self, num_layers: int, num_heads: int, num_kv_heads: int, head_dim: int,
                 block_size: int, num_blocks: int, max_num_seqs: int, max_model_length: int,
                 dtype: torch.dtype = torch.bfloat16, device: Optional[torch.device] = None,
                 enforce_eager: bool = False):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.L, self.H, self.KVH, self.D = num_layers, num_heads, num_kv_heads, head_dim
        self.block_size, self.num_blocks = block_size, num_blocks
        self.max_bs = max_num_seqs
        self.max_blocks = (max_model_length + block_size - 1) // block_size
        self.enforce_eager = enforce_eager
        self.dtype = dtype
        # (2, L, NB, BS, KVH, D): one allocation, per-layer views (model_runner.py:162-168)
        self.kv_cache = torch.zeros(2, num_layers, num_blocks, block_size, num_kv_heads, head_dim,
                                    dtype=dtype, device=self.device)
        self.layers = torch.nn.ModuleList(
            Attention(num_heads, head_dim, 1.0, num_kv_heads, block_size) for _ in range(num_layers))
        # A layer's decode kernel may start while the previous layer's is finishing: that kernel writes
        # only ITS layer's cache and output (MVL_DECODE_OVERLAP_PREV).  Layer 0 keeps plain stream order,
        # whatever ran in front of the step.
        from ..layers import attention as _att
        for layer in list(self.layers)[1:]:
            layer.decode_flags = _att.DEFAULT_DECODE_FLAGS | _ext.DECODE_OVERLAP_PREV
        for i, layer in enumerate(self.layers):
            layer.k_cache, layer.v_cache = self.kv_cache[0, i], self.kv_cache[1, i]
        # Pinned host staging for the step's metadata: a small ring of sets, each guarded by an event
        # recorded after its H2D copies, because the host runs ahead of the GPU under graph replay and
        # must not rewrite a set whose copies are still pending.
        pin = dict(pin_memory=True)
        self._ring = [dict(slot=torch.empty(self.max_bs, dtype=torch.int64, **pin),
                           ctx=torch.empty(self.max_bs, dtype=torch.int64, **pin),
                           tab=torch.empty(self.max_bs, self.max_blocks, dtype=torch.int32, **pin),
                           done=torch.cuda.Event(), used=False) for _ in range(4)]
        for st in self._ring:
            st["np"] = tuple(st[k].numpy() for k in ("slot", "ctx", "tab"))
        self._ring_at = 0
        self.h2d_bytes_per_step = self.max_bs * (8 + 8 + 4 * self.max_blocks)
        self.d_slot = torch.full((self.max_bs,), -1, dtype=torch.int64, device=self.device)
        self.d_ctx = torch.zeros(self.max_bs, dtype=torch.int64, device=self.device)
        self.d_tab = torch.zeros(self.max_bs, self.max_blocks, dtype=torch.int32, device=self.device)
        # static activations the (out-of-scope) projections would write
        self.q = torch.zeros(num_layers, self.max_bs, num_heads, head_dim, dtype=dtype, device=self.device)
        self.k = torch.zeros(num_layers, self.max_bs, num_kv_heads, head_dim, dtype=dtype, device=self.device)
        self.v = torch.zeros_like(self.k)
        self.out = torch.zeros(num_layers, self.max_bs, num_heads * head_dim, dtype=dtype, device=self.device)
        self.graphs: Dict[int, torch.cuda.CUDAGraph] = {}
        self.graph_bs: List[int] = []
        self._bucket = 0

    # ------------------------------------------------------------------ host prep (vectorised)
    def prepare_decode(self, seqs: Sequence[SeqState]) -> int:
        bs = len(seqs)
        if not 0 < bs <= self.max_bs:
            raise ValueError(f"batch {bs} outside (0, {self.max_bs}]")
        BS = self.block_size
        bucket = bs if self.enforce_eager or not self.graphs else next(b for b in self.graph_bs if b >= bs)
        st = self._ring[self._ring_at]
        self._ring_at = (self._ring_at + 1) % len(self._ring)
        if st["used"]:
            st["done"].synchronize()          # its previous H2D copies have been consumed
        np_slot, np_ctx, np_tab = st["np"]
        n = np.fromiter((s.num_tokens for s in seqs), dtype=np.int64, count=bs)
        last_blk = np.fromiter((s.block_table[-1] for s in seqs), dtype=np.int64, count=bs)
        np_ctx[:bs] = n
        np_ctx[bs:bucket] = 0                 # padded rows: zero-length -> zero output, no page reads
        np_slot[:bs] = last_blk * BS + (n - 1) % BS
        np_slot[bs:bucket] = -1               # padded rows never write the cache
        width = self.max_blocks if not self.enforce_eager else max(len(s.block_table) for s in seqs)
        tab = np_tab[:bucket, :width]
        tab.fill(-1)
        for i, s in enumerate(seqs):
            bt = s.block_table
            tab[i, :len(bt)] = bt
        self.d_slot[:bucket].copy_(st["slot"][:bucket], non_blocking=True)
        self.d_ctx[:bucket].copy_(st["ctx"][:bucket], non_blocking=True)
        if width == self.max_blocks:
            self.d_tab[:bucket].copy_(st["tab"][:bucket], non_blocking=True)
        else:
            self.d_tab[:bucket, :width].copy_(st["tab"][:bucket, :width], non_blocking=True)
        st["done"].record()
        st["used"] = True
        self._bucket, self._width = bucket, width
        set_context(False, slot_mapping=self.d_slot[:bucket], context_lens=self.d_ctx[:bucket],
                    block_tables=self.d_tab[:bucket, :width])
        return bucket

    def prepare_prefill(self, seqs: Sequence[SeqState]) -> int:
        """Metadata of a prefill step (reference model_runner.py:180-227): ``cu_seqlens_q`` over the
        tokens still to compute, ``cu_seqlens_k`` over whole prompts, one slot per new token, and —
        when the block manager found cached prefixes — the block tables, so that attention reads the
        cached pages too (model_runner.py:209-215; the reference's kernel then ignores them, A-8).
        One H2D copy per array.  Returns the number of new tokens."""
        m = prefill_metadata(seqs, self.block_size)
        total = int(m.cu_seqlens_q[-1])
        host = [torch.from_numpy(m.cu_seqlens_q).pin_memory(), torch.from_numpy(m.cu_seqlens_k).pin_memory(),
                torch.from_numpy(m.slot_mapping).pin_memory()]
        tables = None
        if m.any_cached:
            width = max(len(s.block_table) for s in seqs)
            tab = np.full((len(seqs), width), -1, dtype=np.int32)
            for i, s in enumerate(seqs):
                tab[i, :len(s.block_table)] = s.block_table
            host.append(torch.from_numpy(tab).pin_memory())
        dev = [h.to(self.device, non_blocking=True) for h in host]
        if m.any_cached:
            tables = dev[3]
        self._prefill_keep = host   # pinned sources stay alive until the copies ran
        set_context(True, cu_seqlens_q=dev[0], cu_seqlens_k=dev[1], max_seqlen_q=m.max_seqlen_q,
                    max_seqlen_k=m.max_seqlen_k, slot_mapping=dev[2], block_tables=tables)
        return total

    @torch.inference_mode()
    def run_prefill(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        """Every layer's prefill attention for the step prepared by ``prepare_prefill``:
        ``q`` (L, T, H, D), ``k``/``v`` (L, T, KVH, D) as the projections would have produced them.
        Stores K/V into the pages and returns (L, T, H*D).  Eager, as in the reference (graphs are
        for decode only, model_runner.py:292-300)."""
        out = torch.empty(self.L, q.shape[1], self.H * self.D, dtype=q.dtype, device=q.device)
        for i, layer in enumerate(self.layers):
            layer(q[i], k[i], v[i], out=out[i].view(-1, self.H, self.D))
        reset_context()
        return out

    # ------------------------------------------------------------------ the layer loop
    def _forward_layers(self, bucket: int) -> None:
        for i, layer in enumerate(self.layers):
            layer(self.q[i, :bucket], self.k[i, :bucket], self.v[i, :bucket],
                  out=self.out[i, :bucket].view(bucket, self.H, self.D))

    @torch.inference_mode()
    def run(self) -> torch.Tensor:
        """Attention of every layer for the step prepared by ``prepare_decode``; reads the static
        ``q/k/v`` buffers, returns ``out[:, :bucket]`` (L, bucket, H*D)."""
        bucket = self._bucket
        if self.enforce_eager or bucket not in self.graphs:
            self._forward_layers(bucket)
        else:
            self.graphs[bucket].replay()
        reset_context()
        return self.out[:, :bucket]

    @torch.inference_mode()
    def capture_cudagraph(self, batch_sizes: Optional[Sequence[int]] = None) -> None:
        """Capture one graph per batch bucket, largest first, sharing one memory pool
        (reference model_runner.py:317-369)."""
        if batch_sizes is None:
            batch_sizes = [b for b in (1, 2, 4, 8) if b <= self.max_bs] + list(range(16, self.max_bs + 1, 16))
        self.graph_bs = sorted(set(batch_sizes))
        pool = None
        for bs in reversed(self.graph_bs):
            set_context(False, slot_mapping=self.d_slot[:bs], context_lens=self.d_ctx[:bs],
                        block_tables=self.d_tab[:bs])
            self._forward_layers(bs)        # eager warm-up: opts kernels in, sizes the workspace
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, pool):
                self._forward_layers(bs)
            if pool is None:
                pool = graph.pool()
            self.graphs[bs] = graph
            torch.cuda.synchronize(self.device)
        reset_context()
