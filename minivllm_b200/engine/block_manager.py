"""Block-table producer with prefix caching.

The part of the reference's ``BlockManager`` (engine/block_manager.py:24-139) that decides which
physical page every 16/256-token block of a sequence lives in, and therefore WHY block tables can
alias (BASELINE config 5): full blocks are keyed by a hash chained over the whole prefix
(``compute_hash``, :39-44), a new sequence whose leading full blocks are already resident shares
those pages (ref-counted) and only its remaining tokens are prefilled (``num_cached_tokens``,
:67-97); pages return to the free list when their last user leaves (:99-108); a generated token that
opens a block takes a page, one that fills a block publishes the block's hash (:120-139).

Same method names and call protocol as the reference (``can_allocate / allocate / deallocate /
can_append / append``).  Repairs, each where the reference cannot run as written:
* a newly taken page starts with ``ref_count = 1`` (the reference resets it to 0 and later
  decrements to -1, so shared pages are never freed, :46-52,99-105);
* the block holding the prompt's LAST token is never taken from the cache, so a prefill always has at
  least one query (a fully cached prompt leaves the reference with an empty prefill batch);
* a page that is recycled for other content drops its stale hash entry; free pages are popped from a
  deque in O(1) (the reference calls ``deque.remove``, O(free pages), per block);
* the 64-bit key is the reference's own (:38-43): xxh64 over (previous key as 8 little-endian bytes,
  int32 token ids), so keys computed here and by the reference agree; any collision is caught by the
  token comparison the reference also does (:76).
"""
from __future__ import annotations

from collections import deque
from typing import Dict, List, Optional, Sequence

import numpy as np
import xxhash


class PrefixSeq:
    """What the manager needs to know about a sequence (reference engine/sequence.py:
    ``token_ids``, ``num_tokens``, ``block_table``, ``num_cached_tokens``, ``num_blocks``,
    ``block(i)``)."""

    __slots__ = ("token_ids", "block_table", "num_cached_tokens", "block_hashes")

    def __init__(self, token_ids: Sequence[int]):
        self.token_ids: List[int] = list(token_ids)
        self.block_table: List[int] = []
        self.num_cached_tokens = 0
        self.block_hashes: List[int] = []   # chained keys of the sequence's full blocks, filled by the manager

    @property
    def num_tokens(self) -> int:
        return len(self.token_ids)

    def __len__(self) -> int:
        return len(self.token_ids)

    def append_token(self, token_id: int) -> None:
        self.token_ids.append(token_id)


class _Block:
    __slots__ = ("hash", "ref_count", "token_ids")

    def __init__(self):
        self.hash: Optional[int] = None
        self.ref_count = 0
        self.token_ids: List[int] = []


class PrefixBlockManager:
    def __init__(self, num_blocks: int, block_size: int):
        self.block_size = block_size
        self.blocks = [_Block() for _ in range(num_blocks)]
        self.hash_to_block_id: Dict[int, int] = {}
        self.free_block_ids = deque(range(num_blocks))
        self.hits = 0          # blocks served from the cache (statistics)

    # ------------------------------------------------------------------ hashing
    @staticmethod
    def compute_hash(token_ids: Sequence[int], prefix_hash: Optional[int]) -> int:
        h = xxhash.xxh64()
        if prefix_hash is not None:
            h.update(prefix_hash.to_bytes(8, "little"))
        h.update(np.asarray(token_ids, dtype=np.int32).tobytes())
        return h.intdigest()

    # ------------------------------------------------------------------ pages
    def _take(self, block_id: Optional[int] = None) -> int:
        """Take a page off the free list (a specific one when reviving a cached page)."""
        if block_id is None:
            block_id = self.free_block_ids.popleft()
        else:
            self.free_block_ids.remove(block_id)
        b = self.blocks[block_id]
        assert b.ref_count == 0
        b.ref_count = 1
        return block_id

    def _recycle(self, block_id: int) -> None:
        """A page about to hold new content forgets what it used to cache."""
        b = self.blocks[block_id]
        if b.hash is not None and self.hash_to_block_id.get(b.hash) == block_id:
            del self.hash_to_block_id[b.hash]
        b.hash, b.token_ids = None, []

    def _blocks_needed(self, n_tokens: int) -> int:
        return (n_tokens + self.block_size - 1) // self.block_size

    # ------------------------------------------------------------------ reference protocol
    def can_allocate(self, seq: PrefixSeq) -> bool:
        return len(self.free_block_ids) >= self._blocks_needed(len(seq))

    def cached_prefix_tokens(self, seq: PrefixSeq) -> int:
        """How many leading tokens ``allocate`` would serve from the cache — a look-up with no side
        effects (the scheduler budgets a prefill step by the tokens that still have to be computed)."""
        bs, n = self.block_size, len(seq)
        h: Optional[int] = None
        cached = 0
        for i in range(self._blocks_needed(n)):
            tokens = seq.token_ids[i * bs:(i + 1) * bs]
            if len(tokens) < bs or (i + 1) * bs >= n:
                break
            h = self.compute_hash(tokens, h)
            cand = self.hash_to_block_id.get(h, -1)
            if cand == -1 or self.blocks[cand].token_ids != tokens:
                break
            cached += bs
        return cached

    def allocate(self, seq: PrefixSeq) -> None:
        """Give every block of the prompt a page; leading full blocks found in the cache are shared
        (``seq.num_cached_tokens`` counts their tokens)."""
        assert not seq.block_table
        bs, n = self.block_size, len(seq)
        h: Optional[int] = None
        cached_run = True
        seq.block_hashes = []
        for i in range(self._blocks_needed(n)):
            tokens = seq.token_ids[i * bs:(i + 1) * bs]
            full = len(tokens) == bs
            h = self.compute_hash(tokens, h) if full else None
            if full:
                seq.block_hashes.append(h)
            hit = -1
            # only an unbroken run of leading blocks can be reused, and never the block of the last token
            if cached_run and full and (i + 1) * bs < n:
                cand = self.hash_to_block_id.get(h, -1)
                if cand != -1 and self.blocks[cand].token_ids == tokens:
                    hit = cand
            if hit != -1:
                b = self.blocks[hit]
                if b.ref_count == 0:
                    self._take(hit)          # still resident on the free list: revive it
                else:
                    b.ref_count += 1
                seq.num_cached_tokens += bs
                self.hits += 1
                seq.block_table.append(hit)
                continue
            cached_run = False
            if not self.free_block_ids:
                raise MemoryError("KV-cache pool exhausted")
            bid = self._take()
            self._recycle(bid)
            if full:
                known = self.hash_to_block_id.get(h, -1)
                if known == -1 or self.blocks[known].token_ids != tokens:   # keep an existing valid entry
                    b = self.blocks[bid]
                    b.hash, b.token_ids = h, tokens
                    self.hash_to_block_id[h] = bid
            seq.block_table.append(bid)

    def deallocate(self, seq: PrefixSeq) -> None:
        for bid in reversed(seq.block_table):
            b = self.blocks[bid]
            b.ref_count -= 1
            assert b.ref_count >= 0
            if b.ref_count == 0:
                self.free_block_ids.append(bid)   # content and hash stay: a later prompt may revive it
        seq.block_table = []
        seq.num_cached_tokens = 0
        seq.block_hashes = []

    def can_append(self, seq: PrefixSeq) -> bool:
        return len(seq) % self.block_size != 1 or len(self.free_block_ids) > 0

    def append(self, seq: PrefixSeq) -> None:
        """Called after a generated token was added to ``seq``: open a page when the token starts a
        block, publish the block's hash when it completes one."""
        bs, n = self.block_size, len(seq)
        if n % bs == 1 or bs == 1:
            if len(seq.block_table) < self._blocks_needed(n):
                if not self.free_block_ids:
                    raise MemoryError("KV-cache pool exhausted")
                bid = self._take()
                self._recycle(bid)
                seq.block_table.append(bid)
        if n % bs == 0:
            i = n // bs - 1
            tokens = seq.token_ids[i * bs:(i + 1) * bs]
            assert len(seq.block_hashes) == i      # the chain over the sequence's own tokens
            h = self.compute_hash(tokens, seq.block_hashes[-1] if i > 0 else None)
            seq.block_hashes.append(h)
            b = self.blocks[seq.block_table[i]]
            known = self.hash_to_block_id.get(h, -1)
            if b.ref_count == 1 and (known == -1 or self.blocks[known].token_ids != tokens):
                b.hash, b.token_ids = h, tokens    # private page, content not cached yet: publish
                self.hash_to_block_id[h] = seq.block_table[i]
