"""Property test of the prefix-caching block manager (minivllm_b200/engine/block_manager.py, the producer of the
aliased block tables of BASELINE config 5; reference engine/block_manager.py:24-139): random interleavings of
allocate / append / deallocate over a small pool, with prompts drawn from a few common stems so that pages are
shared, revived from the free list and recycled all the time.  After every operation the bookkeeping must add
up and no page that a decode step will WRITE may be visible to another sequence.  CPU only."""
from collections import Counter

from hypothesis import given, settings, strategies as st

from minivllm_b200.engine.block_manager import PrefixBlockManager, PrefixSeq

BS, POOL = 4, 20
STEMS = ([1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13], [1, 2, 3, 4, 9, 9, 9, 9, 7], [5, 5, 5, 5, 5, 5, 5, 5, 5, 5])

op = st.one_of(
    st.tuples(st.just("new"), st.integers(0, len(STEMS) - 1), st.integers(1, 13), st.integers(0, 3)),
    st.tuples(st.just("tok"), st.integers(0, 7), st.integers(0, 2)),
    st.tuples(st.just("del"), st.integers(0, 7)),
)


def check_books(m, live, current=None):
    """``live``: every sequence holding pages.  ``current``: those whose last token has been booked (all of them
    unless given: a generated token is booked by the next ``append``, so between an engine's postprocess and the
    sequence's next decode step its table and hash chain are one token behind by design)."""
    BS, POOL = m.block_size, len(m.blocks)
    current = live if current is None else current
    refs = Counter()
    for s in live:
        assert len(set(s.block_table)) == len(s.block_table)          # a sequence never holds a page twice
        refs.update(s.block_table)
    for s in current:
        assert len(s.block_table) == (len(s) + BS - 1) // BS
        assert len(s.block_hashes) == len(s) // BS                    # the chain over its own full blocks
        if len(s) % BS:                                               # the page the next token's K/V go to
            assert m.blocks[s.block_table[-1]].ref_count == 1
    free = list(m.free_block_ids)
    assert len(set(free)) == len(free)
    for bid, b in enumerate(m.blocks):
        assert b.ref_count == refs.get(bid, 0)
        assert (b.ref_count == 0) == (bid in m.free_block_ids)
    assert len(free) + len(refs) == POOL
    for h, bid in m.hash_to_block_id.items():
        assert m.blocks[bid].hash == h and len(m.blocks[bid].token_ids) == BS
    # a page two sequences share sits at the same index of both tables and stands for the same prefix
    where = {}
    for s in live:
        for i, bid in enumerate(s.block_table):
            if bid in where:
                j, other = where[bid]
                assert i == j and other.token_ids[:(i + 1) * BS] == s.token_ids[:(i + 1) * BS]
            else:
                where[bid] = (i, s)
    # a published page's content is what every holder has at that block
    for s in live:
        for i in range(min(len(s.block_hashes), len(s.block_table))):
            b = m.blocks[s.block_table[i]]
            if b.hash is not None:
                assert b.token_ids == s.token_ids[i * BS:(i + 1) * BS] and b.hash == s.block_hashes[i]


@settings(max_examples=300, deadline=None)
@given(st.lists(op, min_size=1, max_size=60))
def test_random_interleavings_keep_the_books(ops):
    m = PrefixBlockManager(POOL, BS)
    live = []
    for o in ops:
        if o[0] == "new":
            stem, n, salt = STEMS[o[1]], o[2], o[3]
            toks = list(stem[:n]) + [100 + salt] * max(0, n - len(stem))
            s = PrefixSeq(toks)
            if not m.can_allocate(s):
                continue
            expect = m.cached_prefix_tokens(s)
            m.allocate(s)
            assert s.num_cached_tokens == expect                      # the side-effect-free look-up agrees
            assert s.num_cached_tokens % BS == 0 and s.num_cached_tokens < len(s)
            assert m.blocks[s.block_table[-1]].ref_count == 1         # the last token's page is never a cached one
            live.append(s)
        elif o[0] == "tok" and live:
            s = live[o[1] % len(live)]
            if not m.can_append(s) and len(s) % BS == 0:
                continue
            s.append_token(200 + o[2])
            if len(s) % BS == 1 and not m.free_block_ids:
                s.token_ids.pop()
                continue
            m.append(s)
        elif o[0] == "del" and live:
            s = live.pop(o[1] % len(live))
            m.deallocate(s)
            assert s.block_table == [] and s.num_cached_tokens == 0
        check_books(m, live)
    for s in live:
        m.deallocate(s)
    check_books(m, [])
    assert len(m.free_block_ids) == POOL
