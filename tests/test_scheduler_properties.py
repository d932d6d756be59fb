"""Property test of the continuous-batching scheduler over the prefix-caching block manager
(minivllm_b200/engine/scheduler.py; reference engine/scheduler.py:6-82 with the repairs of SURVEY.md A-3):
random prompts (common stems, so pages are shared), random completion limits and a pool small enough to force
preemptions.  Every step keeps the limits and the books, nobody is lost or duplicated, a preempted sequence keeps
what it has generated, and the loop always drains — the property the reference's own loop does not have.  CPU only."""
import random

from hypothesis import given, settings, strategies as st

from minivllm_b200.engine.scheduler import FINISHED, RUNNING, WAITING, EngineSeq, Scheduler
from test_block_manager_properties import check_books

BS, EOS = 4, 0
STEMS = ([1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17], [1, 2, 3, 4, 5, 6, 7, 8, 3, 3, 3, 3, 3], [9] * 11)

prompt = st.tuples(st.integers(0, len(STEMS) - 1), st.integers(1, 17), st.integers(1, 9), st.booleans())


@settings(max_examples=150, deadline=None)
@given(st.lists(prompt, min_size=1, max_size=12), st.integers(7, 16), st.integers(1, 5), st.integers(6, 40),
       st.integers(0, 2 ** 31 - 1))
def test_scheduler_drains_and_keeps_its_limits(prompts, pool, max_seqs, max_tokens_step, seed):
    rng = random.Random(seed)
    sch = Scheduler(max_seqs, max_tokens_step, pool, BS, EOS)
    seqs = []
    for stem, n, limit, ignore in prompts:
        # every sequence alone must fit the pool at its longest, or no scheduler could finish it
        n = min(n, (pool - 1) * BS - limit)
        if n < 1:
            continue
        s = EngineSeq(list(STEMS[stem][:n]), max_tokens=limit, ignore_eos=ignore)
        seqs.append(s)
        sch.add_sequence(s)
    steps = 0
    while not sch.is_finished():
        steps += 1
        assert steps < 4000, "the loop does not drain"
        before = {id(s): list(s.token_ids) for s in seqs}
        batch, is_prefill = sch.schedule()
        assert batch, "nothing scheduled although sequences are pending"
        assert len({id(s) for s in batch}) == len(batch)
        assert len(sch.running) <= max_seqs
        assert all(list(s.token_ids) == before[id(s)] for s in seqs)          # scheduling never edits tokens
        for s in batch:
            assert s.status == RUNNING and s in sch.running
            assert len(s.block_table) == (len(s) + BS - 1) // BS
        if is_prefill:
            new = [len(s) - s.num_cached_tokens for s in batch]
            assert sum(new) <= max_tokens_step or len(batch) == 1 or sum(new[:-1]) <= max_tokens_step
            assert all(n >= 1 for n in new)                                   # a prefill is never empty
        else:
            assert [id(s) for s in batch] == [id(s) for s in sch.running]     # every running sequence, once, oldest first
        waiting, running = {id(s) for s in sch.waiting}, {id(s) for s in sch.running}
        assert not (waiting & running)
        for s in seqs:
            if s.status == WAITING:
                assert id(s) in waiting and not s.block_table
            elif s.status == FINISHED:
                assert id(s) not in waiting | running and not s.block_table
        check_books(sch.block_manager, list(sch.running), current=batch)
        toks = [EOS if rng.random() < 0.15 else rng.randint(1, 20) for _ in batch]
        sch.postprocess(batch, toks)
        for s, t in zip(batch, toks):
            assert s.token_ids[-1] == t and s.num_completion_tokens <= s.max_tokens
            done = (t == EOS and not s.ignore_eos) or s.num_completion_tokens >= s.max_tokens
            assert (s.status == FINISHED) == done
        # (the new token's page and hash are booked by the next schedule(): the books are checked there)
    assert all(s.status == FINISHED for s in seqs)
    assert len(sch.block_manager.free_block_ids) == pool
    for s in seqs:
        assert 1 <= s.num_completion_tokens <= s.max_tokens
        if not s.ignore_eos and EOS in s.completion_token_ids:
            assert s.completion_token_ids.index(EOS) == s.num_completion_tokens - 1
