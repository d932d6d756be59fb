"""Continuous-batching scheduler and engine step at the attention-path level.

``Scheduler`` restates reference engine/scheduler.py:6-82 — prefill first, FCFS, under
``max_num_sequences`` / ``max_num_batched_tokens``; otherwise one decode token for every running
sequence, preempting the most recently admitted one when the pool has no page for a token that opens
a block; ``postprocess`` appends the generated tokens and retires sequences at EOS / ``max_tokens`` —
over ``PrefixBlockManager`` and with the repairs without which the reference loop cannot advance
(SURVEY.md Appendix A-3): ``postprocess`` goes through ``append_token`` (the reference appends to
``token_ids`` only, so lengths, slots and block tables never move, scheduler.py:78), a preempted
sequence is prefilled again with everything it has generated so far, and a prompt's admission counts
only its uncached tokens against the token budget.

``AttentionEngine.step`` is reference engine/llm_engine.py:48-60 restricted to the path this repo
owns: schedule -> prepare_prefill / prepare_decode -> the L attention layers -> postprocess.  The
model around attention is out of scope, so the caller supplies two callables standing in for it:
``activations(token_ids, positions) -> (q, k, v)`` (what embedding + projections would produce, per
layer) and ``sample(seqs, attn_out) -> token ids`` (what o_proj ... lm_head + sampler would).
"""
from __future__ import annotations

from collections import deque
from typing import Callable, Deque, List, Sequence, Tuple

import torch

from .block_manager import PrefixBlockManager, PrefixSeq
from .decode_runner import DecodeRunner

WAITING, RUNNING, FINISHED = "waiting", "running", "finished"


class EngineSeq(PrefixSeq):
    """reference engine/sequence.py:15-60: prompt + completion tokens, sampling limits, status."""

    __slots__ = ("seq_id", "status", "num_prompt_tokens", "max_tokens", "ignore_eos")
    _counter = 0

    def __init__(self, token_ids: Sequence[int], max_tokens: int = 64, ignore_eos: bool = False):
        super().__init__(token_ids)
        self.seq_id = EngineSeq._counter
        EngineSeq._counter += 1
        self.status = WAITING
        self.num_prompt_tokens = len(self.token_ids)
        self.max_tokens = max_tokens
        self.ignore_eos = ignore_eos

    @property
    def num_completion_tokens(self) -> int:
        return len(self.token_ids) - self.num_prompt_tokens

    @property
    def completion_token_ids(self) -> List[int]:
        return self.token_ids[self.num_prompt_tokens:]


class Scheduler:
    def __init__(self, max_num_sequences: int, max_num_batched_tokens: int, max_cached_blocks: int,
                 block_size: int, eos: int):
        self.block_manager = PrefixBlockManager(max_cached_blocks, block_size)
        self.max_num_sequences = max_num_sequences
        self.max_num_batched_tokens = max_num_batched_tokens
        self.eos = eos
        self.waiting: Deque[EngineSeq] = deque()
        self.running: Deque[EngineSeq] = deque()
        self.num_preemptions = 0

    def is_finished(self) -> bool:
        return not self.waiting and not self.running

    def add_sequence(self, seq: EngineSeq) -> None:
        self.waiting.append(seq)

    def schedule(self) -> Tuple[List[EngineSeq], bool]:
        """(sequences of this step, is_prefill)."""
        scheduled: List[EngineSeq] = []
        tokens = 0
        while self.waiting and len(self.running) < self.max_num_sequences:
            seq = self.waiting[0]
            if not self.block_manager.can_allocate(seq):
                break
            new_tokens = len(seq) - self.block_manager.cached_prefix_tokens(seq)
            if scheduled and tokens + new_tokens > self.max_num_batched_tokens:
                break                                   # does not fit this step: stays at the head
            self.block_manager.allocate(seq)
            self.waiting.popleft()
            seq.status = RUNNING
            self.running.append(seq)
            scheduled.append(seq)
            tokens += new_tokens
        if scheduled:
            return scheduled, True

        # decode: one token for every running sequence, oldest first
        kept: List[EngineSeq] = []
        while self.running:
            seq = self.running.popleft()
            while not self.block_manager.can_append(seq):
                if self.running:
                    self.preempt(self.running.pop())     # youngest first
                else:
                    self.preempt(seq)
                    seq = None
                    break
            if seq is None:
                break
            self.block_manager.append(seq)
            kept.append(seq)
        self.running.extend(kept)
        if not kept and self.waiting and self.block_manager.can_allocate(self.waiting[0]):
            return self.schedule()   # the only running sequence preempted itself: prefill it again now
        return kept, False

    def preempt(self, seq: EngineSeq) -> None:
        self.block_manager.deallocate(seq)
        seq.status = WAITING
        self.waiting.appendleft(seq)
        self.num_preemptions += 1

    def postprocess(self, seqs: Sequence[EngineSeq], token_ids: Sequence[int]) -> None:
        for seq, tok in zip(seqs, token_ids):
            seq.append_token(int(tok))
            if (not seq.ignore_eos and tok == self.eos) or seq.num_completion_tokens >= seq.max_tokens:
                seq.status = FINISHED
                self.block_manager.deallocate(seq)
                self.running.remove(seq)


class AttentionEngine:
    def __init__(self, runner: DecodeRunner, scheduler: Scheduler,
                 activations: Callable[[List[int], List[int]], Tuple[torch.Tensor, torch.Tensor, torch.Tensor]],
                 sample: Callable[[Sequence[EngineSeq], torch.Tensor], Sequence[int]]):
        assert runner.block_size == scheduler.block_manager.block_size
        self.runner, self.scheduler = runner, scheduler
        self.activations, self.sample = activations, sample
        self.prefill_tokens = self.decode_tokens = 0

    def add_prompt(self, token_ids: Sequence[int], max_tokens: int = 64, ignore_eos: bool = False) -> EngineSeq:
        seq = EngineSeq(token_ids, max_tokens, ignore_eos)
        self.scheduler.add_sequence(seq)
        return seq

    @torch.inference_mode()
    def step(self) -> Tuple[List[EngineSeq], bool, torch.Tensor]:
        """One engine step.  Returns (sequences, is_prefill, attention output of each sequence's LAST
        token, (L, len(sequences), H*D)) — the rows the sampler would see."""
        seqs, is_prefill = self.scheduler.schedule()
        if not seqs:
            raise RuntimeError("nothing to schedule: the KV pool cannot hold a single waiting sequence")
        r = self.runner
        if is_prefill:
            toks: List[int] = []
            pos: List[int] = []
            last: List[int] = []
            for s in seqs:
                c = s.num_cached_tokens
                toks.extend(s.token_ids[c:])
                pos.extend(range(c, len(s)))
                last.append(len(toks) - 1)
            q, k, v = self.activations(toks, pos)              # (L, T, heads, D) each
            r.prepare_prefill(seqs)
            out = r.run_prefill(q, k, v)[:, torch.tensor(last, device=q.device)]
            self.prefill_tokens += len(toks)
        else:
            # the token appended by the last postprocess is this step's input; its K/V go to position len-1
            q, k, v = self.activations([s.token_ids[-1] for s in seqs], [len(s) - 1 for s in seqs])
            bucket = r.prepare_decode(seqs)
            n = len(seqs)
            r.q[:, :n].copy_(q), r.k[:, :n].copy_(k), r.v[:, :n].copy_(v)
            out = r.run()[:, :n].clone()
            assert bucket >= n
            self.decode_tokens += n
        self.scheduler.postprocess(seqs, self.sample(seqs, out))
        return seqs, is_prefill, out

    def run_to_completion(self, max_steps: int = 1 << 30) -> int:
        steps = 0
        while not self.scheduler.is_finished() and steps < max_steps:
            self.step()
            steps += 1
        return steps
