"""KV-head sharding of the decode path over the GPUs of one node, and the output all-gather.

The reference shards attention Megatron-style: rank r owns query heads [r*H/N, (r+1)*H/N) and KV
heads [r*KVH/N, (r+1)*KVH/N) (reference src/myvllm/layers/linear.py:163-165), its KV cache holds
KVH/N heads (model_runner.py:151,162), block tables / context lengths / slot mapping are
replicated, and attention itself needs no communication (qwen3.py:110-111).  A (sequence, KV head)
pair is an independent unit of work, so the same kernels run unchanged on each rank's slice.

The one exchange step is after attention, where a consumer needs all heads: ``gather_heads``
all-gathers the per-rank outputs (B, H/N, D) over NCCL and returns the (B, H, D) layout the
single-GPU operator produces (BASELINE.json north_star: "NCCL all-gather over NVLink only where
the downstream o_proj needs full heads").  One process per GPU, ``torch.distributed`` for the
plumbing; on CPU the same code runs over gloo (tests/test_head_shard_gloo.py).
"""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch
import torch.distributed as dist


@dataclasses.dataclass(frozen=True)
class HeadShard:
    rank: int
    world_size: int
    q_heads: slice    # this rank's query heads
    kv_heads: slice   # this rank's KV heads

    @property
    def num_heads(self) -> int:
        return self.q_heads.stop - self.q_heads.start

    @property
    def num_kv_heads(self) -> int:
        return self.kv_heads.stop - self.kv_heads.start


def shard_plan(num_heads: int, num_kv_heads: int, world_size: int, rank: int) -> HeadShard:
    """Contiguous, equal head ranges per rank.  KV heads are never replicated: world_size must
    divide num_kv_heads (8 KV heads -> at most 8 ranks)."""
    if num_heads % num_kv_heads:
        raise ValueError("num_heads must be a multiple of num_kv_heads")
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank {rank} / world_size {world_size}")
    if num_kv_heads % world_size:
        raise ValueError(f"world_size {world_size} does not divide num_kv_heads {num_kv_heads}")
    kv = num_kv_heads // world_size
    g = num_heads // num_kv_heads
    return HeadShard(rank, world_size, slice(rank * kv * g, (rank + 1) * kv * g),
                     slice(rank * kv, (rank + 1) * kv))


def gather_heads(o_local: torch.Tensor, group: Optional[dist.ProcessGroup] = None,
                 staging: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All-gather per-rank attention outputs (B, H/N, D) into (B, H, D) on every rank.

    NCCL concatenates on the outermost dim, the head dim is the middle one: gather into
    (N, B, H/N, D) (``staging``, reused across steps when given) and permute-copy once.
    """
    world = dist.get_world_size(group)
    if world == 1:
        return o_local
    B, h, D = o_local.shape
    if staging is None:
        staging = torch.empty((world, B, h, D), dtype=o_local.dtype, device=o_local.device)
    dist.all_gather_into_tensor(staging.view(world * B, h, D), o_local.contiguous(), group=group)
    return staging.permute(1, 0, 2, 3).reshape(B, world * h, D)
