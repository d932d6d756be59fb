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


class PeerGatheredDecode:
    """Head-sharded decode whose result stores ARE the all-gather (``mvl_paged_decode_peers``).

    Every rank owns a (batch, total_heads, head_dim) result buffer in symmetric (peer-mapped)
    memory.  A rank's decode kernels write its heads straight into all ranks' buffers over NVLink
    P2P; one barrier on the symmetric-memory signal pads then makes every buffer complete.  Compared
    with ``gather_heads`` this removes the NCCL all-gather launch and the (N,B,h,D)->(B,H,D)
    permute copy from the step.  ``torch.distributed._symmetric_memory`` is used for the plumbing
    only (allocation, handle exchange, barrier); the data path is our kernels' stores.
    """

    def __init__(self, shard: HeadShard, batch: int, total_heads: int, head_dim: int,
                 dtype: torch.dtype, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        import ctypes

        import torch.distributed._symmetric_memory as symm_mem

        self.shard, self.batch, self.total_heads, self.head_dim = shard, batch, total_heads, head_dim
        self.group = group if group is not None else dist.group.WORLD
        self.out = symm_mem.empty((batch, total_heads, head_dim), dtype=dtype, device=device)
        self.handle = symm_mem.rendezvous(self.out, self.group)
        ptrs = list(self.handle.buffer_ptrs)
        if len(ptrs) != shard.world_size:
            raise RuntimeError("symmetric-memory rendezvous returned the wrong number of peers")
        self._ptr_array = (ctypes.c_void_p * len(ptrs))(*ptrs)
        self.device = device

    def __call__(self, query, k_cache, v_cache, block_tables, context_lens, scale, block_size,
                 chunk_tokens: int = 0, flags: Optional[int] = None, barrier: bool = True,
                 store=None) -> torch.Tensor:
        """Decode this rank's heads into every rank's buffer; returns this rank's full
        (batch, total_heads, head_dim) buffer (complete after the barrier).  ``store=(key, value,
        slot_mapping)`` fuses the step's KV-cache write into the same launch (mvl_decode_step_peers)."""
        import ctypes

        from . import _ext
        from .layers import attention as att

        sh = self.shard
        if flags is None:
            flags = att.DEFAULT_DECODE_FLAGS
        batch, max_blocks = query.shape[0], block_tables.shape[1]
        code = att._dtype_code(query)
        lib = _ext.lib()
        with att._on_device(self.device):
            nbytes = lib.mvl_paged_decode_workspace_bytes(batch, sh.num_heads, sh.num_kv_heads,
                                                          self.head_dim, max_blocks, block_size, code,
                                                          chunk_tokens)
            ws = att.workspace_for(self.device, nbytes)
            peers = ctypes.cast(self._ptr_array, ctypes.c_void_p)
            wsp, wsn = (ws.data_ptr(), ws.numel()) if ws is not None else (None, 0)
            idx = att._index_flag("context_lens", context_lens)
            if store is None:
                _ext.check(lib.mvl_paged_decode_peers(
                    peers, sh.world_size, self.total_heads, sh.q_heads.start, query.data_ptr(),
                    k_cache.data_ptr(), v_cache.data_ptr(), block_tables.data_ptr(),
                    context_lens.data_ptr(), idx, batch, max_blocks, block_tables.stride(0),
                    sh.num_heads, sh.num_kv_heads, self.head_dim, block_size, k_cache.shape[0],
                    float(scale), code, wsp, wsn, chunk_tokens, flags, att._stream_ptr(self.device)))
            else:
                key, value, slots = store
                _ext.check(lib.mvl_decode_step_peers(
                    peers, sh.world_size, self.total_heads, sh.q_heads.start, query.data_ptr(),
                    key.data_ptr(), value.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                    slots.data_ptr(), block_tables.data_ptr(), context_lens.data_ptr(), idx, batch,
                    max_blocks, block_tables.stride(0), key.stride(0), value.stride(0), sh.num_heads,
                    sh.num_kv_heads, self.head_dim, block_size, k_cache.shape[0], float(scale), code,
                    wsp, wsn, chunk_tokens, flags, att._stream_ptr(self.device)))
            if barrier:
                self.barrier()
        return self.out

    def barrier(self) -> None:
        """All ranks' stores issued before this point are visible in every buffer after it."""
        self.handle.barrier(channel=0)
