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
    """Head-sharded decode whose result stores ARE the all-gather (``mvl_decode_step_peers_sync``).

    Every rank owns TWO (batch, total_heads, head_dim) result buffers and a 64-byte control block in
    symmetric (peer-mapped) memory.  A rank's decode kernels write its heads straight into all ranks'
    buffers over NVLink P2P and then announce the step in every rank's control block; ``wait()``
    enqueues a one-warp kernel that returns once all ranks have announced this rank's current step.
    Compared with ``gather_heads`` this removes the NCCL all-gather launch, the (N,B,h,D)->(B,H,D)
    permute copy and any barrier collective from the step.

    The two buffers are used in turns, so a fast rank's step t+1 stores never land in the buffer a
    slow rank is still reading for step t; a rank reaches step t+2 (the next writer of step t's
    buffer) only behind its own wait of step t+1, which every rank signals after its step-t consumer
    in stream order.  Contract for the caller: every rank makes the same sequence of calls, and the
    tensor returned for step t is valid from that step's wait until this rank's call of step t+2.

    ``torch.distributed._symmetric_memory`` is used for the plumbing only (allocation, handle
    exchange); the data path and the signalling are this library's kernels.
    """

    def __init__(self, shard: HeadShard, batch: int, total_heads: int, head_dim: int,
                 dtype: torch.dtype, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        import ctypes

        import torch.distributed._symmetric_memory as symm_mem

        self.shard, self.batch, self.total_heads, self.head_dim = shard, batch, total_heads, head_dim
        self.group = group if group is not None else dist.group.WORLD
        self.bufs = symm_mem.empty((2, batch, total_heads, head_dim), dtype=dtype, device=device)
        self.handle = symm_mem.rendezvous(self.bufs, self.group)
        self.ctrl = symm_mem.empty((16,), dtype=torch.int32, device=device)
        self.ctrl.zero_()
        self.ctrl_handle = symm_mem.rendezvous(self.ctrl, self.group)
        ptrs = list(self.handle.buffer_ptrs)
        cptrs = list(self.ctrl_handle.buffer_ptrs)
        if len(ptrs) != shard.world_size or len(cptrs) != shard.world_size:
            raise RuntimeError("symmetric-memory rendezvous returned the wrong number of peers")
        half = batch * total_heads * head_dim * self.bufs.element_size()
        self._ptr_arrays = [(ctypes.c_void_p * len(ptrs))(*[p + par * half for p in ptrs]) for par in (0, 1)]
        self._ctrl_array = (ctypes.c_void_p * len(cptrs))(*cptrs)
        self.device = device
        self.step = 0
        # every control block is zero before any rank can write a step count into it
        torch.cuda.synchronize(device)
        dist.barrier(self.group)

    @property
    def out(self) -> torch.Tensor:
        """The buffer of the most recent step (of step 0 before any call)."""
        return self.bufs[(self.step - 1) & 1 if self.step else 0]

    def __call__(self, query, k_cache, v_cache, block_tables, context_lens, scale, block_size,
                 chunk_tokens: int = 0, flags: Optional[int] = None, barrier: bool = True,
                 store=None) -> torch.Tensor:
        """Decode this rank's heads into every rank's buffer of this step; returns this rank's full
        (batch, total_heads, head_dim) buffer (complete after ``wait()``, which ``barrier=True``
        enqueues).  ``store=(key, value, slot_mapping)`` fuses the step's KV-cache write into the
        same launch."""
        import ctypes

        from . import _ext
        from .layers import attention as att

        sh = self.shard
        if flags is None:
            flags = att.DEFAULT_DECODE_FLAGS
        batch, max_blocks = query.shape[0], block_tables.shape[1]
        code = att._dtype_code(query)
        lib = _ext.lib()
        par = self.step & 1
        with att._on_device(self.device):
            nbytes = lib.mvl_paged_decode_workspace_bytes(batch, sh.num_heads, sh.num_kv_heads,
                                                          self.head_dim, max_blocks, block_size, code,
                                                          chunk_tokens)
            ws = att.workspace_for(self.device, nbytes)
            peers = ctypes.cast(self._ptr_arrays[par], ctypes.c_void_p)
            ctrls = ctypes.cast(self._ctrl_array, ctypes.c_void_p)
            wsp, wsn = (ws.data_ptr(), ws.numel()) if ws is not None else (None, 0)
            idx = att._index_flag("context_lens", context_lens)
            key, value, slots = store if store is not None else (None, None, None)
            _ext.check(lib.mvl_decode_step_peers_sync(
                peers, ctrls, sh.world_size, sh.rank, self.total_heads, sh.q_heads.start, query.data_ptr(),
                key.data_ptr() if key is not None else None, value.data_ptr() if value is not None else None,
                k_cache.data_ptr(), v_cache.data_ptr(), slots.data_ptr() if slots is not None else None,
                block_tables.data_ptr(), context_lens.data_ptr(), idx, batch, max_blocks,
                block_tables.stride(0), key.stride(0) if key is not None else 0,
                value.stride(0) if value is not None else 0, sh.num_heads, sh.num_kv_heads, self.head_dim,
                block_size, k_cache.shape[0], float(scale), code, wsp, wsn, chunk_tokens, flags,
                att._stream_ptr(self.device)))
            self.step += 1
            if barrier:
                self.wait()
        return self.bufs[par]

    def wait(self, timeout_ms: int = 0) -> None:
        """Enqueue the wait for the current step on this device's current stream: work enqueued after
        it sees every rank's rows of the step in this rank's buffer."""
        from . import _ext
        from .layers import attention as att
        with att._on_device(self.device):
            _ext.check(_ext.lib().mvl_peer_wait(self.ctrl.data_ptr(), self.shard.world_size, timeout_ms,
                                                att._stream_ptr(self.device)))

    barrier = wait  # earlier name

    def timed_out(self) -> int:
        """0, or 1 + the rank a wait gave up on (reads the control block: synchronises)."""
        return int(self.ctrl[10].item())
