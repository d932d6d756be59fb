"""CPU, world_size 2 over gloo: the head-shard plan and the all-gather layout.

Each rank evaluates ITS heads of a small golden case with the CPU oracle (the CUDA kernels are
covered by the -m gpu tests), gathers, and every rank must hold the unsharded result bit for
bit — the same check the GPU box runs with the CUDA operator (tests/test_gpu_multi.py).
"""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from minivllm_b200.head_shard import gather_heads, shard_plan


def test_shard_plan_ranges():
    plans = [shard_plan(32, 8, 4, r) for r in range(4)]
    assert [p.kv_heads for p in plans] == [slice(0, 2), slice(2, 4), slice(4, 6), slice(6, 8)]
    assert [p.q_heads for p in plans] == [slice(0, 8), slice(8, 16), slice(16, 24), slice(24, 32)]
    assert plans[0].num_heads == 8 and plans[0].num_kv_heads == 2
    with pytest.raises(ValueError):
        shard_plan(32, 8, 3, 0)
    with pytest.raises(ValueError):
        shard_plan(32, 8, 2, 2)


def _worker(rank, world, port, ret):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from cases import make_inputs
    from oracle import c_oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        inp = make_inputs("ragged_g4")  # H=8, KVH=2
        full = c_oracle.paged_decode(inp["q"], inp["k_cache"], inp["v_cache"], inp["block_tables"],
                                     inp["context_lens"], inp["scale"], inp["H"], inp["KVH"], inp["D"],
                                     inp["BS"])
        sh = shard_plan(inp["H"], inp["KVH"], world, rank)
        local = c_oracle.paged_decode(
            inp["q"][:, sh.q_heads].contiguous(), inp["k_cache"][:, :, sh.kv_heads].contiguous(),
            inp["v_cache"][:, :, sh.kv_heads].contiguous(), inp["block_tables"], inp["context_lens"],
            inp["scale"], sh.num_heads, sh.num_kv_heads, inp["D"], inp["BS"])
        got = gather_heads(local)
        ret[rank] = bool(torch.equal(got, full)) and got.shape == full.shape
    finally:
        dist.destroy_process_group()


def test_gathered_shards_equal_unsharded():
    world = 2
    port = 29500 + os.getpid() % 2000
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {0: True, 1: True}
