"""GPU (-m gpu): the engine-shaped decode loop — several decode steps over
L layers, eager and CUDA-graph replay, against a CPU simulation that stores then attends with the
oracle (order of reference attention.py:498-510 then :544)."""
import numpy as np
import pytest
import torch

from minivllm_b200.engine import BlockPool, DecodeRunner, SeqState
from oracle import c_oracle
from oracle import paged_attention_oracle as po

pytestmark = pytest.mark.gpu
L, H, KVH, D, NB, MAXBS, MAXLEN = 3, 8, 2, 128, 96, 8, 256


def _make(block_size, eager, device):
    r = DecodeRunner(L, H, KVH, D, block_size, NB, MAXBS, MAXLEN, torch.bfloat16, device, enforce_eager=eager)
    if not eager:
        r.capture_cudagraph([2, 4, 8])
    return r


@pytest.mark.parametrize("block_size", [16, 64])
def test_decode_steps_eager_graph_oracle(cuda_device, block_size):
    g = torch.Generator().manual_seed(31)
    lens = [block_size - 1, block_size, 2 * block_size + 3, 1, 70]   # incl. exact page multiples
    runners = {"eager": _make(block_size, True, cuda_device), "graph": _make(block_size, False, cuda_device)}
    pools = {k: BlockPool(NB, block_size, seed=4) for k in runners}
    seqs = {k: [SeqState() for _ in lens] for k in runners}
    for k in runners:
        for s, n in zip(seqs[k], lens):
            pools[k].allocate(s, n)
    # the prompt's K/V: same random cache content everywhere (CPU mirror for the oracle)
    cpu_cache = torch.randn(2, L, NB, block_size, KVH, D, generator=g).to(torch.bfloat16)
    for r in runners.values():
        r.kv_cache.copy_(cpu_cache)
    for step in range(4):
        q = torch.randn(L, len(lens), H, D, generator=g).to(torch.bfloat16)
        k = torch.randn(L, len(lens), KVH, D, generator=g).to(torch.bfloat16)
        v = torch.randn(L, len(lens), KVH, D, generator=g).to(torch.bfloat16)
        outs = {}
        for name, r in runners.items():
            for s in seqs[name]:
                pools[name].append_token(s)
            bucket = r.prepare_decode(seqs[name])
            assert bucket == (len(lens) if name == "eager" else 8)
            r.q[:, :len(lens)].copy_(q), r.k[:, :len(lens)].copy_(k), r.v[:, :len(lens)].copy_(v)
            outs[name] = r.run()[:, :len(lens)].float().cpu().clone()
        assert torch.equal(outs["eager"], outs["graph"])
        # CPU simulation
        ss = seqs["eager"]
        n = torch.tensor([s.num_tokens for s in ss])
        width = max(len(s.block_table) for s in ss)
        tab = torch.full((len(ss), width), -1, dtype=torch.int32)
        for i, s in enumerate(ss):
            tab[i, :len(s.block_table)] = torch.tensor(s.block_table, dtype=torch.int32)
        slots = torch.tensor([s.block_table[-1] * block_size + (s.num_tokens - 1) % block_size for s in ss])
        for layer in range(L):
            po.store_kvcache_oracle(k[layer], v[layer], cpu_cache[0, layer], cpu_cache[1, layer], slots)
            want = c_oracle.paged_decode(q[layer], cpu_cache[0, layer], cpu_cache[1, layer], tab, n,
                                         1.0 / D ** 0.5, H, KVH, D, block_size)
            got = outs["eager"][layer].view(len(ss), H, D)
            err = (got - want).abs() - (1e-3 + 2.0 ** -8 * want.abs())
            assert err.max() <= 0
    for name, r in runners.items():   # caches ended bit-identical to the CPU mirror
        assert torch.equal(r.kv_cache.cpu().view(torch.uint8), cpu_cache.view(torch.uint8))
