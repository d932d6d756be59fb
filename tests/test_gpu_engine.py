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


@pytest.mark.parametrize("block_size", [16, 256])
def test_prefill_then_decode(cuda_device, block_size):
    """Whole-prompt prefill through the runner (store + causal attention, reference
    attention.py:514-531), then decode steps that read what the prefill stored."""
    from oracle import prefill_oracle as pf
    g = torch.Generator().manual_seed(77)
    nb = 24 if block_size == 256 else NB
    lens = [block_size + 5, 3, 2 * block_size, 40]
    maxlen = 3 * block_size + 8
    r = DecodeRunner(L, H, KVH, D, block_size, nb, MAXBS, maxlen, torch.bfloat16, cuda_device, enforce_eager=True)
    pool = BlockPool(nb, block_size, seed=9)
    seqs = [SeqState() for _ in lens]
    for s, n in zip(seqs, lens):
        pool.allocate(s, n)
    T = sum(lens)
    q = torch.randn(L, T, H, D, generator=g).to(torch.bfloat16)
    k = torch.randn(L, T, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(L, T, KVH, D, generator=g).to(torch.bfloat16)
    assert r.prepare_prefill(seqs) == T
    got = r.run_prefill(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device)).float().cpu()
    cu = torch.tensor(np.concatenate([[0], np.cumsum(lens)]), dtype=torch.int32)
    cpu_cache = torch.zeros(2, L, nb, block_size, KVH, D, dtype=torch.bfloat16)
    slots = torch.tensor([s.block_table[j // block_size] * block_size + j % block_size
                          for s in seqs for j in range(len(s))])
    for layer in range(L):
        want = pf.causal_varlen_attention(q[layer], k[layer], v[layer], cu, 1.0 / D ** 0.5, H, KVH, D)
        want = torch.as_tensor(want, dtype=torch.float32).reshape(T, H * D)
        err = (got[layer] - want).abs() - (4e-3 + 2.0 ** -7 * want.abs())
        assert err.max() <= 0
        po.store_kvcache_oracle(k[layer], v[layer], cpu_cache[0, layer], cpu_cache[1, layer], slots)
    assert torch.equal(r.kv_cache.cpu().view(torch.uint8), cpu_cache.view(torch.uint8))
    for step in range(2):
        qd = torch.randn(L, len(lens), H, D, generator=g).to(torch.bfloat16)
        kd = torch.randn(L, len(lens), KVH, D, generator=g).to(torch.bfloat16)
        vd = torch.randn(L, len(lens), KVH, D, generator=g).to(torch.bfloat16)
        for s in seqs:
            pool.append_token(s)
        bucket = r.prepare_decode(seqs)
        r.q[:, :bucket].copy_(qd), r.k[:, :bucket].copy_(kd), r.v[:, :bucket].copy_(vd)
        out = r.run().float().cpu()
        n = torch.tensor([s.num_tokens for s in seqs])
        width = max(len(s.block_table) for s in seqs)
        tab = torch.full((len(seqs), width), -1, dtype=torch.int32)
        for i, s in enumerate(seqs):
            tab[i, :len(s.block_table)] = torch.tensor(s.block_table, dtype=torch.int32)
        dslots = torch.tensor([s.block_table[-1] * block_size + (s.num_tokens - 1) % block_size for s in seqs])
        for layer in range(L):
            po.store_kvcache_oracle(kd[layer], vd[layer], cpu_cache[0, layer], cpu_cache[1, layer], dslots)
            want = c_oracle.paged_decode(qd[layer], cpu_cache[0, layer], cpu_cache[1, layer], tab, n,
                                         1.0 / D ** 0.5, H, KVH, D, block_size)
            err = (out[layer].view(len(seqs), H, D) - want).abs() - (1e-3 + 2.0 ** -8 * want.abs())
            assert err.max() <= 0


def test_prefix_cached_prefill_through_manager_and_runner(cuda_device):
    """N2 + N3 together: the prefix block manager gives three prompts with a common 64-token prefix
    aliased pages; the first prompt is prefilled whole, the others only past their cached blocks
    (attention over the cached pages), and every prompt's outputs equal those of an engine without
    prefix caching (every prompt prefilled whole into private pages).  A decode step follows."""
    from minivllm_b200.engine import PrefixBlockManager, PrefixSeq
    from oracle import prefill_oracle as pf
    BS, nb = 16, 64
    g = torch.Generator().manual_seed(5)
    common = torch.randint(0, 1000, (64,), generator=g).tolist()
    prompts = [common + torch.randint(0, 1000, (n,), generator=g).tolist() for n in (23, 40, 1)]
    # token -> (q, k, v) per layer: a deterministic "model" so equal tokens at equal positions give equal K/V
    emb = torch.randn(1000 + 200, L, H + 2 * KVH, D, generator=g)

    def acts(tokens, start):
        x = torch.stack([emb[t] + 0.01 * emb[1000 + start + i] for i, t in enumerate(tokens)])   # (n, L, H+2KVH, D)
        x = x.to(torch.bfloat16).permute(1, 0, 2, 3)
        return x[:, :, :H], x[:, :, H:H + KVH], x[:, :, H + KVH:]

    def run(prefix_caching):
        r = DecodeRunner(L, H, KVH, D, BS, nb, MAXBS, 256, torch.bfloat16, cuda_device, enforce_eager=True)
        mgr = PrefixBlockManager(nb, BS)
        seqs = [PrefixSeq(p) for p in prompts]
        outs = []
        for s in seqs:                       # one prefill step per prompt, in arrival order
            mgr.allocate(s)
            if not prefix_caching:
                s.num_cached_tokens = 0      # same tables, but recompute (and re-store) everything
            c = s.num_cached_tokens
            q, k, v = acts(s.token_ids[c:], c)
            assert r.prepare_prefill([s]) == len(s) - c
            outs.append(r.run_prefill(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device)).float().cpu())
        return r, mgr, seqs, outs

    r1, mgr1, seqs1, out1 = run(True)
    r0, _, seqs0, out0 = run(False)
    assert [s.num_cached_tokens for s in seqs1] == [0, 64, 64] and mgr1.hits == 8
    assert seqs1[1].block_table[:4] == seqs1[0].block_table[:4] == seqs1[2].block_table[:4]
    assert torch.equal(r1.kv_cache.cpu().view(torch.uint8), r0.kv_cache.cpu().view(torch.uint8))
    for s, a, b in zip(seqs1, out1, out0):
        c = s.num_cached_tokens
        assert (a - b[:, c:]).abs().max() <= 2e-2          # same attention, the cached run read from pages
    # and against the exact oracle for the second prompt
    s = seqs1[1]
    q, _, _ = acts(s.token_ids[64:], 64)
    tab = torch.tensor([s.block_table], dtype=torch.int32)
    for layer in range(L):
        want = pf.causal_paged_attention(q[layer], r1.kv_cache[0, layer].cpu(), r1.kv_cache[1, layer].cpu(),
                                         torch.tensor([0, len(s) - 64]), tab, torch.tensor([len(s)]),
                                         1.0 / D ** 0.5, H, KVH, D, BS)
        err = (out1[1][layer].view(-1, H, D) - want).abs() - (4e-3 + 2.0 ** -7 * want.abs())
        assert err.max() <= 0
    # one decode step over the aliased tables
    for s in seqs1:
        s.append_token(7)
        mgr1.append(s)
    bucket = r1.prepare_decode(seqs1)
    qd = torch.randn(L, bucket, H, D, generator=g).to(torch.bfloat16)
    kd = torch.randn(L, bucket, KVH, D, generator=g).to(torch.bfloat16)
    vd = torch.randn(L, bucket, KVH, D, generator=g).to(torch.bfloat16)
    r1.q[:, :bucket].copy_(qd), r1.k[:, :bucket].copy_(kd), r1.v[:, :bucket].copy_(vd)
    out = r1.run().float().cpu()
    cache = r1.kv_cache.cpu()
    n = torch.tensor([len(s) for s in seqs1])
    width = max(len(s.block_table) for s in seqs1)
    tab = torch.full((len(seqs1), width), -1, dtype=torch.int32)
    for i, s in enumerate(seqs1):
        tab[i, :len(s.block_table)] = torch.tensor(s.block_table, dtype=torch.int32)
    for layer in range(L):
        want = c_oracle.paged_decode(qd[layer], cache[0, layer], cache[1, layer], tab, n, 1.0 / D ** 0.5, H, KVH, D, BS)
        err = (out[layer].view(len(seqs1), H, D) - want).abs() - (1e-3 + 2.0 ** -8 * want.abs())
        assert err.max() <= 0


def test_attention_engine_continuous_batching(cuda_device):
    """The whole loop — scheduler (prefill first, then decode, preemption), prefix-caching block
    manager, prepare_prefill / prepare_decode, dense and paged prefill kernels, fused decode — on a
    synthetic model (tokens -> q/k/v by table look-up, next token a function of (prompt, length)).
    A roomy pool and one so small that sequences are preempted and prefilled again must give every
    (prompt, position) the same attention output: scheduling, page placement, prefix reuse and
    recomputation after preemption do not change results.  A sample of rows is checked against the
    float64 oracle evaluated directly from the token streams."""
    from minivllm_b200.engine import AttentionEngine, Scheduler
    from oracle import prefill_oracle as pf
    BS = 16
    g = torch.Generator().manual_seed(17)
    emb = torch.randn(1500, L, H + 2 * KVH, D, generator=g)
    common = torch.randint(0, 1000, (40,), generator=g).tolist()
    prompts = [common + torch.randint(0, 1000, (n,), generator=g).tolist() for n in (25, 3, 60)] + \
              [torch.randint(0, 1000, (n,), generator=g).tolist() for n in (70, 9)]
    max_tokens = [12, 30, 8, 20, 25]

    def activations(toks, pos):
        x = torch.stack([emb[t] + 0.05 * emb[1000 + p] for t, p in zip(toks, pos)])     # (T, L, heads, D)
        x = x.to(torch.bfloat16).permute(1, 0, 2, 3).to(cuda_device)
        return x[:, :, :H].contiguous(), x[:, :, H:H + KVH].contiguous(), x[:, :, H + KVH:].contiguous()

    def run(num_blocks):
        runner = DecodeRunner(L, H, KVH, D, BS, num_blocks, 8, 160, torch.bfloat16, cuda_device, enforce_eager=True)
        sch = Scheduler(max_num_sequences=8, max_num_batched_tokens=256, max_cached_blocks=num_blocks,
                        block_size=BS, eos=-1)
        tag = {}
        eng = AttentionEngine(runner, sch, activations,
                              lambda seqs, out: [(tag[id(s)] * 7919 + len(s) * 104729) % 1000 for s in seqs])
        seqs = []
        for i, (p, m) in enumerate(zip(prompts, max_tokens)):
            s = eng.add_prompt(p, max_tokens=m, ignore_eos=True)
            tag[id(s)] = i
            seqs.append(s)
        rows = {}
        steps = 0
        while not sch.is_finished():
            batch, is_prefill, out = eng.step()
            for i, s in enumerate(batch):
                # the row of the token at position len(s) - 2 (postprocess has appended its successor)
                rows.setdefault((tag[id(s)], len(s) - 2), []).append(out[:, i].float().cpu())
            steps += 1
            assert steps < 500
        return seqs, rows, sch, eng

    seqs_a, rows_a, sch_a, eng_a = run(64)
    seqs_b, rows_b, sch_b, eng_b = run(12)          # 12 pages of 16 slots: the running set outgrows the pool
    assert sch_a.num_preemptions == 0 and sch_b.num_preemptions > 0
    assert sch_a.block_manager.hits > 0            # the common 40-token prefix: two blocks shared
    for a, b, p, m in zip(seqs_a, seqs_b, prompts, max_tokens):
        assert a.token_ids == b.token_ids and len(a) == len(p) + m and a.status == "finished"
    assert set(rows_a) == set(rows_b)
    for key in rows_a:
        for r in rows_a[key] + rows_b[key]:        # a preempted sequence computes some positions twice
            assert (r - rows_a[key][0]).abs().max() <= 2e-2
    assert len(sch_b.block_manager.free_block_ids) == 12
    # oracle on the final token streams: position len-2 of every sequence, layer 0
    for i, s in enumerate(seqs_a):
        n = len(s) - 1                              # tokens whose K/V exist
        q, k, v = activations(s.token_ids[:n], list(range(n)))
        want = pf.causal_varlen_attention(q[0].cpu(), k[0].cpu(), v[0].cpu(), torch.tensor([0, n]),
                                          1.0 / D ** 0.5, H, KVH, D)[-1].reshape(-1)
        got = rows_a[(i, n - 1)][0][0]
        assert ((got - want).abs() - (4e-3 + 2.0 ** -7 * want.abs())).max() <= 0
