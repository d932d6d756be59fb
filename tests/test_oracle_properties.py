"""The two CPU oracles against each other on random small problems (the golden cases pin them to the reference's
outputs, this pins them to each other everywhere else): oracle/paged_oracle.c — the checker behind bench.py's
`parity` and the full-size GPU tests — against the float64 Python evaluation and the vectorised port of the
reference's decode_torch_optimized (benchmark_decoding.py:136-189), over group sizes, head dims, page sizes,
ragged lengths including 0 and 1, padded tables with stale tails, int32 / int64 lengths and all three dtypes;
and the C store against the Python one.  CPU only."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from oracle import c_oracle
from oracle import paged_attention_oracle as po

DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}


@settings(max_examples=120, deadline=None)
@given(st.integers(1, 4), st.integers(1, 3), st.sampled_from([1, 2, 3, 4, 8]), st.sampled_from([64, 128]),
       st.sampled_from([1, 4, 16, 32]), st.sampled_from(["f32", "bf16", "f16"]), st.booleans(),
       st.integers(0, 2 ** 31 - 1))
def test_c_oracle_equals_python_oracles(B, KVH, G, D, BS, dtype, idx64, seed):
    rng = np.random.RandomState(seed)
    g = torch.Generator().manual_seed(seed)
    H, dt = KVH * G, DT[dtype]
    ctx = rng.randint(0, 71, size=B)
    if seed % 3 == 0:
        ctx[rng.randint(B)] = 1
    nblk = (ctx + BS - 1) // BS
    width = int(nblk.max()) + 2
    pool = int(nblk.sum()) + 3
    perm = rng.permutation(pool)
    tables = np.full((B, width), -1, dtype=np.int32)
    at = 0
    for b in range(B):
        tables[b, :nblk[b]] = perm[at:at + nblk[b]]
        at += nblk[b]
        if seed % 2:                       # stale entries behind a sequence's last block (graph replay buffers)
            tables[b, nblk[b]:] = rng.randint(0, pool, size=width - nblk[b])
    q = torch.randn(B, H, D, generator=g).to(dt)
    kc = torch.randn(pool, BS, KVH, D, generator=g).to(dt)
    vc = torch.randn(pool, BS, KVH, D, generator=g).to(dt)
    bt = torch.from_numpy(tables)
    cl = torch.from_numpy(ctx.astype(np.int64 if idx64 else np.int32))
    scale = 1.0 / D ** 0.5
    args = (q, kc, vc, bt, cl, scale, H, KVH, D, BS)
    c_out = c_oracle.paged_decode(*args).numpy()
    exact = po.decode_attention_exact(*args)
    assert c_out.shape == (B, H, D) and np.isfinite(c_out).all()
    assert np.abs(c_out - exact).max() <= 2e-6
    assert not c_out[ctx == 0].any()                                  # documented deviation: zero rows, not NaN
    live = ctx > 0
    if live.any():
        sel = torch.from_numpy(np.nonzero(live)[0])
        padded = po.decode_attention_padded(q[sel].float(), kc.float(), vc.float(), bt[sel], cl[sel], scale, H, KVH,
                                            D, BS).numpy()
        assert np.abs(padded - c_out[live]).max() <= 2e-5
    one = ctx == 1                                                    # one token: its value row, whatever the scores
    for b in np.nonzero(one)[0]:
        row = vc[tables[b, 0], 0].float().numpy()                     # (KVH, D)
        assert np.abs(c_out[b].reshape(KVH, G, D) - row[:, None, :]).max() <= 1e-6


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 40), st.integers(1, 3), st.sampled_from([64, 128]), st.sampled_from([1, 16, 256]),
       st.sampled_from(["f32", "bf16", "f16"]), st.integers(0, 2 ** 31 - 1))
def test_c_store_equals_python_store(N, KVH, D, BS, dtype, seed):
    rng = np.random.RandomState(seed)
    g = torch.Generator().manual_seed(seed)
    dt = DT[dtype]
    NB = (N + BS - 1) // BS + 2
    slots = rng.permutation(NB * BS)[:N].astype(np.int64)
    slots[rng.rand(N) < 0.2] = -1
    key = torch.randn(N, KVH, D, generator=g).to(dt)
    val = torch.randn(N, KVH, D, generator=g).to(dt)
    kc = torch.randn(NB, BS, KVH, D, generator=g).to(dt)
    vc = torch.randn(NB, BS, KVH, D, generator=g).to(dt)
    k1, v1, k2, v2 = kc.clone(), vc.clone(), kc.clone(), vc.clone()
    sm = torch.from_numpy(slots)
    po.store_kvcache_oracle(key, val, k1, v1, sm)
    c_oracle.store_kvcache(key, val, k2, v2, sm)
    assert torch.equal(k1.view(torch.uint8), k2.view(torch.uint8)) and torch.equal(v1.view(torch.uint8), v2.view(torch.uint8))
    untouched = np.setdiff1d(np.arange(NB * BS), slots[slots >= 0])
    assert torch.equal(k1.view(-1, KVH, D)[untouched], kc.view(-1, KVH, D)[untouched])
