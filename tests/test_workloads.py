"""The synthetic workloads behind bench.py and the full-size GPU tests (workloads.py): the
recipes are what BASELINE.json's configs say, the byte counts follow §8(d)'s formula, and the page gather
that feeds the oracle at full size returns what the oracle would have seen in the whole batch.  CPU only."""
import numpy as np
import pytest
import torch

import workloads
from oracle import c_oracle


def _valid_entries(t):
    nblk = (t.ctx + t.spec.BS - 1) // t.spec.BS
    for b, n in enumerate(nblk):
        yield b, int(n), t.tables[b]


@pytest.mark.parametrize("name,B,H,KVH,D,BS,dtype,lo,hi", [
    ("c1", 4, 16, 8, 128, 16, "f32", 512, 512),
    ("c2", 64, 16, 8, 128, 256, "bf16", 2048, 2048),
    ("c3", 256, 32, 8, 128, 16, "bf16", 1024, 8192),
    ("c4", 512, 64, 8, 128, 16, "bf16", 8192, 16384),
    ("c5", 32, 32, 8, 128, 16, "bf16", 16384 + 16, 131072),
])
def test_recipes_are_baseline_configs(name, B, H, KVH, D, BS, dtype, lo, hi):
    t = workloads.build_tables(name)
    sp = t.spec
    assert (len(t.ctx), sp.H, sp.KVH, sp.D, sp.BS, sp.dtype) == (B, H, KVH, D, BS, dtype)
    assert lo <= t.ctx.min() and t.ctx.max() <= hi
    assert t.tables.dtype == np.int32 and t.ctx.dtype == np.int64
    assert t.tables.shape[1] == int(((t.ctx + BS - 1) // BS).max())
    for b, n, row in _valid_entries(t):
        assert (row[:n] >= 0).all() and (row[:n] < t.num_blocks).all() and (row[n:] == -1).all()
    again = workloads.build_tables(name)                      # seeded: the same batch every time
    assert np.array_equal(t.ctx, again.ctx) and np.array_equal(t.tables, again.tables)


def test_headline_draws_are_pinned():
    """The draws the committed bench lines and profiles/traffic.json were measured on."""
    c3, c4, c5 = (workloads.build_tables(n) for n in ("c3", "c4", "c5"))
    assert (int(c3.ctx.sum()), int(c3.ctx.min()), int(c3.ctx.max())) == (1188633, 1037, 8185)
    assert (int(c4.ctx.sum()), int(c4.ctx[0])) == (6261466, 16384)
    assert (int(c5.ctx.sum()), int(c5.ctx[-1]), c5.unique_tokens) == (1953606, 131072, 462662)
    assert workloads.algorithmic_bytes(c3) == 4873134740
    assert workloads.algorithmic_bytes(c5, unique=True) == 1896076564


def test_private_pools_do_not_alias_and_c5_groups_do():
    for name in ("c2", "c3"):
        t = workloads.build_tables(name)
        seen = np.concatenate([row[:n] for _, n, row in _valid_entries(t)])
        assert len(np.unique(seen)) == len(seen) == t.num_blocks     # a permutation of the pool
        assert t.unique_tokens == int(t.ctx.sum())
    t = workloads.build_tables("c5")
    BS, per = t.spec.BS, len(t.ctx) // 4
    for g, prefix in enumerate((16384, 32768, 65536, 98304)):
        rows = t.tables[g * per:(g + 1) * per]
        pb = prefix // BS
        assert (rows[:, :pb] == rows[0, :pb]).all()                      # the group's members alias the prefix
        own = np.concatenate([rows[i, pb:(t.ctx[g * per + i] + BS - 1) // BS] for i in range(per)])
        assert len(np.unique(own)) == len(own)                           # ... and nothing after it
        assert not np.intersect1d(own, rows[0, :pb]).size
    used = np.unique(t.tables[t.tables >= 0])
    assert len(used) == t.num_blocks
    assert t.unique_tokens < int(t.ctx.sum()) // 4


@pytest.mark.parametrize("name", ["c1", "c3", "c5"])
def test_algorithmic_bytes_is_section_8d(name):
    """bytes = sum_i ctx_i * KVH * D * 2 * es  +  2 * B * H * D * es  +  4 * sum_i ceil(ctx_i / BS)  +  idx * B."""
    t = workloads.build_tables(name)
    sp = t.spec
    es = 4 if sp.dtype == "f32" else 2
    B = len(t.ctx)
    entries = sum(n for _, n, _ in _valid_entries(t))
    for kvh, h in ((sp.KVH, sp.H), (1, sp.H // sp.KVH)):                  # whole problem, one kv head's shard
        want = int(t.ctx.sum()) * kvh * sp.D * 2 * es + 2 * B * h * sp.D * es + 4 * entries + 8 * B
        assert workloads.algorithmic_bytes(t, kvh_local=kvh, h_local=h) == want
        want_u = want - (int(t.ctx.sum()) - t.unique_tokens) * kvh * sp.D * 2 * es
        assert workloads.algorithmic_bytes(t, kvh_local=kvh, h_local=h, unique=True) == want_u
    assert workloads.store_bytes(t) == 4 * B * sp.KVH * sp.D * es + 8 * B


def test_oracle_sample_sees_what_the_whole_batch_holds():
    """The rows the oracle computes from a gathered, re-numbered sample are bit for bit the rows it computes
    from the whole batch (this is how bench.py's `parity` and the full-size tests feed it)."""
    t = workloads.build_tables("c3", batch=6)
    w = workloads.materialize(t, torch.device("cpu"))
    full = c_oracle.paged_decode(w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"],
                                 w["scale"], w["H"], w["KVH"], w["D"], w["BS"])
    pick = [4, 1, 5]
    s = workloads.oracle_sample(t, w, pick)
    assert s["k_cache"].shape[0] < w["k_cache"].shape[0]
    assert int(s["block_tables"].max()) == s["k_cache"].shape[0] - 1
    part = c_oracle.paged_decode(s["q"], s["k_cache"], s["v_cache"], s["block_tables"], s["context_lens"],
                                 s["scale"], s["H"], s["KVH"], s["D"], s["BS"])
    assert torch.equal(part, full[pick])
    # a head shard materialises that shard's heads of the same draw sizes, tables and lengths replicated
    w1 = workloads.materialize(t, torch.device("cpu"), kv_heads=slice(2, 4))
    assert (w1["KVH"], w1["H"]) == (2, 2 * (t.spec.H // t.spec.KVH))
    assert torch.equal(w1["block_tables"], w["block_tables"]) and torch.equal(w1["context_lens"], w["context_lens"])
    # the step's new token sits in the last slot of each sequence (reference model_runner.py:238-239)
    last = t.ctx - 1
    want = t.tables[np.arange(len(t.ctx)), last // t.spec.BS].astype(np.int64) * t.spec.BS + last % t.spec.BS
    assert np.array_equal(w["slot_mapping"].numpy(), want)


def test_highest_page_sequences():
    t = workloads.build_tables("c3", batch=32)
    top = workloads.highest_page_sequences(t, 3)
    peak = [int(t.tables[b].max()) for b in range(len(t.ctx))]
    assert sorted((peak[b] for b in top), reverse=True) == sorted(peak, reverse=True)[:3]
