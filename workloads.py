"""Synthetic decode-attention workloads of BASELINE.json's configs.

Harness code shared by bench.py and the GPU tests — not part of the product package.  Everything
is seeded; values are standard normal like the reference's setup_test_data
(reference benchmark_decoding.py:273-295), scale = 1/sqrt(head_dim) (:293).

  c1  16Q/8KV/D128, B=4,   ctx=512,           fp32, BS=16, identity tables   (the reference's CPU case)
  c2  16Q/8KV/D128, B=64,  ctx=2048,          bf16, BS=256, free-list tables (one layer of config 2)
  c3  32Q/8KV/D128, B=256, ctx~U[1024,8192],  bf16, BS=16, permuted pool     (headline, 1 GPU)
  c4  64Q/8KV/D128, B=512, ctx~U[8192,16384], bf16, BS=16, head-sharded at 2/4/8 GPUs
  c5  32Q/8KV/D128, B=32,  shared-prefix tables, ctx<=131072, bf16, BS=16
"""
from __future__ import annotations

import dataclasses
from typing import Optional

import numpy as np
import torch

_DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}


@dataclasses.dataclass
class Spec:
    name: str
    H: int
    KVH: int
    D: int
    BS: int
    dtype: str
    seed: int
    B: int


SPECS = {
    "c1": Spec("c1", 16, 8, 128, 16, "f32", 1, 4),
    "c2": Spec("c2", 16, 8, 128, 256, "bf16", 2, 64),
    "c3": Spec("c3", 32, 8, 128, 16, "bf16", 3, 256),
    "c4": Spec("c4", 64, 8, 128, 16, "bf16", 4, 512),
    "c5": Spec("c5", 32, 8, 128, 16, "bf16", 5, 32),
    # experiment: same bytes as c3 but ONE kv head (pages fully contiguous, 4 KB): isolates the
    # effect of the 2 KB row pitch of the 8-head layout on DRAM efficiency
    "c3x": Spec("c3x", 4, 1, 128, 16, "bf16", 3, 2048),
}


@dataclasses.dataclass
class Tables:
    """Host-side description of a batch: context lengths and block tables (numpy)."""
    spec: Spec
    ctx: np.ndarray          # (B,) int64
    tables: np.ndarray       # (B, width) int32, pad -1
    num_blocks: int          # physical blocks in the pool
    unique_tokens: int       # tokens stored in distinct physical slots (== ctx.sum() unless aliased)


def build_tables(name: str, batch: Optional[int] = None) -> Tables:
    sp = SPECS[name]
    B = batch or sp.B
    rng = np.random.RandomState(sp.seed)
    BS = sp.BS
    if name == "c1":
        ctx = np.full(B, 512, dtype=np.int64)
    elif name == "c2":
        ctx = np.full(B, 2048, dtype=np.int64)
    elif name in ("c3", "c3x"):
        ctx = rng.randint(1024, 8192 + 1, size=B).astype(np.int64)
    elif name == "c4":
        ctx = rng.randint(8192, 16384 + 1, size=B).astype(np.int64)
        ctx[0] = 16384
    elif name == "c5":
        # 4 groups x (B/4) sequences sharing block-aligned prefixes of 16k/32k/64k/96k tokens;
        # suffix lengths log-uniform; one sequence is exactly 131072 long.
        prefixes = [16384, 32768, 65536, 98304]
        per = max(B // 4, 1)
        ctx = np.zeros(B, dtype=np.int64)
        for b in range(B):
            pre = prefixes[min(b // per, 3)]
            room = 131072 - pre
            ctx[b] = pre + int(np.exp(rng.uniform(np.log(16), np.log(room))))
        ctx[B - 1] = 131072
    else:
        raise KeyError(name)
    nblk = (ctx + BS - 1) // BS
    width = int(nblk.max())
    tables = np.full((B, width), -1, dtype=np.int32)
    if name == "c1":
        total = B * width
        tables[:] = np.arange(total, dtype=np.int32).reshape(B, width)
        unique = int(ctx.sum())
    elif name == "c5":
        prefixes = [16384, 32768, 65536, 98304]
        per = max(B // 4, 1)
        pre_blocks = [p // BS for p in prefixes]
        total = int(sum(pre_blocks) + sum(int(nblk[b]) - pre_blocks[min(b // per, 3)] for b in range(B)))
        perm = rng.permutation(total).astype(np.int32)
        at = 0
        shared = []
        for pb in pre_blocks:
            shared.append(perm[at:at + pb])
            at += pb
        unique = int(sum(prefixes))
        for b in range(B):
            g = min(b // per, 3)
            pb = pre_blocks[g]
            own = int(nblk[b]) - pb
            tables[b, :pb] = shared[g]
            tables[b, pb:pb + own] = perm[at:at + own]
            at += own
            unique += int(ctx[b]) - prefixes[g]
    else:
        # engine-shaped: blocks handed out from a shuffled free list (block_manager.py:99-118)
        total = int(nblk.sum())
        perm = rng.permutation(total).astype(np.int32)
        at = 0
        for b in range(B):
            tables[b, :nblk[b]] = perm[at:at + nblk[b]]
            at += int(nblk[b])
        unique = int(ctx.sum())
    return Tables(sp, ctx, tables, total, unique)


def algorithmic_bytes(t: Tables, kvh_local: Optional[int] = None, h_local: Optional[int] = None,
                      idx_bytes: int = 8, unique: bool = False) -> int:
    """K and V rows of each sequence's own tokens counted once per sequence (or once
    per distinct physical slot when ``unique``), + Q read + O written + block-table entries + lengths."""
    sp = t.spec
    es = torch.empty((), dtype=_DT[sp.dtype]).element_size()
    kvh = kvh_local or sp.KVH
    h = h_local or sp.H
    toks = t.unique_tokens if unique else int(t.ctx.sum())
    B = len(t.ctx)
    nblk = int(((t.ctx + sp.BS - 1) // sp.BS).sum())
    return toks * kvh * sp.D * 2 * es + 2 * B * h * sp.D * es + 4 * nblk + idx_bytes * B


def store_bytes(t: Tables, kvh_local: Optional[int] = None, idx_bytes: int = 8) -> int:
    sp = t.spec
    es = torch.empty((), dtype=_DT[sp.dtype]).element_size()
    B = len(t.ctx)
    return 4 * B * (kvh_local or sp.KVH) * sp.D * es + idx_bytes * B


def materialize(t: Tables, device, kv_heads: Optional[slice] = None, seed_offset: int = 0,
                ctx_dtype=torch.int64):
    """Device tensors for one rank.  ``kv_heads`` selects this rank's KV heads (head sharding,
    SURVEY.md §8(e)); query heads follow (G per KV head).  Returns a dict with q, k, v (the new
    token's rows), k_cache, v_cache, block_tables, context_lens, slot_mapping, scale and sizes."""
    sp = t.spec
    dt = _DT[sp.dtype]
    G = sp.H // sp.KVH
    kv_heads = kv_heads or slice(0, sp.KVH)
    kvh = kv_heads.stop - kv_heads.start
    h = kvh * G
    B = len(t.ctx)
    gen = torch.Generator(device=device).manual_seed(sp.seed * 1000 + seed_offset)

    def randn(*shape):
        out = torch.empty(*shape, dtype=dt, device=device)
        # fill in slabs so the fp32 temporaries of normal_() stay small for multi-GB caches
        flat = out.view(-1)
        step = 1 << 28
        for s in range(0, flat.numel(), step):
            flat[s:s + step].normal_(generator=gen)
        return out

    k_cache = randn(t.num_blocks, sp.BS, kvh, sp.D)
    v_cache = randn(t.num_blocks, sp.BS, kvh, sp.D)
    q = randn(B, h, sp.D)
    k = randn(B, kvh, sp.D)
    v = randn(B, kvh, sp.D)
    tables = torch.from_numpy(t.tables).to(device)
    ctx = torch.from_numpy(t.ctx).to(device=device, dtype=ctx_dtype)
    # the step's new token is the last one of each sequence (model_runner.py:238-239)
    last = t.ctx - 1
    slots = t.tables[np.arange(B), last // sp.BS].astype(np.int64) * sp.BS + last % sp.BS
    slot_mapping = torch.from_numpy(slots).to(device=device, dtype=ctx_dtype)
    return dict(q=q, k=k, v=v, k_cache=k_cache, v_cache=v_cache, block_tables=tables,
                context_lens=ctx, slot_mapping=slot_mapping, scale=1.0 / (sp.D ** 0.5),
                H=h, KVH=kvh, D=sp.D, BS=sp.BS, B=B)


def oracle_sample(t: Tables, w: dict, pick):
    """CPU copies of sequences ``pick`` of a materialised batch ``w`` for the oracle: only the pages
    those sequences reference are gathered (on the device) and the tables re-numbered, so a 26 GB
    cache costs a few hundred MB of host memory.  Returns the oracle's argument dict."""
    sp = t.spec
    pick = np.asarray(pick, dtype=np.int64)
    nblk = (t.ctx[pick] + sp.BS - 1) // sp.BS
    used = np.unique(np.concatenate([t.tables[b, :n] for b, n in zip(pick, nblk)]))
    remap = np.full(t.num_blocks, -1, dtype=np.int32)
    remap[used] = np.arange(len(used), dtype=np.int32)
    tables = np.full((len(pick), int(nblk.max())), -1, dtype=np.int32)
    for i, (b, n) in enumerate(zip(pick, nblk)):
        tables[i, :n] = remap[t.tables[b, :n]]
    idx = torch.from_numpy(used.astype(np.int64)).to(w["k_cache"].device)
    sel = torch.from_numpy(pick).to(w["q"].device)
    return dict(q=w["q"][sel].cpu(), k_cache=w["k_cache"][idx].cpu(), v_cache=w["v_cache"][idx].cpu(),
                block_tables=torch.from_numpy(tables), context_lens=torch.from_numpy(t.ctx[pick].copy()),
                scale=w["scale"], H=w["H"], KVH=w["KVH"], D=w["D"], BS=w["BS"])


def highest_page_sequences(t: Tables, k: int = 2):
    """The ``k`` sequences whose block tables reach the highest physical pages of the pool (largest
    byte offsets into the caches)."""
    top = np.where(t.tables >= 0, t.tables, -1).max(axis=1)
    return [int(i) for i in np.argsort(-top)[:k]]
