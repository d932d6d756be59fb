"""Seeded input recipes for the golden vectors (shared by make_golden.py and the tests).

Inputs are regenerated from the recipe (torch CPU generator, fixed seed) instead of being
committed; each golden file stores a SHA-256 of the regenerated input bytes so that a drift of
the generator is detected instead of silently comparing different problems.
Edge cases follow SURVEY.md §8(c) / Appendix B: lengths 1, BS-1, BS, BS+1, non-multiples of 64,
permuted and aliased (shared-prefix) block tables, G in {1,2,4,5,8}, D in {64,128},
BS in {16,256}, int32 and int64 context_lens, fp32 / bf16 / fp16.
"""
from __future__ import annotations

import hashlib

import numpy as np
import torch

_DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}

# name -> recipe.  tables: "identity" (benchmark_decoding.py:287), "permuted", "aliased".
CASES = {
    "c1_bench_shape": dict(seed=1, H=16, KVH=8, D=128, BS=16, ctx=[512] * 4, dtype="f32",
                           tables="identity", ctx_dtype="i32"),
    "ragged_g4": dict(seed=11, H=8, KVH=2, D=128, BS=16, ctx=[1, 15, 16, 17, 63, 130], dtype="f32",
                      tables="permuted", ctx_dtype="i64"),
    "g1_d64": dict(seed=12, H=4, KVH=4, D=64, BS=16, ctx=[33, 1, 64, 100], dtype="f32",
                   tables="permuted", ctx_dtype="i32"),
    "g2_bs256": dict(seed=13, H=4, KVH=2, D=128, BS=256, ctx=[300, 256, 257, 1], dtype="f32",
                     tables="permuted", ctx_dtype="i64"),
    "g8_ragged": dict(seed=14, H=16, KVH=2, D=128, BS=16, ctx=[200, 7, 129, 64, 65], dtype="f32",
                      tables="permuted", ctx_dtype="i32"),
    "g5_d64": dict(seed=15, H=10, KVH=2, D=64, BS=16, ctx=[90, 31, 32], dtype="f32",
                   tables="permuted", ctx_dtype="i64"),
    "aliased_bf16": dict(seed=16, H=8, KVH=2, D=128, BS=16, ctx=[400, 420, 390, 385], dtype="bf16",
                         tables="aliased", prefix=384, ctx_dtype="i64"),
    "ragged_bf16": dict(seed=17, H=32, KVH=8, D=128, BS=16, ctx=[700, 1, 129, 512, 333, 64, 640, 17],
                        dtype="bf16", tables="permuted", ctx_dtype="i64"),
    "ragged_f16": dict(seed=18, H=8, KVH=4, D=64, BS=16, ctx=[257, 100, 5], dtype="f16",
                       tables="permuted", ctx_dtype="i32"),
    "long_split": dict(seed=19, H=8, KVH=2, D=128, BS=16, ctx=[5000, 3333], dtype="f32",
                       tables="permuted", ctx_dtype="i64"),
    "bs256_bf16": dict(seed=20, H=16, KVH=8, D=128, BS=256, ctx=[2048, 2047, 1500], dtype="bf16",
                       tables="permuted", ctx_dtype="i64"),
}


def make_inputs(name: str):
    """-> dict(q, k_cache, v_cache, block_tables, context_lens, scale, H, KVH, D, BS) on CPU."""
    c = CASES[name]
    H, KVH, D, BS = c["H"], c["KVH"], c["D"], c["BS"]
    ctx = list(c["ctx"])
    B = len(ctx)
    gen = torch.Generator().manual_seed(c["seed"])
    rng = np.random.RandomState(c["seed"])
    nblk = [(n + BS - 1) // BS for n in ctx]
    width = max(nblk)
    tables = np.full((B, width), -1, dtype=np.int32)
    if c["tables"] == "identity":
        total = B * width
        tables[:] = np.arange(total, dtype=np.int32).reshape(B, width)
    elif c["tables"] == "permuted":
        total = sum(nblk) + 3  # a few never-referenced blocks
        perm = rng.permutation(total)
        at = 0
        for b in range(B):
            tables[b, :nblk[b]] = perm[at:at + nblk[b]]
            at += nblk[b]
    elif c["tables"] == "aliased":
        pre = c["prefix"] // BS
        total = pre + sum(n - pre for n in nblk) + 2
        perm = rng.permutation(total)
        shared, at = perm[:pre], pre
        for b in range(B):
            tables[b, :pre] = shared
            tables[b, pre:nblk[b]] = perm[at:at + nblk[b] - pre]
            at += nblk[b] - pre
    else:
        raise KeyError(c["tables"])
    dt = _DT[c["dtype"]]
    q = torch.randn(B, H, D, generator=gen).to(dt)
    k_cache = torch.randn(total, BS, KVH, D, generator=gen).to(dt)
    v_cache = torch.randn(total, BS, KVH, D, generator=gen).to(dt)
    ctx_t = torch.tensor(ctx, dtype=torch.int32 if c["ctx_dtype"] == "i32" else torch.int64)
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, block_tables=torch.from_numpy(tables),
                context_lens=ctx_t, scale=1.0 / (D ** 0.5), H=H, KVH=KVH, D=D, BS=BS)


def input_digest(inp: dict) -> str:
    h = hashlib.sha256()
    for key in ("q", "k_cache", "v_cache", "block_tables", "context_lens"):
        t = inp[key].contiguous()
        h.update(t.view(torch.uint8).numpy().tobytes() if t.dtype.is_floating_point
                 else t.numpy().tobytes())
    return h.hexdigest()
