"""Seeded inputs for the prefill goldens (shared by make_golden_prefill.py and the tests)."""
import hashlib

import torch

_DT = {"bf16": torch.bfloat16, "f16": torch.float16}
PREFILL_CASES = {
    "bf16_g4": dict(seed=51, lens=[1, 63, 64, 65, 130], H=4, KVH=1, D=128, dtype="bf16"),
    "bf16_g1_d64": dict(seed=52, lens=[17, 128, 5], H=2, KVH=2, D=64, dtype="bf16"),
    "f16_g8": dict(seed=53, lens=[96, 33, 129], H=8, KVH=1, D=128, dtype="f16"),
    "bf16_long": dict(seed=54, lens=[450, 70], H=2, KVH=1, D=128, dtype="bf16"),
}


def make_prefill_inputs(name):
    c = PREFILL_CASES[name]
    g = torch.Generator().manual_seed(c["seed"])
    total = sum(c["lens"])
    dt = _DT[c["dtype"]]
    q = torch.randn(total, c["H"], c["D"], generator=g).to(dt)
    k = torch.randn(total, c["KVH"], c["D"], generator=g).to(dt)
    v = torch.randn(total, c["KVH"], c["D"], generator=g).to(dt)
    cu = torch.tensor([0] + list(torch.tensor(c["lens"]).cumsum(0)), dtype=torch.int32)
    return dict(q=q, k=k, v=v, cu_seqlens=cu, scale=1.0 / c["D"] ** 0.5, H=c["H"], KVH=c["KVH"], D=c["D"],
                max_seqlen=max(c["lens"]))


def prefill_digest(inp):
    h = hashlib.sha256()
    for key in ("q", "k", "v"):
        h.update(inp[key].contiguous().view(torch.uint8).numpy().tobytes())
    h.update(inp["cu_seqlens"].numpy().tobytes())
    return h.hexdigest()
