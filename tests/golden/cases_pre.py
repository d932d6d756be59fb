"""Seeded inputs for the pre-attention goldens (shared by make_golden_pre.py and the tests)."""
import hashlib

import torch

_DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}

PRE_CASES = {
    "f32_d128": dict(seed=41, N=7, H=8, KVH=2, D=128, dtype="f32", norm=True, eps=1e-5, base=10000,
                     max_position=512),
    "f32_d64": dict(seed=42, N=5, H=4, KVH=4, D=64, dtype="f32", norm=True, eps=1e-5, base=10000,
                    max_position=128),
    "f32_nonorm": dict(seed=43, N=4, H=16, KVH=8, D=128, dtype="f32", norm=False, eps=1e-5,
                       base=1000000, max_position=4096),
    "bf16_d128": dict(seed=44, N=9, H=32, KVH=8, D=128, dtype="bf16", norm=True, eps=1e-5, base=10000,
                      max_position=8192),
}


def make_pre_inputs(name):
    c = PRE_CASES[name]
    g = torch.Generator().manual_seed(c["seed"])
    dt = _DT[c["dtype"]]
    q = (torch.randn(c["N"], c["H"], c["D"], generator=g) * 2).to(dt)
    k = (torch.randn(c["N"], c["KVH"], c["D"], generator=g) * 0.5).to(dt)
    v = torch.randn(c["N"], c["KVH"], c["D"], generator=g).to(dt)
    gamma = (1 + 0.1 * torch.randn(c["D"], generator=g)).to(dt)
    positions = torch.randint(0, c["max_position"], (c["N"],), generator=g)
    positions[0], positions[-1] = 0, c["max_position"] - 1
    return dict(q=q, k=k, v=v, gamma=gamma, positions=positions)


def pre_digest(inp):
    h = hashlib.sha256()
    for key in ("q", "k", "v", "gamma"):
        h.update(inp[key].contiguous().view(torch.uint8).numpy().tobytes())
    h.update(inp["positions"].numpy().tobytes())
    return h.hexdigest()
