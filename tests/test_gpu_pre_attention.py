"""GPU (-m gpu): the fused pre-attention pass (q/k RMSNorm + RoPE + KV store)
against the reference's own modules (goldens) and the float64 oracle, then the whole attention
core (pre-attention + paged decode) against oracle composition.

Tolerances: fp32 2e-5 relative to the tensor's scale (fp32 arithmetic both sides, different
association); 16-bit outputs |err| <= A + R*|want| with R the output dtype's unit roundoff (one
rounding of an fp32 result); v rows and untouched cache bytes bit-exact.
"""
import os

import numpy as np
import pytest
import torch

from cases_pre import PRE_CASES, make_pre_inputs
from minivllm_b200 import reset_context, set_context
from minivllm_b200.layers.pre_attention import Qwen3AttentionCore, fused_qk_norm_rope_store, rope_cache
from oracle import c_oracle
from oracle import pre_attention_oracle as pre

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
ROUND = {torch.float32: (2e-5, 2e-6), torch.bfloat16: (1e-3, 2.0 ** -8), torch.float16: (2e-4, 2.0 ** -11)}


def _rope_table_f64(base, d, max_position):
    """rotary_embedding.py:61-71 evaluated entirely in float64 (no fp32 rounding of the angles)."""
    inv_freq = 1.0 / (float(base) ** (torch.arange(0, d, 2, dtype=torch.float64) / d))
    freqs = torch.arange(max_position, dtype=torch.float64)[:, None] * inv_freq[None, :]
    return torch.cat([freqs.cos(), freqs.sin()], dim=-1)


def _close(got, want, dtype):
    got, want = got.double().cpu().numpy(), want.double().cpu().numpy()
    a, r = ROUND[dtype]
    excess = np.abs(got - want) - (a * max(1.0, np.abs(want).max()) + r * np.abs(want))
    assert excess.max() <= 0, f"max excess {excess.max():.3e}"


@pytest.mark.parametrize("name", sorted(PRE_CASES))
def test_pre_attention_matches_reference_and_oracle(cuda_device, name):
    c = PRE_CASES[name]
    inp = make_pre_inputs(name)
    z = np.load(os.path.join(GOLD, f"pre_{name}.npz"))
    dt = inp["q"].dtype
    N, KVH, D, BS, NB = c["N"], c["KVH"], c["D"], 16, 6
    g = torch.Generator().manual_seed(c["seed"] + 1)
    slots = torch.randperm(NB * BS, generator=g)[:N]
    slots[1] = -1
    kc0 = torch.randn(NB, BS, KVH, D, generator=g).to(dt)
    vc0 = torch.randn(NB, BS, KVH, D, generator=g).to(dt)
    table = pre.rope_table(c["base"], D, c["max_position"])
    # the product's own table: same expression, cos/sin correctly rounded (host-independent).  The
    # reference's fp32 host cos/sin (and powf) may be off by an fp32 ulp of the angle: <= ulp(max_position)
    own = rope_cache(c["base"], D, c["max_position"])
    assert own.dtype == torch.float32 and own.shape == table.shape
    assert (own - table).abs().max().item() <= 2.0 ** -23 * c["max_position"]
    exact = _rope_table_f64(c["base"], D, c["max_position"])
    assert (own.double() - exact).abs().max().item() <= 2.0 ** -23 * c["max_position"]
    gamma = inp["gamma"] if c["norm"] else None
    want_kc, want_vc = kc0.clone(), vc0.clone()
    q_want, k_want = pre.pre_attention(inp["q"], inp["k"], inp["v"], inp["positions"], gamma, c["eps"],
                                       table, want_kc, want_vc, slots)
    kc, vc = kc0.to(cuda_device), vc0.to(cuda_device)
    q_out = fused_qk_norm_rope_store(inp["q"].to(cuda_device), inp["k"].to(cuda_device),
                                     inp["v"].to(cuda_device), inp["positions"].to(cuda_device),
                                     gamma.to(cuda_device) if gamma is not None else None, c["eps"],
                                     table.to(cuda_device), kc, vc, slots.to(cuda_device))
    assert q_out.dtype == dt and q_out.shape == inp["q"].shape
    _close(q_out, q_want, dt)
    # against the reference's own modules (goldens); the bf16 reference rounds after every op
    ref_tol = 1e-5 if dt == torch.float32 else 2e-2
    assert np.abs(q_out.float().cpu().numpy() - z["q_ref"]).max() <= ref_tol * max(1.0, np.abs(z["q_ref"]).max())
    # caches: v rows bit-exact, k rows = rounded oracle, everything else untouched
    assert torch.equal(vc.cpu().view(torch.uint8), want_vc.view(torch.uint8))
    keep = slots != -1
    flat_k = kc.cpu().view(-1, KVH, D)
    _close(flat_k[slots[keep]], k_want[keep], dt)
    untouched = torch.ones(NB * BS, dtype=torch.bool)
    untouched[slots[keep]] = False
    assert torch.equal(flat_k[untouched].view(torch.uint8), kc0.view(-1, KVH, D)[untouched].view(torch.uint8))


def test_pre_attention_on_fused_qkv_views(cuda_device):
    H, KVH, D, N = 8, 2, 128, 5
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn(N, (H + 2 * KVH) * D, generator=g).to(torch.bfloat16).to(cuda_device)
    q, k, v = qkv.split([H * D, KVH * D, KVH * D], dim=-1)
    q, k, v = q.view(N, H, D), k.view(N, KVH, D), v.view(N, KVH, D)
    assert not q.is_contiguous()
    pos = torch.arange(N, device=cuda_device)
    table = rope_cache(10000, D, 64, device=cuda_device)
    gamma = torch.ones(D, device=cuda_device)
    a = fused_qk_norm_rope_store(q, k, v, pos, gamma, 1e-5, table, None, None, None)
    b = fused_qk_norm_rope_store(q.contiguous(), k.contiguous(), v.contiguous(), pos, gamma, 1e-5, table,
                                 None, None, None)
    assert torch.equal(a, b)
    with pytest.raises(ValueError):
        fused_qk_norm_rope_store(q.cpu(), k, v, pos, gamma, 1e-5, table, None, None, None)


def test_attention_core_decode_step(cuda_device):
    """Qwen3AttentionCore (pre-attention pass + paged decode) vs the oracles composed the same way
    (reference models/qwen3.py:86-111, decode; positions = context_lens - 1, qwen3.py:212)."""
    H, KVH, D, BS, B = 16, 8, 128, 16, 5
    g = torch.Generator().manual_seed(77)
    ctx = torch.tensor([1, 16, 17, 130, 300])
    nblk = (ctx + BS - 1) // BS
    total = int(nblk.sum())
    tables = torch.full((B, int(nblk.max())), -1, dtype=torch.int32)
    perm = torch.randperm(total, generator=g).to(torch.int32)
    at = 0
    for b in range(B):
        tables[b, :nblk[b]] = perm[at:at + nblk[b]]
        at += int(nblk[b])
    kc = torch.randn(total, BS, KVH, D, generator=g).to(torch.bfloat16)
    vc = torch.randn(total, BS, KVH, D, generator=g).to(torch.bfloat16)
    qkv = torch.randn(B, (H + 2 * KVH) * D, generator=g).to(torch.bfloat16)
    last = ctx - 1
    slots = tables[torch.arange(B), last // BS].long() * BS + last % BS
    core = Qwen3AttentionCore(H, KVH, D, scale=1.0, max_position=512, block_size=BS)
    core.gamma.copy_(1 + 0.1 * torch.randn(D, generator=g))
    core = core.to(cuda_device)
    core.attention.k_cache, core.attention.v_cache = kc.to(cuda_device), vc.to(cuda_device)
    set_context(False, slot_mapping=slots.to(cuda_device), context_lens=ctx.to(cuda_device),
                block_tables=tables.to(cuda_device))
    try:
        out = core(qkv.to(cuda_device), last.to(cuda_device))
    finally:
        reset_context()
    assert out.shape == (B, H * D)
    # oracle composition on CPU
    q, k, v = qkv.split([H * D, KVH * D, KVH * D], dim=-1)
    table = pre.rope_table(10000, D, 512)
    okc, ovc = kc.clone(), vc.clone()
    qn, _ = pre.pre_attention(q.reshape(B, H, D), k.reshape(B, KVH, D), v.reshape(B, KVH, D), last,
                              core.gamma.cpu(), 1e-5, table, okc, ovc, slots)
    want = c_oracle.paged_decode(qn.to(torch.bfloat16), okc, ovc, tables, ctx, 1.0 / D ** 0.5, H, KVH, D, BS)
    got = out.float().cpu().view(B, H, D)
    err = (got - want).abs() - (4e-3 + 2.0 ** -7 * want.abs())   # two bf16 roundings (q, out) + k ulp flips
    assert err.max() <= 0
    assert torch.equal(core.attention.v_cache.cpu().view(torch.uint8), ovc.view(torch.uint8))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("D,norm", [(128, True), (64, True), (128, False)])
def test_token_per_warp_kernel_equals_row_kernel(cuda_device, dtype, D, norm):
    """From 8192 tokens the pass runs the token-per-warp kernel; in chunks below that size it runs the
    row kernel.  Same fp32 operations in the same order -> bit-identical q and caches.  H + 2 KVH = 10
    rows per token, so the groups of 4 rows a warp handles mix q, k and v rows; some tokens have no
    slot.  The first tokens are also checked against the float64 oracle."""
    H, KVH, N, BS = 6, 2, 8192 + 37, 16
    g = torch.Generator().manual_seed(D + int(norm))
    q = torch.randn(N, H, D, generator=g).to(dtype)
    k = torch.randn(N, KVH, D, generator=g).to(dtype)
    v = torch.randn(N, KVH, D, generator=g).to(dtype)
    pos = torch.randint(0, 512, (N,), generator=g)
    nb = (N + BS - 1) // BS + 2
    slots = torch.randperm(nb * BS, generator=g)[:N]
    slots[::97] = -1
    gamma = (1 + 0.1 * torch.randn(D, generator=g)) if norm else None
    table = pre.rope_table(1e4, D, 512)
    dq, dk, dv, dpos, dslots = (x.to(cuda_device) for x in (q, k, v, pos, slots))
    dgamma = gamma.to(cuda_device) if norm else None
    dtable = table.to(cuda_device)

    def run(chunk):
        kc = torch.zeros(nb, BS, KVH, D, dtype=dtype, device=cuda_device)
        vc = torch.zeros_like(kc)
        outs = [fused_qk_norm_rope_store(dq[s:s + chunk], dk[s:s + chunk], dv[s:s + chunk], dpos[s:s + chunk],
                                         dgamma, 1e-6, dtable, kc, vc, dslots[s:s + chunk])
                for s in range(0, N, chunk)]
        return torch.cat(outs), kc, vc

    q1, kc1, vc1 = run(N)         # token-per-warp kernel
    q2, kc2, vc2 = run(4096)      # row kernel
    assert torch.equal(q1.view(torch.uint8), q2.view(torch.uint8))
    assert torch.equal(kc1.view(torch.uint8), kc2.view(torch.uint8))
    assert torch.equal(vc1.view(torch.uint8), vc2.view(torch.uint8))
    M = 64
    want_kc = torch.zeros(nb, BS, KVH, D, dtype=dtype)
    want_vc = torch.zeros_like(want_kc)
    q_want, _ = pre.pre_attention(q[:M], k[:M], v[:M], pos[:M], gamma, 1e-6, table, want_kc, want_vc, slots[:M])
    _close(q1[:M], q_want, dtype)
    keep = slots[:M][slots[:M] >= 0]
    assert torch.equal(vc1.cpu().view(-1, KVH, D)[keep].view(torch.uint8),
                       want_vc.view(-1, KVH, D)[keep].view(torch.uint8))
