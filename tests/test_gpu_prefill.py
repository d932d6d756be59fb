"""GPU (-m gpu): prefill attention (varlen causal GQA flash attention) against
the reference's own PyTorch implementation (goldens) and the float64 oracle, and the prefill
branch of Attention.forward (store, then attend; reference attention.py:498-525).

Tolerances: 2e-2 abs against the reference run in the 16-bit dtype (north star); against the
exact oracle |err| <= A + R |want| with (A, R) = (4e-3, 2^-7) for bf16 and (1e-3, 2^-10) for fp16:
probabilities are rounded to the 16-bit type once before the P.V product (unit roundoff 2^-8 /
2^-11 times the p-weighted mean of |v|, as in FlashAttention) plus the output rounding."""
import os

import numpy as np
import pytest
import torch

from cases_prefill import PREFILL_CASES, make_prefill_inputs
from minivllm_b200 import reset_context, set_context
from minivllm_b200.layers.attention import Attention, flash_attention_prefill
from oracle import prefill_oracle

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _check(got, want):
    a, r = (4e-3, 2.0 ** -7) if got.dtype == torch.bfloat16 else (1e-3, 2.0 ** -10)
    got, want = got.double().cpu().numpy(), want.double().cpu().numpy()
    excess = np.abs(got - want) - (a + r * np.abs(want))
    assert np.isfinite(got).all() and excess.max() <= 0, f"max excess {excess.max():.3e}"


KERNELS = {"auto": 0, "mma_sync": 1, "tcgen05_one_tile": 2, "tcgen05_pair": 4}   # mvl_flash_prefill flags


@pytest.mark.parametrize("name", sorted(PREFILL_CASES))
@pytest.mark.parametrize("pass_max", [True, False])
@pytest.mark.parametrize("kernel", sorted(KERNELS))
def test_prefill_matches_reference_and_oracle(cuda_device, name, pass_max, kernel):
    inp = make_prefill_inputs(name)
    z = np.load(os.path.join(GOLD, f"prefill_{name}.npz"))
    out = flash_attention_prefill(inp["q"].to(cuda_device), inp["k"].to(cuda_device), inp["v"].to(cuda_device),
                                  inp["cu_seqlens"].to(cuda_device), inp["scale"], inp["H"], inp["KVH"], inp["D"],
                                  max_seqlen=inp["max_seqlen"] if pass_max else None, flags=KERNELS[kernel])
    assert out.shape == inp["q"].shape and out.dtype == inp["q"].dtype
    native = torch.from_numpy(z["ref_out_bits"]).view(inp["q"].dtype).float()
    assert (out.float().cpu() - native).abs().max() <= 2e-2
    want = prefill_oracle.causal_varlen_attention(inp["q"], inp["k"], inp["v"], inp["cu_seqlens"], inp["scale"],
                                                  inp["H"], inp["KVH"], inp["D"])
    _check(out, want)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H,KVH,D", [(8, 8, 64), (8, 2, 128), (6, 2, 128), (16, 1, 64)])
@pytest.mark.parametrize("kernel", sorted(KERNELS))
def test_prefill_random_ragged_and_strided(cuda_device, dtype, H, KVH, D, kernel):
    g = torch.Generator().manual_seed(H * 100 + KVH * 10 + D)
    lens = [int(x) for x in torch.randint(1, 300, (5,), generator=g)] + [64, 128, 1]
    total = sum(lens)
    qkv = torch.randn(total, (H + 2 * KVH) * D, generator=g).to(dtype)
    q, k, v = qkv.split([H * D, KVH * D, KVH * D], dim=-1)       # strided views of a fused projection
    q, k, v = q.view(total, H, D), k.view(total, KVH, D), v.view(total, KVH, D)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    dq = qkv.to(cuda_device)
    a, b, c = dq.split([H * D, KVH * D, KVH * D], dim=-1)
    out = flash_attention_prefill(a.view(total, H, D), b.view(total, KVH, D), c.view(total, KVH, D),
                                  cu.to(cuda_device), 1.0 / D ** 0.5, H, KVH, D, flags=KERNELS[kernel])
    want = prefill_oracle.causal_varlen_attention(q, k, v, cu, 1.0 / D ** 0.5, H, KVH, D)
    _check(out, want)
    # the first token of every sequence attends to itself only: output == its v row
    for s in cu[:-1].tolist():
        assert torch.equal(out[s].cpu(), v[s].repeat_interleave(H // KVH, dim=0))


def test_attention_forward_prefill_then_decode(cuda_device):
    """Prefill two prompts through Attention.forward (store + causal attention), then one decode step
    over the cache the prefill wrote: the engine's real sequence of calls."""
    from oracle import c_oracle
    H, KVH, D, BS = 8, 2, 128, 16
    g = torch.Generator().manual_seed(9)
    lens = [37, 64]
    total = sum(lens)
    q = torch.randn(total, H, D, generator=g).to(torch.bfloat16)
    k = torch.randn(total, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(total, KVH, D, generator=g).to(torch.bfloat16)
    cu = torch.tensor([0, 37, 101], dtype=torch.int32)
    tables = torch.tensor([[5, 2, 7, -1, -1], [0, 9, 3, 1, 8]], dtype=torch.int32)  # room for the decode token
    slots = torch.cat([tables[b, torch.arange(n) // BS].long() * BS + torch.arange(n) % BS
                       for b, n in enumerate(lens)])
    layer = Attention(H, D, scale=1.0, num_kv_heads=KVH, block_size=BS)
    layer.k_cache = torch.zeros(10, BS, KVH, D, dtype=torch.bfloat16, device=cuda_device)
    layer.v_cache = torch.zeros_like(layer.k_cache)
    set_context(True, cu_seqlens_q=cu.to(cuda_device), cu_seqlens_k=cu.to(cuda_device), max_seqlen_q=64,
                max_seqlen_k=64, slot_mapping=slots.to(cuda_device))
    try:
        out = layer(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device))
    finally:
        reset_context()
    assert out.shape == (total, H * D)
    want = prefill_oracle.causal_varlen_attention(q, k, v, cu, 1.0 / D ** 0.5, H, KVH, D)
    _check(out.view(total, H, D), want)
    flat_k = layer.k_cache.cpu().view(-1, KVH, D)
    assert torch.equal(flat_k[slots], k)
    # decode step on top of the prefilled cache
    qd = torch.randn(2, H, D, generator=g).to(torch.bfloat16)
    kd = torch.randn(2, KVH, D, generator=g).to(torch.bfloat16)
    vd = torch.randn(2, KVH, D, generator=g).to(torch.bfloat16)
    ctx = torch.tensor([38, 65])
    dslots = tables[torch.arange(2), (ctx - 1) // BS].long() * BS + (ctx - 1) % BS
    set_context(False, slot_mapping=dslots.to(cuda_device), context_lens=ctx.to(cuda_device),
                block_tables=tables.to(cuda_device))
    try:
        od = layer(qd.to(cuda_device), kd.to(cuda_device), vd.to(cuda_device))
    finally:
        reset_context()
    wd = c_oracle.paged_decode(qd, layer.k_cache.cpu(), layer.v_cache.cpu(), tables, ctx, 1.0 / D ** 0.5, H, KVH, D, BS)
    err = (od.float().cpu().view(2, H, D) - wd).abs() - (1e-3 + 2.0 ** -8 * wd.abs())
    assert err.max() <= 0


def test_prefill_rejects_unsupported_dtypes(cuda_device):
    q = torch.zeros(4, 2, 64, device=cuda_device, dtype=torch.float64)
    with pytest.raises(NotImplementedError):
        flash_attention_prefill(q, q, q, torch.tensor([0, 4], device=cuda_device), 1.0, 2, 2, 64)
    q16 = q.to(torch.bfloat16)
    with pytest.raises(NotImplementedError):   # mixed dtypes
        flash_attention_prefill(q16, q16, q.float(), torch.tensor([0, 4], device=cuda_device), 1.0, 2, 2, 64)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("D", [64, 128])
def test_prefill_long_sequences_all_kernels(cuda_device, dtype, D):
    """Sequences long enough for many key tiles per query tile (the tcgen05 kernels' steady state:
    double-buffered K/V, lazy rescale, ping-pong), with lengths around the 128 / 256 tile edges and a
    batch whose last tile overhangs the end of the packed tensors.  Every kernel against the oracle;
    large scores (x4) make the running maximum grow late, which exercises the O rescale in TMEM."""
    H, KVH = 4, 2
    g = torch.Generator().manual_seed(D)
    lens = [1500, 257, 256, 255, 129, 1, 2049]
    total = sum(lens)
    q = (torch.randn(total, H, D, generator=g) * 4).to(dtype)
    k = torch.randn(total, KVH, D, generator=g).to(dtype)
    v = torch.randn(total, KVH, D, generator=g).to(dtype)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    want = prefill_oracle.causal_varlen_attention(q, k, v, cu, 1.0 / D ** 0.5, H, KVH, D)
    outs = {}
    for name, flags in KERNELS.items():
        outs[name] = flash_attention_prefill(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device),
                                             cu.to(cuda_device), 1.0 / D ** 0.5, H, KVH, D, max_seqlen=max(lens),
                                             flags=flags)
        _check(outs[name], want)
    # same arithmetic in the two tcgen05 kernels, and auto picked the pair kernel here (max_seqlen >= 2048)
    assert torch.equal(outs["tcgen05_one_tile"], outs["tcgen05_pair"])
    assert torch.equal(outs["auto"], outs["tcgen05_pair"])


def test_prefill_kernel_choice_and_scale_sign(cuda_device):
    """A non-positive scale cannot use the raw-score maximum of the tcgen05 kernels: the call falls
    back to the mma.sync kernel and still matches the oracle."""
    H, KVH, D = 4, 2, 64
    g = torch.Generator().manual_seed(5)
    lens = [300, 200]
    total = sum(lens)
    q = torch.randn(total, H, D, generator=g).to(torch.bfloat16)
    k = torch.randn(total, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(total, KVH, D, generator=g).to(torch.bfloat16)
    cu = torch.tensor([0, 300, 500], dtype=torch.int32)
    for scale in (-0.125, 0.0):
        out = flash_attention_prefill(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device), cu.to(cuda_device),
                                      scale, H, KVH, D, max_seqlen=300)
        _check(out, prefill_oracle.causal_varlen_attention(q, k, v, cu, scale, H, KVH, D))


# ------------------------------------------------------------------ prefill over cached pages (N3 / A-8)
from minivllm_b200.layers.attention import flash_attention_prefill_paged  # noqa: E402
from test_oracle_golden import _paged_case  # noqa: E402

PAGED_KERNELS = {"auto": 0, "tcgen05_one_tile": 2, "tcgen05_pair": 4}


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("BS", [16, 64, 256])
@pytest.mark.parametrize("D", [64, 128])
def test_paged_prefill_with_cached_prefix(cuda_device, dtype, BS, D):
    """New queries over the paged cache (cached prefix + their own K/V already stored): prefixes that
    are page-aligned or not, shorter and longer than a key tile, no prefix at all, one new token
    (= a decode step), chunks around the 128 / 256 tile edges; both tcgen05 kernels vs the oracle."""
    H, KVH = 4, 2
    lens_ctx = [300, 128, 1000, 513, 2049, 77, 640]
    lens_new = [300, 128, 257, 1, 300, 13, 256]
    c = _paged_case(D + BS, lens_ctx, lens_new, H, KVH, D, BS, dtype)
    want = prefill_oracle.causal_paged_attention(c["q"], c["kc"], c["vc"], c["cu_q"], c["bt"], c["ctx"],
                                                 1.0 / D ** 0.5, H, KVH, D, BS)
    outs = {}
    for name, flags in PAGED_KERNELS.items():
        outs[name] = flash_attention_prefill_paged(
            c["q"].to(cuda_device), c["kc"].to(cuda_device), c["vc"].to(cuda_device), c["cu_q"].to(cuda_device),
            c["bt"].to(cuda_device), c["ctx"].to(cuda_device), 1.0 / D ** 0.5, H, KVH, D, BS,
            max_seqlen_q=max(lens_new), flags=flags)
        _check(outs[name], want)
    assert torch.equal(outs["tcgen05_one_tile"], outs["tcgen05_pair"])


def test_paged_prefill_without_prefix_equals_dense_prefill(cuda_device):
    """context_lens == new lengths: the paged entry point must return exactly what the dense one does
    (same kernels, same tiles, K/V merely fetched through the block table)."""
    H, KVH, D, BS = 8, 2, 128, 16
    lens = [700, 129, 256]
    c = _paged_case(3, lens, lens, H, KVH, D, BS, torch.bfloat16)
    k = torch.cat(c["k_full"]).to(cuda_device)
    v = torch.cat(c["v_full"]).to(cuda_device)
    for flags in (2, 4):
        a = flash_attention_prefill_paged(c["q"].to(cuda_device), c["kc"].to(cuda_device), c["vc"].to(cuda_device),
                                          c["cu_q"].to(cuda_device), c["bt"].to(cuda_device), c["ctx"].to(cuda_device),
                                          1.0 / D ** 0.5, H, KVH, D, BS, flags=flags)
        b = flash_attention_prefill(c["q"].to(cuda_device), k, v, c["cu_q"].to(cuda_device), 1.0 / D ** 0.5, H, KVH, D,
                                    flags=flags)
        assert torch.equal(a, b)


def test_attention_forward_prefill_with_cached_prefix(cuda_device):
    """Attention.forward in the prefill branch with block_tables set (what reference
    model_runner.py:209-215 builds when the block manager found a cached prefix): the new tokens' K/V
    are stored, then the queries attend over prefix + new tokens."""
    H, KVH, D, BS = 8, 2, 128, 16
    lens_ctx, lens_new = [200, 96], [72, 96]          # 128 cached + 72 new; nothing cached + 96 new
    c = _paged_case(11, lens_ctx, lens_new, H, KVH, D, BS, torch.bfloat16)
    # remove the new tokens' rows from the cache: forward() has to store them itself
    kc, vc = c["kc"].clone(), c["vc"].clone()
    k_new, v_new, slots = [], [], []
    for s, (ctx, n) in enumerate(zip(lens_ctx, lens_new)):
        pos = torch.arange(ctx - n, ctx)
        sl = c["bt"][s, pos // BS].long() * BS + pos % BS
        kc.view(-1, KVH, D)[sl] = 0
        vc.view(-1, KVH, D)[sl] = 0
        k_new.append(c["k_full"][s][ctx - n:]), v_new.append(c["v_full"][s][ctx - n:]), slots.append(sl)
    cu_k = torch.tensor([0] + list(np.cumsum(lens_ctx)), dtype=torch.int32)
    layer = Attention(H, D, scale=1.0, num_kv_heads=KVH, block_size=BS)
    layer.k_cache, layer.v_cache = kc.to(cuda_device), vc.to(cuda_device)
    set_context(True, cu_seqlens_q=c["cu_q"].to(cuda_device), cu_seqlens_k=cu_k.to(cuda_device),
                max_seqlen_q=max(lens_new), max_seqlen_k=max(lens_ctx), slot_mapping=torch.cat(slots).to(cuda_device),
                block_tables=c["bt"].to(cuda_device))
    try:
        out = layer(c["q"].to(cuda_device), torch.cat(k_new).to(cuda_device), torch.cat(v_new).to(cuda_device))
    finally:
        reset_context()
    assert torch.equal(layer.k_cache.cpu().view(torch.uint8), c["kc"].view(torch.uint8))
    want = prefill_oracle.causal_paged_attention(c["q"], c["kc"], c["vc"], c["cu_q"], c["bt"], c["ctx"],
                                                 1.0 / D ** 0.5, H, KVH, D, BS)
    _check(out.view(-1, H, D), want)


# ------------------------------------------------------------------ fp32 (the reference's default dtype)
def _check_f32(got, want):
    got, want = got.double().cpu().numpy(), want.double().cpu().numpy()
    excess = np.abs(got - want) - (2e-5 + 1e-5 * np.abs(want))
    assert np.isfinite(got).all() and excess.max() <= 0, f"max excess {excess.max():.3e}"


@pytest.mark.parametrize("H,KVH,D", [(8, 2, 128), (6, 6, 64)])
@pytest.mark.parametrize("scale_sign", [1.0, -1.0])
def test_prefill_fp32(cuda_device, H, KVH, D, scale_sign):
    """fp32 q/k/v through the CUDA-core kernel, inside north_star's 1e-4 for fp32 (here 2e-5 + 1e-5 |w|
    against the float64 oracle): ragged lengths around the 64-row tile edges, strided views of a fused
    projection, either scale sign (the maximum is taken on scaled scores)."""
    g = torch.Generator().manual_seed(H + D)
    lens = [1, 63, 64, 65, 130, 700]
    total = sum(lens)
    qkv = torch.randn(total, (H + 2 * KVH) * D, generator=g)
    q, k, v = (x.view(total, -1, D) for x in qkv.split([H * D, KVH * D, KVH * D], dim=-1))
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    scale = scale_sign / D ** 0.5
    dq, dk, dv = (x.view(total, -1, D) for x in qkv.to(cuda_device).split([H * D, KVH * D, KVH * D], dim=-1))
    out = flash_attention_prefill(dq, dk, dv, cu.to(cuda_device), scale, H, KVH, D)
    assert out.dtype == torch.float32
    _check_f32(out, prefill_oracle.causal_varlen_attention(q, k, v, cu, scale, H, KVH, D))


@pytest.mark.parametrize("BS", [16, 256])
def test_paged_prefill_fp32(cuda_device, BS):
    H, KVH, D = 4, 2, 128
    lens_ctx, lens_new = [300, 128, 1000, 65, 77], [300, 64, 257, 1, 13]
    c = _paged_case(BS, lens_ctx, lens_new, H, KVH, D, BS, torch.float32)
    out = flash_attention_prefill_paged(c["q"].to(cuda_device), c["kc"].to(cuda_device), c["vc"].to(cuda_device),
                                        c["cu_q"].to(cuda_device), c["bt"].to(cuda_device), c["ctx"].to(cuda_device),
                                        1.0 / D ** 0.5, H, KVH, D, BS)
    _check_f32(out, prefill_oracle.causal_paged_attention(c["q"], c["kc"], c["vc"], c["cu_q"], c["bt"], c["ctx"],
                                                          1.0 / D ** 0.5, H, KVH, D, BS))


def test_attention_forward_prefill_fp32_then_decode(cuda_device):
    """The reference's default dtype end to end through Attention.forward: fp32 prefill (store + causal
    attention), then an fp32 decode step over what it stored."""
    from oracle import c_oracle
    H, KVH, D, BS = 4, 2, 64, 16
    g = torch.Generator().manual_seed(21)
    lens = [50, 17]
    total = sum(lens)
    q, k, v = (torch.randn(total, n, D, generator=g) for n in (H, KVH, KVH))
    cu = torch.tensor([0, 50, 67], dtype=torch.int32)
    tables = torch.tensor([[3, 0, 5, 6], [2, 4, -1, -1]], dtype=torch.int32)
    slots = torch.cat([tables[b, torch.arange(n) // BS].long() * BS + torch.arange(n) % BS for b, n in enumerate(lens)])
    layer = Attention(H, D, scale=1.0, num_kv_heads=KVH, block_size=BS)
    layer.k_cache = torch.zeros(8, BS, KVH, D, device=cuda_device)
    layer.v_cache = torch.zeros_like(layer.k_cache)
    set_context(True, cu_seqlens_q=cu.to(cuda_device), cu_seqlens_k=cu.to(cuda_device), max_seqlen_q=50,
                max_seqlen_k=50, slot_mapping=slots.to(cuda_device))
    try:
        out = layer(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device))
    finally:
        reset_context()
    _check_f32(out.view(total, H, D), prefill_oracle.causal_varlen_attention(q, k, v, cu, 1.0 / D ** 0.5, H, KVH, D))
    qd, kd, vd = (torch.randn(2, n, D, generator=g) for n in (H, KVH, KVH))
    ctx = torch.tensor([51, 18])
    dslots = tables[torch.arange(2), (ctx - 1) // BS].long() * BS + (ctx - 1) % BS
    set_context(False, slot_mapping=dslots.to(cuda_device), context_lens=ctx.to(cuda_device),
                block_tables=tables.to(cuda_device))
    try:
        od = layer(qd.to(cuda_device), kd.to(cuda_device), vd.to(cuda_device))
    finally:
        reset_context()
    wd = c_oracle.paged_decode(qd, layer.k_cache.cpu(), layer.v_cache.cpu(), tables, ctx, 1.0 / D ** 0.5, H, KVH, D, BS)
    assert (od.cpu().view(2, H, D) - wd).abs().max() <= 1e-4


def test_prefill_full_size_properties(cuda_device):
    """At a size the CPU oracle cannot finish (2 x 8192 tokens, 32Q/8KV/D128 bf16: the bench shape),
    properties that hold at any size: (1) V = 1 gives 1 (the weights of a row sum to one, up to the
    16-bit rounding of P); (2) causality — changing the second half of a sequence leaves the outputs of
    its first half bit-identical; (3) sequences are independent — the batch equals the sequences run
    alone; (4) the paged entry point over a randomly permuted cache equals the dense one bit for bit;
    (5) a strided sample of rows against the float64 oracle evaluated on those rows only."""
    H, KVH, D, n, BS = 32, 8, 128, 8192, 16
    dev = cuda_device
    g = torch.Generator(device=dev).manual_seed(8)
    q = torch.randn(2 * n, H, D, device=dev, generator=g).to(torch.bfloat16)
    k = torch.randn(2 * n, KVH, D, device=dev, generator=g).to(torch.bfloat16)
    v = torch.randn(2 * n, KVH, D, device=dev, generator=g).to(torch.bfloat16)
    cu = torch.tensor([0, n, 2 * n], dtype=torch.int32, device=dev)
    scale = 1.0 / D ** 0.5
    base = flash_attention_prefill(q, k, v, cu, scale, H, KVH, D, max_seqlen=n)
    assert torch.isfinite(base.float()).all()
    # (1)
    ones = flash_attention_prefill(q, k, torch.ones_like(v), cu, scale, H, KVH, D, max_seqlen=n)
    assert (ones.float() - 1).abs().max().item() <= 4e-3
    # (2)
    k2, v2 = k.clone(), v.clone()
    k2[n // 2:n].normal_(generator=g)
    v2[n // 2:n].normal_(generator=g)
    pert = flash_attention_prefill(q, k2, v2, cu, scale, H, KVH, D, max_seqlen=n)
    assert torch.equal(pert[:n // 2], base[:n // 2]) and torch.equal(pert[n:], base[n:])
    assert not torch.equal(pert[n // 2:n], base[n // 2:n])
    # (3)
    alone = flash_attention_prefill(q[n:], k[n:], v[n:], cu[:2], scale, H, KVH, D, max_seqlen=n)
    assert torch.equal(alone, base[n:])
    # (4)
    nb = 2 * n // BS
    perm = torch.randperm(nb, device=dev, generator=g)
    kc = torch.empty(nb, BS, KVH, D, dtype=torch.bfloat16, device=dev)
    vc = torch.empty_like(kc)
    kc[perm] = k.view(nb, BS, KVH, D)
    vc[perm] = v.view(nb, BS, KVH, D)
    bt = perm.to(torch.int32).view(2, n // BS)
    ctx = torch.tensor([n, n], dtype=torch.int32, device=dev)
    paged = flash_attention_prefill_paged(q, kc, vc, cu, bt, ctx, scale, H, KVH, D, BS, max_seqlen_q=n)
    assert torch.equal(paged, base)
    # (5) rows 0, 127, 128, 4095, 8191 of sequence 1, heads 0 and 31
    for row in (0, 127, 128, 4095, n - 1):
        for h in (0, H - 1):
            kk = k[n:n + row + 1, h // (H // KVH)].double().cpu()
            vv = v[n:n + row + 1, h // (H // KVH)].double().cpu()
            s = (kk @ q[n + row, h].double().cpu()) * scale
            want = torch.softmax(s, 0) @ vv
            err = (base[n + row, h].double().cpu() - want).abs() - (4e-3 + 2.0 ** -7 * want.abs())
            assert err.max() <= 0
