"""GPU (-m gpu): the CUDA path, called through the C ABI, against the oracle and the goldens.

Bars (BASELINE.json north_star): KV-cache write bit-exact; attention within 1e-4 abs when both
sides are fp32, within 2e-2 abs for bf16/fp16 against the reference run in that dtype — plus the
stricter bound against the exact oracle on the same rounded inputs: fp32 accumulation error plus
ONE rounding of the result to the output dtype, |got - want| <= A + R*|want| with R the dtype's unit
roundoff (bf16 2^-8, fp16 2^-11) and A = 1e-3 / 2e-4 / 1e-4 (bf16 / fp16 / fp32).
"""
import os

import numpy as np
import pytest
import torch

import workloads
from cases import CASES, make_inputs
from minivllm_b200 import reset_context, set_context
from minivllm_b200.layers.attention import Attention, paged_attention_decode, store_kvcache
from oracle import c_oracle
from oracle import paged_attention_oracle as po

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL_EXACT = {torch.float32: (1e-4, 0.0), torch.bfloat16: (1e-3, 2.0 ** -8), torch.float16: (2e-4, 2.0 ** -11)}


def _assert_close(got, want, dtype):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    a, r = TOL_EXACT[dtype]
    excess = np.abs(got - want) - (a + r * np.abs(want))
    assert excess.max() <= 0, f"max excess {excess.max():.3e}, max abs err {np.abs(got - want).max():.3e}"


def _dev(inp, device):
    return {k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in inp.items()}


def _run(inp, **kw):
    return paged_attention_decode(inp["q"], inp["k_cache"], inp["v_cache"], inp["block_tables"],
                                  inp["context_lens"], inp["scale"], inp["H"], inp["KVH"], inp["D"],
                                  inp["BS"], **kw)


def _oracle(inp):
    return c_oracle.paged_decode(inp["q"], inp["k_cache"], inp["v_cache"], inp["block_tables"],
                                 inp["context_lens"], inp["scale"], inp["H"], inp["KVH"], inp["D"],
                                 inp["BS"])


# ------------------------------------------------------------------ store_kvcache: bit-exact
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("idx", [torch.int64, torch.int32])
def test_store_bit_exact(cuda_device, dtype, idx):
    g = torch.Generator().manual_seed(7)
    NB, BS, KVH, D, N = 40, 16, 8, 128, 300
    key = torch.randn(N, KVH, D, generator=g).to(dtype)
    val = torch.randn(N, KVH, D, generator=g).to(dtype)
    slots = torch.randperm(NB * BS, generator=g)[:N].to(torch.int64)
    slots[::7] = -1
    kc = torch.randn(NB, BS, KVH, D, generator=g).to(dtype)
    vc = torch.randn(NB, BS, KVH, D, generator=g).to(dtype)
    want_k, want_v = kc.clone(), vc.clone()
    po.store_kvcache_oracle(key, val, want_k, want_v, slots)
    dk, dv = kc.to(cuda_device), vc.to(cuda_device)
    assert store_kvcache(key.to(cuda_device), val.to(cuda_device), dk, dv,
                         slots.to(cuda_device, idx), BS) is None
    # whole-cache equality: untouched slots must stay untouched
    assert torch.equal(dk.cpu().view(torch.uint8), want_k.view(torch.uint8))
    assert torch.equal(dv.cpu().view(torch.uint8), want_v.view(torch.uint8))


def test_store_edge_cases(cuda_device):
    BS, KVH, D = 256, 2, 64
    kc = torch.zeros(3, BS, KVH, D, device=cuda_device, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    # empty input, all -1, first and last slot of the pool
    store_kvcache(kc.new_zeros(0, KVH, D), kc.new_zeros(0, KVH, D), kc, vc,
                  torch.zeros(0, dtype=torch.int64, device=cuda_device), BS)
    k = torch.randn(3, KVH, D, device=cuda_device).to(torch.bfloat16)
    v = torch.randn(3, KVH, D, device=cuda_device).to(torch.bfloat16)
    store_kvcache(k, v, kc, vc, torch.full((3,), -1, dtype=torch.int64, device=cuda_device), BS)
    assert torch.count_nonzero(kc) == 0 and torch.count_nonzero(vc) == 0
    store_kvcache(k, v, kc, vc, torch.tensor([0, -1, 3 * BS - 1], device=cuda_device), BS)
    assert torch.equal(kc[0, 0], k[0]) and torch.equal(vc[2, BS - 1], v[2])
    assert torch.count_nonzero(kc) == torch.count_nonzero(k[0]) + torch.count_nonzero(k[2])
    # strided views (k/v slices of a fused qkv projection) are consumed without a copy
    H = 4
    qkv = torch.randn(5, (H + 2 * KVH) * D, device=cuda_device).to(torch.bfloat16)
    k_view = qkv[:, H * D:(H + KVH) * D].view(5, KVH, D)
    v_view = qkv[:, (H + KVH) * D:].view(5, KVH, D)
    assert not k_view.is_contiguous()
    slots = torch.tensor([10, 11, -1, 300, 12], device=cuda_device)
    kc.zero_(), vc.zero_()
    store_kvcache(k_view, v_view, kc, vc, slots, BS)
    flat_k, flat_v = kc.view(-1, KVH, D), vc.view(-1, KVH, D)
    for t, s in enumerate(slots.tolist()):
        if s >= 0:
            assert torch.equal(flat_k[s], k_view[t]) and torch.equal(flat_v[s], v_view[t])


# ------------------------------------------------------------------ decode vs reference goldens
@pytest.mark.parametrize("name", sorted(CASES))
def test_decode_matches_reference_golden(cuda_device, name):
    inp = make_inputs(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    out = _run(_dev(inp, cuda_device))
    assert out.dtype == inp["q"].dtype and out.shape == inp["q"].shape
    got = out.float().cpu().numpy()
    if inp["q"].dtype == torch.float32:
        assert np.abs(got - z["ref_out"]).max() <= 1e-4          # both sides fp32
    else:
        assert np.abs(got - z["ref_out"]).max() <= 2e-2          # reference run in bf16/fp16
        _assert_close(got, z["ref_out_f32"], inp["q"].dtype)


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("chunk", [64, 128, 100000])
def test_decode_split_sizes_agree_with_oracle(cuda_device, name, chunk):
    inp = make_inputs(name)
    want = _oracle(inp).numpy()
    got = _run(_dev(inp, cuda_device), chunk_tokens=chunk).float().cpu().numpy()
    _assert_close(got, want, inp["q"].dtype)


def _random_case(seed, dtype, H, KVH, D, BS, B, max_ctx, ctx_dtype=torch.int64):
    g = torch.Generator().manual_seed(seed)
    rng = np.random.RandomState(seed)
    ctx = rng.randint(1, max_ctx + 1, size=B)
    nblk = (ctx + BS - 1) // BS
    total = int(nblk.sum()) + 5
    perm = rng.permutation(total)
    tables = np.full((B, int(nblk.max()) + 2), -1, dtype=np.int32)
    at = 0
    for b in range(B):
        tables[b, :nblk[b]] = perm[at:at + nblk[b]]
        at += nblk[b]
        # stale garbage beyond the valid range must never be dereferenced (Appendix B-6)
        tables[b, nblk[b]:] = 2 ** 30
    return dict(q=torch.randn(B, H, D, generator=g).to(dtype),
                k_cache=torch.randn(total, BS, KVH, D, generator=g).to(dtype),
                v_cache=torch.randn(total, BS, KVH, D, generator=g).to(dtype),
                block_tables=torch.from_numpy(tables), context_lens=torch.from_numpy(ctx).to(ctx_dtype),
                scale=1.0 / D ** 0.5, H=H, KVH=KVH, D=D, BS=BS)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H,KVH", [(4, 4), (8, 4), (32, 8), (16, 2), (10, 2), (32, 2), (6, 2)])
@pytest.mark.parametrize("D,BS", [(128, 16), (64, 16), (128, 256), (64, 8), (128, 4)])
def test_decode_random_ragged(cuda_device, dtype, H, KVH, D, BS):
    inp = _random_case(1000 + H * 7 + KVH + D + BS, dtype, H, KVH, D, BS, B=5, max_ctx=600,
                       ctx_dtype=torch.int32 if BS == 16 else torch.int64)
    want = _oracle(inp).numpy()
    got = _run(_dev(inp, cuda_device)).float().cpu().numpy()
    assert np.isfinite(got).all()
    _assert_close(got, want, dtype)


def test_decode_edge_lengths(cuda_device):
    # 1, BS-1, BS, BS+1, tile and chunk boundaries, and a zero-length (graph padding) row
    inp = _random_case(77, torch.bfloat16, 32, 8, 128, 16, B=12, max_ctx=10)
    ctx = torch.tensor([1, 15, 16, 17, 63, 64, 65, 127, 128, 129, 1024, 1025])
    BS = 16
    nblk = (ctx + BS - 1) // BS
    total = int(nblk.sum())
    g = torch.Generator().manual_seed(3)
    tables = torch.full((12, int(nblk.max())), -1, dtype=torch.int32)
    perm = torch.randperm(total, generator=g).to(torch.int32)
    at = 0
    for b in range(12):
        tables[b, :nblk[b]] = perm[at:at + nblk[b]]
        at += int(nblk[b])
    inp.update(k_cache=torch.randn(total, BS, 8, 128, generator=g).to(torch.bfloat16),
               v_cache=torch.randn(total, BS, 8, 128, generator=g).to(torch.bfloat16),
               block_tables=tables, context_lens=ctx)
    want = _oracle(inp).numpy()
    got = _run(_dev(inp, cuda_device)).float().cpu().numpy()
    _assert_close(got, want, torch.bfloat16)
    # ctx == 1 -> output is exactly that token's V row
    blk = int(tables[0, 0])
    assert torch.equal(torch.from_numpy(got[0, 5]), inp["v_cache"][blk, 0, 1].float())
    # zero-length rows give zeros and do not disturb their neighbours
    inp["context_lens"] = ctx.clone()
    inp["context_lens"][3] = 0
    got0 = _run(_dev(inp, cuda_device)).float().cpu().numpy()
    assert np.abs(got0[3]).max() == 0
    keep = [i for i in range(12) if i != 3]
    assert np.array_equal(got0[keep], got[keep])


def test_decode_large_scores_are_stable(cuda_device):
    # a large scale drives scores far apart: exercises the lazy-max refresh and masking
    inp = _random_case(5, torch.float32, 8, 2, 128, 16, B=3, max_ctx=900)
    inp["scale"] = 3.0
    want = _oracle(inp).numpy()
    got = _run(_dev(inp, cuda_device)).float().cpu().numpy()
    assert np.isfinite(got).all() and np.abs(got - want).max() <= 1e-4


def test_decode_is_deterministic_and_respects_out(cuda_device):
    inp = _dev(make_inputs("ragged_bf16"), cuda_device)
    a = _run(inp)
    buf = torch.full_like(a, 7)
    b = _run(inp, out=buf)
    assert b is buf and torch.equal(a, b)


# ------------------------------------------------------------------ Attention module + Context
@pytest.mark.parametrize("block_size", [16, 256])
def test_attention_forward_decode_step(cuda_device, block_size):
    """store-then-attend in one forward, as reference attention.py:493-557 (decode branch)."""
    H, KVH, D, B = 16, 8, 128, 6
    g = torch.Generator().manual_seed(21)
    ctx = torch.tensor([5, 16, 17, 300, 256, 257])
    nblk = (ctx + block_size - 1) // block_size
    total = int(nblk.sum())
    tables = torch.full((B, int(nblk.max())), -1, dtype=torch.int32)
    perm = torch.randperm(total, generator=g).to(torch.int32)
    at = 0
    for b in range(B):
        tables[b, :nblk[b]] = perm[at:at + nblk[b]]
        at += int(nblk[b])
    kc = torch.randn(total, block_size, KVH, D, generator=g).to(torch.bfloat16)
    vc = torch.randn(total, block_size, KVH, D, generator=g).to(torch.bfloat16)
    q = torch.randn(B, H, D, generator=g).to(torch.bfloat16)
    k = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
    last = ctx - 1
    slots = tables[torch.arange(B), last // block_size].long() * block_size + last % block_size
    # oracle: write the new token, then attend over ctx tokens (Appendix B-9)
    ok, ov = kc.clone(), vc.clone()
    po.store_kvcache_oracle(k, v, ok, ov, slots)
    want = c_oracle.paged_decode(q, ok, ov, tables, ctx, 1.0 / D ** 0.5, H, KVH, D, block_size)

    layer = Attention(H, D, scale=1.0, num_kv_heads=KVH)  # block_size left at the default 16
    layer.k_cache, layer.v_cache = kc.to(cuda_device), vc.to(cuda_device)
    set_context(False, slot_mapping=slots.to(cuda_device), context_lens=ctx.to(cuda_device),
                block_tables=tables.to(cuda_device))
    try:
        out = layer(q.to(cuda_device).unsqueeze(1), k.to(cuda_device), v.to(cuda_device))
    finally:
        reset_context()
    assert out.shape == (B, H * D) and out.dtype == torch.bfloat16
    assert torch.equal(layer.k_cache.cpu().view(torch.uint8), ok.view(torch.uint8))
    assert torch.equal(layer.v_cache.cpu().view(torch.uint8), ov.view(torch.uint8))
    _assert_close(out.float().cpu().view(B, H, D).numpy(), want.numpy(), torch.bfloat16)


def test_bad_arguments_raise(cuda_device):
    inp = _dev(make_inputs("g1_d64"), cuda_device)
    with pytest.raises(ValueError):
        _run(dict(inp, block_tables=inp["block_tables"].long()))
    with pytest.raises(ValueError):
        _run(dict(inp, q=inp["q"].double()))
    with pytest.raises(ValueError):  # head_dim 32 is not built
        paged_attention_decode(inp["q"][..., :32].contiguous(), inp["k_cache"][..., :32].contiguous(),
                               inp["v_cache"][..., :32].contiguous(), inp["block_tables"],
                               inp["context_lens"], 1.0, 4, 4, 32, 16)
    with pytest.raises(ValueError):  # block size 12 is not a power of two
        kc = torch.zeros(4, 12, 4, 64, device=cuda_device)
        paged_attention_decode(inp["q"], kc, kc, inp["block_tables"], inp["context_lens"], 1.0, 4, 4, 64, 12)


# ------------------------------------------------------------------ full-size properties (C3 shape)
def test_c3_full_size_properties(cuda_device):
    """BASELINE config 3 at full size: sub-batch parity against the oracle, then size-independent
    properties: split-size invariance, block-permutation invariance, V-linearity, ones-V row sums."""
    t = workloads.build_tables("c3")
    w = workloads.materialize(t, cuda_device)
    args = (w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"],
            w["KVH"], w["D"], w["BS"])
    out = paged_attention_decode(w["q"], *args)
    assert torch.isfinite(out.float()).all()
    # (1) oracle parity on a strided sample of sequences (the oracle needs seconds per sequence)
    pick = [0, 85, 170, 255]
    sub_want = c_oracle.paged_decode(w["q"][pick].cpu(), w["k_cache"].cpu(), w["v_cache"].cpu(),
                                     w["block_tables"][pick].cpu(), w["context_lens"][pick].cpu(),
                                     w["scale"], w["H"], w["KVH"], w["D"], w["BS"])
    _assert_close(out[pick].float().cpu().numpy(), sub_want.numpy(), torch.bfloat16)
    # (2) the split size changes the reduction tree, not the result (beyond fp32 rounding)
    out_512 = paged_attention_decode(w["q"], *args, chunk_tokens=512)
    assert (out.float() - out_512.float()).abs().max() <= 2e-2
    # (3) V == 1 everywhere -> every output element is exactly 1 (softmax rows sum to 1)
    ones = torch.ones_like(w["v_cache"])
    o1 = paged_attention_decode(w["q"], w["k_cache"], ones, *args[2:])
    assert (o1.float() - 1).abs().max() <= 1e-2
    # (4) permuting physical blocks (cache rows and table entries together) changes nothing
    perm = torch.randperm(t.num_blocks, device=cuda_device)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(t.num_blocks, device=cuda_device)
    bt = w["block_tables"].long()
    bt2 = torch.where(bt >= 0, inv[bt.clamp(min=0)], bt).to(torch.int32)
    kc2, vc2 = w["k_cache"][perm], w["v_cache"][perm]
    out_p = paged_attention_decode(w["q"], kc2, vc2, bt2, *args[3:])
    assert torch.equal(out_p, out)


def _sample_parity(t, w, out, pick, dtype=torch.bfloat16):
    s = workloads.oracle_sample(t, w, pick)
    want = c_oracle.paged_decode(s["q"], s["k_cache"], s["v_cache"], s["block_tables"], s["context_lens"],
                                 s["scale"], s["H"], s["KVH"], s["D"], s["BS"])
    _assert_close(out[pick].float().cpu().numpy(), want.numpy(), dtype)


def test_c2_full_size_properties(cuda_device):
    """BASELINE config 2, one layer at full size (16Q/8KV/D128, batch 64, 2048-token contexts, 256-slot
    pages): sample parity against the oracle, split-size invariance, ones-V, and the fused step."""
    from minivllm_b200.layers.attention import decode_step
    t = workloads.build_tables("c2")
    w = workloads.materialize(t, cuda_device)
    args = (w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"],
            w["KVH"], w["D"], w["BS"])
    out = paged_attention_decode(w["q"], *args)
    assert torch.isfinite(out.float()).all()
    _sample_parity(t, w, out, [0, 21, 42, 63] + workloads.highest_page_sequences(t, 1))
    for chunk in (256, 100000):
        assert (out.float() - paged_attention_decode(w["q"], *args, chunk_tokens=chunk).float()).abs().max() <= 2e-2
    o1 = paged_attention_decode(w["q"], w["k_cache"], torch.ones_like(w["v_cache"]), *args[2:])
    assert (o1.float() - 1).abs().max() <= 1e-2
    # the one-launch step (store fused in) on the same batch: caches bit-exact, output as above + new row
    kc, vc = w["k_cache"].clone(), w["v_cache"].clone()
    got = decode_step(w["q"], w["k"], w["v"], kc, vc, w["slot_mapping"], w["block_tables"], w["context_lens"],
                      w["scale"], w["H"], w["KVH"], w["D"], w["BS"])
    want_k, want_v = w["k_cache"].clone(), w["v_cache"].clone()
    store_kvcache(w["k"], w["v"], want_k, want_v, w["slot_mapping"], w["BS"])
    assert torch.equal(kc, want_k) and torch.equal(vc, want_v)
    w2 = dict(w, k_cache=kc, v_cache=vc)
    _sample_parity(t, w2, got, [0, 31, 63])


def test_c4_full_size_properties(cuda_device):
    """BASELINE config 4 at full size on one GPU (64Q/8KV/D128, batch 512, contexts 8k-16k): two 13 GB
    caches whose byte offsets pass 2^32 (where the reference's Triton kernel faults).  The sample
    deliberately holds the sequences that reach the highest pages of the pool."""
    free, _ = torch.cuda.mem_get_info(cuda_device)
    if free < 60e9:
        pytest.skip("needs ~45 GB of free HBM")
    t = workloads.build_tables("c4")
    w = workloads.materialize(t, cuda_device)
    assert w["k_cache"].numel() * w["k_cache"].element_size() > 2 ** 32
    args = (w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"],
            w["KVH"], w["D"], w["BS"])
    out = paged_attention_decode(w["q"], *args)
    assert torch.isfinite(out.float()).all()
    high = workloads.highest_page_sequences(t, 2)
    assert int(t.tables[high[0]].max()) * w["BS"] * w["KVH"] * w["D"] * 2 > 2 ** 32
    _sample_parity(t, w, out, [0, 255, 511] + high)
    assert (out.float() - paged_attention_decode(w["q"], *args, chunk_tokens=1024).float()).abs().max() <= 2e-2
    # CUDA-core kernel on the same batch (other arithmetic, same tolerance)
    from minivllm_b200 import _ext
    ffma = paged_attention_decode(w["q"], *args, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_CUDA_CORES)
    _sample_parity(t, w, ffma, high)
    w["v_cache"].fill_(1)
    o1 = paged_attention_decode(w["q"], *args)
    assert (o1.float() - 1).abs().max() <= 1e-2


def test_c5_full_size_properties(cuda_device):
    """BASELINE config 5 at full size (batch 32, contexts up to 131072, four groups aliasing 16k-96k-token
    prefixes): unshared, prefix sharing narrow (bit-identical) and wide (group kernel), each against the
    oracle on a sample holding the 131072-token sequence, one member of every group and the sequences on
    the highest pages."""
    from minivllm_b200 import _ext
    t = workloads.build_tables("c5")
    w = workloads.materialize(t, cuda_device)
    assert int(t.ctx.max()) == 131072 and t.unique_tokens < int(t.ctx.sum())
    args = (w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"],
            w["KVH"], w["D"], w["BS"])
    pick = sorted(set([0, 9, 18, 31] + workloads.highest_page_sequences(t, 1)))
    base = paged_attention_decode(w["q"], *args, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SPLIT_GRID)
    _sample_parity(t, w, base, pick)
    stream = paged_attention_decode(w["q"], *args, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_STREAM)
    assert (stream.float() - base.float()).abs().max() <= 2e-2
    _sample_parity(t, w, stream, pick)
    narrow = paged_attention_decode(w["q"], *args, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX |
                                    _ext.DECODE_SHARE_NARROW)
    assert torch.equal(narrow, base)
    for _ in range(2):   # twice: counters and group lists must be left clean
        wide = paged_attention_decode(w["q"], *args, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX)
        assert torch.isfinite(wide.float()).all()
        assert (wide.float() - base.float()).abs().max() <= 2e-2
        _sample_parity(t, w, wide, pick)
    o1 = paged_attention_decode(w["q"], w["k_cache"], torch.ones_like(w["v_cache"]), *args[2:],
                                flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX)
    assert (o1.float() - 1).abs().max() <= 1e-2


# ------------------------------------------------------------------ persistent stream kernel
def _stream_flags():
    from minivllm_b200 import _ext
    return _ext.DECODE_FUSED_COMBINE | _ext.DECODE_STREAM


@pytest.mark.parametrize("name", sorted(n for n in CASES if CASES[n]["dtype"] != "f32"))
def test_stream_kernel_matches_reference_and_oracle(cuda_device, name):
    """The persistent, evenly cut form of the tensor-core kernel (MVL_DECODE_STREAM) on every 16-bit
    golden case: same bars as the chunk-grid kernel, twice in a row (its arrival counters must come
    back to zero), and deterministic."""
    inp = make_inputs(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    dev = _dev(inp, cuda_device)
    outs = [_run(dev, flags=_stream_flags()) for _ in range(2)]
    assert torch.equal(outs[0], outs[1])
    got = outs[1].float().cpu().numpy()
    assert np.abs(got - z["ref_out"]).max() <= 2e-2
    _assert_close(got, z["ref_out_f32"], inp["q"].dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H,KVH", [(4, 4), (32, 8), (16, 2), (10, 2), (24, 2)])
@pytest.mark.parametrize("D,BS", [(128, 16), (64, 16), (128, 256), (64, 8)])
@pytest.mark.parametrize("B,max_ctx", [(5, 600), (37, 3000), (1, 20000)])
def test_stream_kernel_random_ragged(cuda_device, dtype, H, KVH, D, BS, B, max_ctx):
    """Pieces that start and end inside sequences, units spread over many pieces (one long sequence),
    pieces holding many whole sequences (short ones), G > 8 (two head tiles), pages smaller than a tile."""
    inp = _random_case(5000 + H * 7 + KVH + D + BS + B, dtype, H, KVH, D, BS, B=B, max_ctx=max_ctx,
                       ctx_dtype=torch.int32 if BS == 16 else torch.int64)
    want = _oracle(inp).numpy()
    got = _run(_dev(inp, cuda_device), flags=_stream_flags()).float().cpu().numpy()
    assert np.isfinite(got).all()
    _assert_close(got, want, dtype)


def test_stream_kernel_edges_and_fused_step(cuda_device):
    """Zero-length rows give zeros, lengths around the 64-token row and the page size, the fused
    KV-cache write (caches bit-exact against the store kernel), and a batch past the kernel's
    1024-sequence limit falls back to the chunk grid with the same results."""
    from minivllm_b200.layers.attention import decode_step
    inp = _random_case(78, torch.bfloat16, 32, 8, 128, 16, B=14, max_ctx=10)
    ctx = torch.tensor([1, 15, 16, 17, 63, 64, 65, 127, 128, 129, 1024, 1025, 0, 4000])
    BS, B = 16, 14
    nblk = (ctx + BS - 1) // BS
    total = int(nblk.sum())
    g = torch.Generator().manual_seed(5)
    tables = torch.full((B, int(nblk.max())), -1, dtype=torch.int32)
    perm = torch.randperm(total, generator=g).to(torch.int32)
    at = 0
    for b in range(B):
        tables[b, :nblk[b]] = perm[at:at + nblk[b]]
        at += int(nblk[b])
    inp.update(k_cache=torch.randn(total, BS, 8, 128, generator=g).to(torch.bfloat16),
               v_cache=torch.randn(total, BS, 8, 128, generator=g).to(torch.bfloat16),
               block_tables=tables, context_lens=ctx)
    dev = _dev(inp, cuda_device)
    got = _run(dev, flags=_stream_flags()).float().cpu().numpy()
    keep = ctx.numpy() > 0
    _assert_close(got[keep], _oracle(inp).numpy()[keep], torch.bfloat16)
    assert (got[~keep] == 0).all()
    # fused step
    k = torch.randn(B, 8, 128, generator=g).to(torch.bfloat16)
    v = torch.randn(B, 8, 128, generator=g).to(torch.bfloat16)
    last = (ctx - 1).clamp(min=0)
    slots = tables[torch.arange(B), last // BS].long() * BS + last % BS
    slots[12] = -1
    kc, vc = dev["k_cache"].clone(), dev["v_cache"].clone()
    out = decode_step(dev["q"], k.to(cuda_device), v.to(cuda_device), kc, vc, slots.to(cuda_device),
                      dev["block_tables"], dev["context_lens"], inp["scale"], 32, 8, 128, BS, flags=_stream_flags())
    want_k, want_v = inp["k_cache"].clone(), inp["v_cache"].clone()
    po.store_kvcache_oracle(k, v, want_k, want_v, slots)
    assert torch.equal(kc.cpu().view(torch.uint8), want_k.view(torch.uint8))
    assert torch.equal(vc.cpu().view(torch.uint8), want_v.view(torch.uint8))
    want = c_oracle.paged_decode(inp["q"], want_k, want_v, tables, ctx, inp["scale"], 32, 8, 128, BS).numpy()
    _assert_close(out.float().cpu().numpy()[keep], want[keep], torch.bfloat16)
    # 1100 short sequences: beyond the stream kernel's batch limit -> chunk grid, same answer
    big = _random_case(79, torch.bfloat16, 8, 2, 128, 16, B=1100, max_ctx=90)
    got = _run(_dev(big, cuda_device), flags=_stream_flags()).float().cpu().numpy()
    _assert_close(got, _oracle(big).numpy(), torch.bfloat16)


def test_stream_kernel_full_size_c3_shard(cuda_device):
    """One kv head of C3 (an 8-way head shard, the shape the stream kernel is for) at full size: sample
    parity, agreement with the chunk-grid kernel, block-permutation invariance (bit-identical)."""
    from minivllm_b200 import _ext
    t = workloads.build_tables("c3")
    w = workloads.materialize(t, cuda_device, kv_heads=slice(0, 1))
    args = (w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"],
            w["KVH"], w["D"], w["BS"])
    out = paged_attention_decode(w["q"], *args, flags=_stream_flags())
    _sample_parity(t, w, out, [0, 85, 170, 255] + workloads.highest_page_sequences(t, 1))
    grid = paged_attention_decode(w["q"], *args, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SPLIT_GRID)
    assert (out.float() - grid.float()).abs().max() <= 2e-2
    perm = torch.randperm(t.num_blocks, device=cuda_device)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(t.num_blocks, device=cuda_device)
    bt = w["block_tables"].long()
    bt2 = torch.where(bt >= 0, inv[bt.clamp(min=0)], bt).to(torch.int32)
    out_p = paged_attention_decode(w["q"], w["k_cache"][perm], w["v_cache"][perm], bt2, *args[3:],
                                   flags=_stream_flags())
    assert torch.equal(out_p, out)


# ------------------------------------------------------------------ fused step (one launch)
@pytest.mark.parametrize("name", ["ragged_bf16", "long_split", "bs256_bf16", "g8_ragged", "g5_d64"])
def test_fused_step_equals_separate_launches(cuda_device, name):
    """mvl_decode_step (store + split + combine in one launch) against the three separate launches
    (store_kvcache, split kernel, combine kernel): identical cache, same output to fp32 rounding;
    three fused calls in a row (arrival counters must return to zero) give identical results."""
    from minivllm_b200 import _ext
    from minivllm_b200.layers.attention import decode_step
    inp = make_inputs(name)
    B, KVH, D, BS = inp["q"].shape[0], inp["KVH"], inp["D"], inp["BS"]
    g = torch.Generator().manual_seed(123)
    k = torch.randn(B, KVH, D, generator=g).to(inp["q"].dtype)
    v = torch.randn(B, KVH, D, generator=g).to(inp["q"].dtype)
    last = inp["context_lens"].long() - 1
    slots = (inp["block_tables"][torch.arange(B), last // BS].long() * BS + last % BS).to(inp["context_lens"].dtype)
    dev = _dev(inp, cuda_device)
    kc_a, vc_a = dev["k_cache"].clone(), dev["v_cache"].clone()
    store_kvcache(k.to(cuda_device), v.to(cuda_device), kc_a, vc_a, slots.to(cuda_device), BS)
    want = paged_attention_decode(dev["q"], kc_a, vc_a, dev["block_tables"], dev["context_lens"],
                                  inp["scale"], inp["H"], KVH, D, BS, flags=0)
    outs = []
    kc_b, vc_b = dev["k_cache"].clone(), dev["v_cache"].clone()
    for _ in range(3):
        outs.append(decode_step(dev["q"], k.to(cuda_device), v.to(cuda_device), kc_b, vc_b,
                                slots.to(cuda_device), dev["block_tables"], dev["context_lens"],
                                inp["scale"], inp["H"], KVH, D, BS).clone())
    assert torch.equal(kc_a, kc_b) and torch.equal(vc_a, vc_b)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
    a, r = TOL_EXACT[inp["q"].dtype]
    diff = (outs[0].float() - want.float()).abs() - (a + r * want.float().abs())
    assert diff.max() <= 0
    # oracle on the updated cache
    ref = c_oracle.paged_decode(inp["q"], kc_a.cpu(), vc_a.cpu(), inp["block_tables"], inp["context_lens"],
                                inp["scale"], inp["H"], KVH, D, BS)
    _assert_close(outs[0].float().cpu().numpy(), ref.numpy(), inp["q"].dtype)


def test_fused_step_skips_rows_without_a_slot(cuda_device):
    from minivllm_b200.layers.attention import decode_step
    inp = _dev(make_inputs("ragged_bf16"), cuda_device)
    B, KVH, D, BS = inp["q"].shape[0], inp["KVH"], inp["D"], inp["BS"]
    k = torch.randn(B, KVH, D, device=cuda_device).to(torch.bfloat16)
    slots = torch.full((B,), -1, dtype=inp["context_lens"].dtype, device=cuda_device)
    kc, vc = inp["k_cache"].clone(), inp["v_cache"].clone()
    out = decode_step(inp["q"], k, k, kc, vc, slots, inp["block_tables"], inp["context_lens"],
                      inp["scale"], inp["H"], KVH, D, BS)
    assert torch.equal(kc, inp["k_cache"]) and torch.equal(vc, inp["v_cache"])
    assert torch.equal(out, _run(inp))


# ------------------------------------------------------------------ TMA load path
@pytest.mark.parametrize("name", sorted(CASES))
def test_tma_load_path_matches_cp_async_bitwise(cuda_device, name):
    """MVL_DECODE_LOAD_TMA only changes how tiles reach shared memory: results must be identical
    bit for bit to the cp.async path (and so inherit its parity with the reference)."""
    from minivllm_b200 import _ext
    inp = _dev(make_inputs(name), cuda_device)
    base = _run(inp, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_CUDA_CORES)
    tma = _run(inp, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_LOAD_TMA)
    assert torch.equal(base, tma)
    tma2 = _run(inp, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_LOAD_TMA, chunk_tokens=64)
    _assert_close(tma2.float().cpu().numpy(), _oracle(make_inputs(name)).numpy(), inp["q"].dtype)


def test_tma_load_path_ignores_garbage_past_the_end(cuda_device):
    """Rows of the last page beyond context_len may hold NaN bit patterns (uninitialised cache
    memory); the TMA path loads them with the tile and must blank them."""
    from minivllm_b200 import _ext
    inp = make_inputs("ragged_bf16")
    BS = inp["BS"]
    kc, vc = inp["k_cache"].clone(), inp["v_cache"].clone()
    for b, n in enumerate(inp["context_lens"].tolist()):
        if n % BS:
            blk = int(inp["block_tables"][b, (n - 1) // BS])
            kc[blk, n % BS:] = float("nan")
            vc[blk, n % BS:] = float("nan")
    poisoned = _dev(dict(inp, k_cache=kc, v_cache=vc), cuda_device)
    clean = _dev(inp, cuda_device)
    for flags in (_ext.DECODE_FUSED_COMBINE, _ext.DECODE_FUSED_COMBINE | _ext.DECODE_LOAD_TMA):
        assert torch.equal(_run(poisoned, flags=flags), _run(clean, flags=flags))


# ------------------------------------------------------------------ tensor-core kernel (default for 16-bit caches)
def _mma_flags():
    from minivllm_b200 import _ext
    return _ext.DECODE_FUSED_COMBINE  # tensor cores are the default for 16-bit caches


def _ffma_flags():
    from minivllm_b200 import _ext
    return _ext.DECODE_FUSED_COMBINE | _ext.DECODE_CUDA_CORES


@pytest.mark.parametrize("name", sorted(n for n in CASES if CASES[n]["dtype"] != "f32"))
@pytest.mark.parametrize("chunk", [0, 64, 100000])
def test_mma_variant_matches_reference_and_oracle(cuda_device, name, chunk):
    inp = make_inputs(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    out = _run(_dev(inp, cuda_device), flags=_mma_flags(), chunk_tokens=chunk)
    got = out.float().cpu().numpy()
    assert np.abs(got - z["ref_out"]).max() <= 2e-2          # reference run in bf16/fp16
    _assert_close(got, z["ref_out_f32"], inp["q"].dtype)     # same bound as the FFMA2 kernel


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H,KVH", [(4, 4), (8, 4), (32, 8), (16, 2), (10, 2), (32, 2), (6, 2)])
@pytest.mark.parametrize("D,BS", [(128, 16), (64, 16), (128, 256), (64, 8), (128, 4)])
def test_mma_variant_random_ragged(cuda_device, dtype, H, KVH, D, BS):
    inp = _random_case(2000 + H * 7 + KVH + D + BS, dtype, H, KVH, D, BS, B=5, max_ctx=600)
    want = _oracle(inp).numpy()
    got = _run(_dev(inp, cuda_device), flags=_mma_flags()).float().cpu().numpy()
    assert np.isfinite(got).all()
    _assert_close(got, want, dtype)


def test_mma_variant_fused_step_and_fp32_fallback(cuda_device):
    from minivllm_b200.layers.attention import decode_step
    inp = make_inputs("ragged_bf16")
    B, KVH, D, BS = inp["q"].shape[0], inp["KVH"], inp["D"], inp["BS"]
    g = torch.Generator().manual_seed(5)
    k = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
    last = inp["context_lens"].long() - 1
    slots = inp["block_tables"][torch.arange(B), last // BS].long() * BS + last % BS
    dev = _dev(inp, cuda_device)
    kc, vc = dev["k_cache"].clone(), dev["v_cache"].clone()
    out = decode_step(dev["q"], k.to(cuda_device), v.to(cuda_device), kc, vc, slots.to(cuda_device),
                      dev["block_tables"], dev["context_lens"], inp["scale"], inp["H"], KVH, D, BS,
                      flags=_mma_flags())
    okc, ovc = inp["k_cache"].clone(), inp["v_cache"].clone()
    po.store_kvcache_oracle(k, v, okc, ovc, slots)
    assert torch.equal(kc.cpu().view(torch.uint8), okc.view(torch.uint8))
    want = c_oracle.paged_decode(inp["q"], okc, ovc, inp["block_tables"], inp["context_lens"],
                                 inp["scale"], inp["H"], KVH, D, BS)
    _assert_close(out.float().cpu().numpy(), want.numpy(), torch.bfloat16)
    # fp32 caches have no 16-bit MMA: they always run the FFMA2 kernel, flag or not
    f32 = _dev(make_inputs("ragged_g4"), cuda_device)
    assert torch.equal(_run(f32, flags=_ffma_flags()), _run(f32))


@pytest.mark.parametrize("name", sorted(n for n in CASES if CASES[n]["dtype"] != "f32"))
def test_cuda_core_kernel_still_selectable_for_16bit(cuda_device, name):
    """MVL_DECODE_CUDA_CORES: the FFMA2 kernel on 16-bit caches, same bound."""
    inp = make_inputs(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    got = _run(_dev(inp, cuda_device), flags=_ffma_flags()).float().cpu().numpy()
    assert np.abs(got - z["ref_out"]).max() <= 2e-2
    _assert_close(got, z["ref_out_f32"], inp["q"].dtype)


# ------------------------------------------------------------------ host-buffer step (e2e path)
@pytest.mark.parametrize("idx", [torch.int64, torch.int32])
@pytest.mark.parametrize("pieces", [1, 3, 8])
def test_host_decode_step_matches_device_path(cuda_device, idx, pieces):
    """mvl_decode_step_host (pinned H2D of the step's inputs, fused store + decode, D2H) against the
    device-resident fused step on the same inputs; two steps in a row reuse every buffer.  With
    pieces > 1 (mvl_decode_step_host_overlapped) the batch is cut into runs that overlap their copies
    with each other's kernels: same caches bit for bit, same rows to fp32 rounding (a piece is a
    smaller launch, so the split points may differ)."""
    from minivllm_b200.layers.attention import HostDecodeStep, decode_step
    inp = make_inputs("ragged_bf16")
    B, H, KVH, D, BS = inp["q"].shape[0], inp["H"], inp["KVH"], inp["D"], inp["BS"]
    kc = inp["k_cache"].to(cuda_device)
    vc = inp["v_cache"].to(cuda_device)
    kc2, vc2 = kc.clone(), vc.clone()
    hs = HostDecodeStep(kc, vc, B, inp["block_tables"].shape[1], H, inp["scale"], index_dtype=idx, pieces=pieces)
    assert hs.pieces == pieces
    g = torch.Generator().manual_seed(8)
    last = inp["context_lens"].long() - 1
    slots = inp["block_tables"][torch.arange(B), last // BS].long() * BS + last % BS
    for step in range(2):
        q = torch.randn(B, H, D, generator=g).to(torch.bfloat16)
        k = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
        v = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
        hs.q_h.copy_(q), hs.k_h.copy_(k), hs.v_h.copy_(v)
        hs.slot_h.copy_(slots.to(idx)), hs.ctx_h.copy_(inp["context_lens"].to(idx))
        hs.bt_h.copy_(inp["block_tables"])
        got = hs().clone()
        want = decode_step(q.to(cuda_device), k.to(cuda_device), v.to(cuda_device), kc2, vc2,
                           slots.to(cuda_device, idx), inp["block_tables"].to(cuda_device),
                           inp["context_lens"].to(cuda_device, idx), inp["scale"], H, KVH, D, BS)
        if pieces == 1:
            assert torch.equal(got, want.cpu())
        else:
            assert (got.float() - want.float().cpu()).abs().max() <= 2e-2
            exact = c_oracle.paged_decode(q, kc2.cpu(), vc2.cpu(), inp["block_tables"], inp["context_lens"],
                                          inp["scale"], H, KVH, D, BS)
            _assert_close(got.float().numpy(), exact.numpy(), torch.bfloat16)
        assert torch.equal(kc, kc2) and torch.equal(vc, vc2)
    assert hs.h2d_bytes == hs._h_arena.numel()


# ------------------------------------------------------------------ prefix sharing (MVL_DECODE_SHARE_PREFIX)
def _aliased_case(seed, dtype, H, KVH, D, BS, groups):
    """groups: list of (shared_tokens, [context lengths]); members of a group alias the blocks of the
    shared prefix (block aligned), each has private blocks after it."""
    g = torch.Generator().manual_seed(seed)
    rng = np.random.RandomState(seed)
    ctx = [n for _, lens in groups for n in lens]
    B = len(ctx)
    nblk = [(n + BS - 1) // BS for n in ctx]
    total = sum(nblk) + 4
    perm = list(rng.permutation(total))
    tables = np.full((B, max(nblk) + 1), -1, dtype=np.int32)
    b = 0
    for shared, lens in groups:
        sb = shared // BS
        shared_ids = [perm.pop() for _ in range(sb)]
        for n in lens:
            need = (n + BS - 1) // BS
            assert need >= sb
            tables[b, :sb] = shared_ids
            tables[b, sb:need] = [perm.pop() for _ in range(need - sb)]
            b += 1
    return dict(q=torch.randn(B, H, D, generator=g).to(dtype),
                k_cache=torch.randn(total, BS, KVH, D, generator=g).to(dtype),
                v_cache=torch.randn(total, BS, KVH, D, generator=g).to(dtype),
                block_tables=torch.from_numpy(tables), context_lens=torch.tensor(ctx, dtype=torch.int64),
                scale=1.0 / D ** 0.5, H=H, KVH=KVH, D=D, BS=BS)


@pytest.mark.parametrize("H,KVH,D", [(8, 2, 128), (4, 2, 128), (2, 2, 64), (6, 2, 128), (16, 2, 128)])
@pytest.mark.parametrize("chunk", [64, 128, 0])
def test_prefix_sharing_is_bit_identical(cuda_device, H, KVH, D, chunk):
    """Sequences aliasing a prefix scored together (one copy of the K/V tiles) must give exactly the
    bits of scoring each alone; groups of 1..5, shared lengths off the chunk grid, a member whose
    whole context is the shared part plus one token, identical twins, and G = 8 (flag ignored)."""
    from minivllm_b200 import _ext
    groups = [(256, [300, 257, 700, 290, 512]), (0, [130]), (128, [129, 129]), (384, [385, 1000, 400]),
              (64, [200]), (512, [600, 520]), (320, [330, 321, 400, 900, 350, 322, 640, 333, 500, 321])]
    inp = _aliased_case(900 + H + D, torch.bfloat16, H, KVH, D, 16, groups)
    dev = _dev(inp, cuda_device)
    base = _run(dev, chunk_tokens=chunk, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SPLIT_GRID)
    want = _oracle(inp).numpy()
    narrow = _ext.DECODE_SHARE_PREFIX | _ext.DECODE_SHARE_NARROW
    for flags in (_ext.DECODE_FUSED_COMBINE | narrow, narrow,
                  _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX, _ext.DECODE_SHARE_PREFIX):
        for _ in range(2):  # twice: the arrival counters must come back to zero
            got = _run(dev, chunk_tokens=chunk, flags=flags)
            if flags == _ext.DECODE_FUSED_COMBINE | narrow:
                # sub-groups of 8/G sequences: the same warps add the same tiles in the same order
                assert torch.equal(got, base)
            else:
                # separate combine kernel, or the group kernel (one warp walks a shared chunk's tiles
                # that four warps split otherwise): another summation order, same tolerance
                assert (got.float() - base.float()).abs().max() <= 2e-2
                _assert_close(got.float().cpu().numpy(), want, torch.bfloat16)
    _assert_close(base.float().cpu().numpy(), want, torch.bfloat16)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("H,KVH,D", [(8, 2, 128), (4, 4, 64), (6, 2, 128)])
def test_prefix_sharing_wide_form_on_tcgen05(cuda_device, dtype, H, KVH, D):
    """The shared chunks of large sub-groups on the tcgen05 group kernel (chunks of whole 128-token multiples,
    16-slot pages): fp16 and bf16, D 64 and 128, G = 4 / 1 / 3 (a spare head row per member), sub-groups of 3
    to 10 sequences with shared lengths of several chunks, against the oracle, the unshared run and the
    mma.sync group kernel; twice (counters and work list left clean)."""
    from minivllm_b200 import _ext
    groups = [(1024, [1100, 1025, 1900, 1030, 1500]), (0, [130]), (512, [513, 700, 600]), (256, [300, 290]),
              (768, [800, 769, 900, 1000, 770, 1200, 1700, 771, 905, 800])]
    inp = _aliased_case(1200 + H + D, dtype, H, KVH, D, 16, groups)
    dev = _dev(inp, cuda_device)
    want = _oracle(inp).numpy()
    base = _run(dev, chunk_tokens=128, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SPLIT_GRID)
    _assert_close(base.float().cpu().numpy(), want, dtype)
    for chunk in (128, 256):
        for flags in (_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX,
                      _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX | _ext.DECODE_GROUP_MMA,
                      _ext.DECODE_SHARE_PREFIX):
            for _ in range(2):
                got = _run(dev, chunk_tokens=chunk, flags=flags)
                assert (got.float() - base.float()).abs().max() <= 2e-2
                _assert_close(got.float().cpu().numpy(), want, dtype)


def test_prefix_sharing_tcgen05_many_small_items(cuda_device):
    """The tcgen05 group kernel with several hundred two- and four-tile work items (32k- and 16k-token shared
    prefixes cut into 128- / 256-token chunks: up to 6 items per CTA, 130-260 partials per row): its roles run
    whole items ahead of each other there — copies up to three items in front of the softmax, the finisher
    warps behind it with merges longer than an item — so every hand-over between them is under load."""
    from minivllm_b200 import _ext
    groups = [(32768, [32800, 32769, 33000, 34000, 32900, 33500, 32773, 33333]),
              (16384, [16400, 16500, 17000, 16385, 16390]), (0, [300])]
    inp = _aliased_case(77, torch.bfloat16, 8, 2, 128, 16, groups)
    dev = _dev(inp, cuda_device)
    want = _oracle(inp).numpy()
    base = _run(dev, chunk_tokens=256, flags=_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SPLIT_GRID)
    _assert_close(base.float().cpu().numpy(), want, torch.bfloat16)
    for chunk in (128, 256):
        for flags in (_ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX, _ext.DECODE_SHARE_PREFIX):
            for _ in range(2):
                got = _run(dev, chunk_tokens=chunk, flags=flags)
                assert (got.float() - base.float()).abs().max() <= 2e-2
                _assert_close(got.float().cpu().numpy(), want, torch.bfloat16)


def test_prefix_sharing_golden_and_fused_step(cuda_device):
    from minivllm_b200 import _ext
    from minivllm_b200.layers.attention import decode_step
    inp = make_inputs("aliased_bf16")   # 4 sequences sharing a 384-token prefix, G = 4
    z = np.load(os.path.join(GOLD, "aliased_bf16.npz"))
    flags = _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX
    for chunk in (64, 128, 0):
        got = _run(_dev(inp, cuda_device), flags=flags, chunk_tokens=chunk).float().cpu().numpy()
        assert np.abs(got - z["ref_out"]).max() <= 2e-2
        _assert_close(got, z["ref_out_f32"], torch.bfloat16)
    # fused store + sharing: the new token's page is private, its chunk is never scored jointly
    B, KVH, D, BS = 4, inp["KVH"], inp["D"], inp["BS"]
    g = torch.Generator().manual_seed(4)
    k = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(B, KVH, D, generator=g).to(torch.bfloat16)
    last = inp["context_lens"].long() - 1
    slots = inp["block_tables"][torch.arange(B), last // BS].long() * BS + last % BS
    dev = _dev(inp, cuda_device)
    outs = []
    for fl in (_ext.DECODE_FUSED_COMBINE, flags | _ext.DECODE_SHARE_NARROW):
        kc, vc = dev["k_cache"].clone(), dev["v_cache"].clone()
        outs.append(decode_step(dev["q"], k.to(cuda_device), v.to(cuda_device), kc, vc, slots.to(cuda_device),
                                dev["block_tables"], dev["context_lens"], inp["scale"], inp["H"], KVH, D, BS,
                                chunk_tokens=64, flags=fl))
    assert torch.equal(outs[0], outs[1])
    kc, vc = dev["k_cache"].clone(), dev["v_cache"].clone()
    wide = decode_step(dev["q"], k.to(cuda_device), v.to(cuda_device), kc, vc, slots.to(cuda_device),
                       dev["block_tables"], dev["context_lens"], inp["scale"], inp["H"], KVH, D, BS,
                       chunk_tokens=64, flags=flags)     # 4 sequences > 8/G = 2: the group kernel's case
    assert (wide.float() - outs[0].float()).abs().max() <= 2e-2

def test_random_problems_against_the_oracle(cuda_device):
    """Eighty seeded random problems through the C ABI against oracle/paged_oracle.c: group sizes 1-8, both head
    dims, page sizes from 1 slot to 256, ragged lengths with empty and one-token sequences, stale table tails,
    aliased prefixes, int32 / int64 lengths, all dtypes, every split size and kernel flag that is a no-op where
    it does not apply."""
    from minivllm_b200 import _ext
    rng = np.random.RandomState(20261107)
    flag_sets = [0, _ext.DECODE_FUSED_COMBINE, _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX,
                 _ext.DECODE_SHARE_PREFIX, _ext.DECODE_CUDA_CORES, _ext.DECODE_CUDA_CORES | _ext.DECODE_FUSED_COMBINE,
                 _ext.DECODE_CUDA_CORES | _ext.DECODE_LOAD_TMA, _ext.DECODE_FUSED_COMBINE | _ext.DECODE_STREAM,
                 _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX | _ext.DECODE_GROUP_MMA,
                 _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX | _ext.DECODE_SHARE_NARROW]
    for case in range(80):
        B, KVH = int(rng.randint(1, 10)), int(rng.randint(1, 5))
        G = int(rng.choice([1, 2, 3, 4, 6, 8]))
        D = int(rng.choice([64, 128]))
        BS = int(rng.choice([1, 2, 8, 16, 16, 32, 256]))
        dtype = [torch.float32, torch.bfloat16, torch.float16][int(rng.randint(3))]
        H = KVH * G
        hi = int(rng.choice([40, 300, 1500]))
        ctx = rng.randint(0, hi + 1, size=B)
        if case % 4 == 0:
            ctx[rng.randint(B)] = 1
        nblk = (ctx + BS - 1) // BS
        width = int(nblk.max()) + int(rng.randint(0, 3)) + 1
        pool = int(nblk.sum()) + 2
        perm = rng.permutation(pool)
        tables = np.full((B, width), -1, dtype=np.int32)
        at = 0
        for b in range(B):
            tables[b, :nblk[b]] = perm[at:at + nblk[b]]
            at += nblk[b]
        if case % 3 == 0 and B > 1:          # aliased prefixes: later sequences start with the first one's whole blocks
            for b in range(1, B):
                share = min(int(ctx[0]) // BS, max(int(ctx[b]) - 1, 0) // BS)
                tables[b, :share] = tables[0, :share]
        if case % 2:                         # stale entries behind a sequence's last block
            for b in range(B):
                tables[b, nblk[b]:] = rng.randint(0, pool, size=width - nblk[b])
        g = torch.Generator().manual_seed(case)
        inp = dict(q=torch.randn(B, H, D, generator=g).to(dtype),
                   k_cache=torch.randn(pool, BS, KVH, D, generator=g).to(dtype),
                   v_cache=torch.randn(pool, BS, KVH, D, generator=g).to(dtype),
                   block_tables=torch.from_numpy(tables),
                   context_lens=torch.from_numpy(ctx.astype(np.int64 if case % 5 else np.int32)),
                   scale=float(rng.choice([1.0 / D ** 0.5, 0.05, 0.3])), H=H, KVH=KVH, D=D, BS=BS)
        want = _oracle(inp).numpy()
        flags = flag_sets[case % len(flag_sets)]
        chunk = int(rng.choice([0, 64, 128, 256, 1024]))
        what = f"case {case}: B={B} KVH={KVH} G={G} D={D} BS={BS} {dtype} ctx={ctx.tolist()} flags={flags} chunk={chunk}"
        dev = _dev(inp, cuda_device)
        for _ in range(2):                   # twice: counters and work lists must be left clean
            got = _run(dev, flags=flags, chunk_tokens=chunk).float().cpu().numpy()
            assert np.isfinite(got).all(), what
            a, r = TOL_EXACT[dtype]
            excess = np.abs(got.astype(np.float64) - want) - (a + r * np.abs(want))
            assert excess.max() <= 0, f"{what}: max abs err {np.abs(got - want).max():.3e}"
            assert not got[ctx == 0].any(), what
