"""CPU: the oracles against the golden vectors written by the reference's own function.

Golden outputs come from reference benchmark_decoding.py:136-189 (decode_torch_optimized), run by
tests/golden/make_golden.py.  Tolerances: fp32 1e-5 abs (same math, different summation order);
16-bit dtypes 2e-2 abs for the in-dtype port (BASELINE.json north_star), 1e-5 for the exact
oracles against the reference fed fp32-upcast inputs.
"""
import os

import numpy as np
import pytest
import torch

from cases import CASES, input_digest, make_inputs
from oracle import c_oracle
from oracle import paged_attention_oracle as po

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return z["ref_out"], z["ref_out_f32"], float(z["naive_maxdiff"]), str(z["digest"])


def _args(inp):
    return (inp["q"], inp["k_cache"], inp["v_cache"], inp["block_tables"], inp["context_lens"],
            inp["scale"], inp["H"], inp["KVH"], inp["D"], inp["BS"])


@pytest.mark.parametrize("name", sorted(CASES))
def test_inputs_reproduce(name):
    inp = make_inputs(name)
    assert input_digest(inp) == _load(name)[3], "seeded input generator drifted from the goldens"


@pytest.mark.parametrize("name", sorted(CASES))
def test_reference_variants_agree(name):
    assert _load(name)[2] <= 1e-6  # decode_torch_optimized vs naive_decode_attention


@pytest.mark.parametrize("name", sorted(CASES))
def test_padded_port_matches_reference(name):
    inp = make_inputs(name)
    ref_out, ref_f32, _, _ = _load(name)
    got = po.decode_attention_padded(*_args(inp)).float().numpy()
    tol = 1e-5 if CASES[name]["dtype"] == "f32" else 2e-2
    assert np.abs(got - ref_out).max() <= tol
    up = dict(inp, q=inp["q"].float(), k_cache=inp["k_cache"].float(), v_cache=inp["v_cache"].float())
    got32 = po.decode_attention_padded(*_args(up)).numpy()
    assert np.abs(got32 - ref_f32).max() <= 1e-5


@pytest.mark.parametrize("name", sorted(CASES))
def test_exact_oracles_match_reference(name):
    inp = make_inputs(name)
    ref_f32 = _load(name)[1]
    exact = po.decode_attention_exact(*_args(inp))
    assert np.abs(exact - ref_f32).max() <= 1e-5
    c_out = c_oracle.paged_decode(*_args(inp)).numpy()
    assert np.abs(c_out - ref_f32).max() <= 1e-5
    assert np.abs(c_out - exact).max() <= 2e-6


def test_context_len_zero_is_zero_row():
    inp = make_inputs("g1_d64")
    inp["context_lens"] = inp["context_lens"].clone()
    inp["context_lens"][1] = 0
    out = c_oracle.paged_decode(*_args(inp))
    assert torch.count_nonzero(out[1]) == 0
    assert np.abs(po.decode_attention_exact(*_args(inp))[1]).max() == 0


def test_context_len_one_returns_value_row():
    inp = make_inputs("ragged_g4")  # sequence 0 has one token
    out = c_oracle.paged_decode(*_args(inp))
    blk = int(inp["block_tables"][0, 0])
    group = inp["H"] // inp["KVH"]
    for h in range(inp["H"]):
        assert torch.equal(out[0, h], inp["v_cache"][blk, 0, h // group].float())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_store_oracles_follow_kernel_index_math(dtype):
    # reference attention.py:33-34,52-56: block = slot // BS, offset = slot % BS
    g = torch.Generator().manual_seed(5)
    NB, BS, KVH, D, N = 7, 16, 2, 64, 9
    key = torch.randn(N, KVH, D, generator=g).to(dtype)
    val = torch.randn(N, KVH, D, generator=g).to(dtype)
    slots = torch.tensor([3, -1, 111, 0, 16, -1, 15, 64, 90], dtype=torch.int64)
    base_k = torch.randn(NB, BS, KVH, D, generator=g).to(dtype)
    base_v = torch.randn(NB, BS, KVH, D, generator=g).to(dtype)
    want_k, want_v = base_k.clone(), base_v.clone()
    for t, s in enumerate(slots.tolist()):
        if s == -1:
            continue
        want_k[s // BS, s % BS] = key[t]
        want_v[s // BS, s % BS] = val[t]
    k1, v1 = base_k.clone(), base_v.clone()
    po.store_kvcache_oracle(key, val, k1, v1, slots)
    assert torch.equal(k1.view(torch.uint8), want_k.view(torch.uint8))
    assert torch.equal(v1.view(torch.uint8), want_v.view(torch.uint8))
    k2, v2 = base_k.clone(), base_v.clone()
    c_oracle.store_kvcache(key, val, k2, v2, slots)
    assert torch.equal(k2.view(torch.uint8), want_k.view(torch.uint8))
    assert torch.equal(v2.view(torch.uint8), want_v.view(torch.uint8))


# ------------------------------------------------------------------ pre-attention pass (N4)
import hashlib  # noqa: E402

from cases_pre import PRE_CASES, make_pre_inputs, pre_digest  # noqa: E402
from oracle import pre_attention_oracle as pre  # noqa: E402


@pytest.mark.parametrize("name", sorted(PRE_CASES))
def test_pre_attention_oracle_matches_reference_modules(name):
    """oracle/pre_attention_oracle.py against LayerNorm.rms_forward + RotaryEmbedding.forward of the
    reference (goldens written by tests/golden/make_golden_pre.py).  fp32 cases: 1e-5; the bf16
    case: the reference rounds after every op, the oracle computes exactly -> 2e-2."""
    c = PRE_CASES[name]
    z = np.load(os.path.join(GOLD, f"pre_{name}.npz"))
    inp = make_pre_inputs(name)
    assert pre_digest(inp) == str(z["digest"])
    table = pre.rope_table(c["base"], c["D"], c["max_position"])
    assert hashlib.sha256(table.numpy().tobytes()).hexdigest() == str(z["table_digest"])
    qn, kn = pre.pre_attention(inp["q"], inp["k"], inp["v"], inp["positions"],
                               inp["gamma"] if c["norm"] else None, c["eps"], table, None, None, None)
    tol = 1e-5 if c["dtype"] == "f32" else 2e-2
    assert np.abs(qn.numpy() - z["q_ref"]).max() <= tol * max(1.0, np.abs(z["q_ref"]).max())
    assert np.abs(kn.numpy() - z["k_ref"]).max() <= tol * max(1.0, np.abs(z["k_ref"]).max())


# ------------------------------------------------------------------ prefill attention (N3)
from cases_prefill import PREFILL_CASES, make_prefill_inputs, prefill_digest  # noqa: E402
from oracle import prefill_oracle  # noqa: E402


@pytest.mark.parametrize("name", sorted(PREFILL_CASES))
def test_prefill_oracle_matches_reference(name):
    """oracle/prefill_oracle.py against the reference's pytorch_standard_attention
    (benchmark_prefilling.py:9-52): 1e-5 vs its fp32 run, 2e-2 vs its run in the 16-bit dtype."""
    z = np.load(os.path.join(GOLD, f"prefill_{name}.npz"))
    inp = make_prefill_inputs(name)
    assert prefill_digest(inp) == str(z["digest"])
    got = prefill_oracle.causal_varlen_attention(inp["q"], inp["k"], inp["v"], inp["cu_seqlens"], inp["scale"],
                                                 inp["H"], inp["KVH"], inp["D"]).numpy()
    assert np.abs(got[::3] - z["ref_out_f32"]).max() <= 1e-5
    native = torch.from_numpy(z["ref_out_bits"]).view(inp["q"].dtype).float().numpy()
    assert np.abs(got - native).max() <= 2e-2


def _paged_case(seed, lens_ctx, lens_new, H, KVH, D, BS, dtype=torch.float32):
    """Random sequences laid out in a paged cache (random page order) + the queries of their last
    lens_new tokens.  Returns everything both prefill oracles need."""
    g = torch.Generator().manual_seed(seed)
    pages = [(c + BS - 1) // BS for c in lens_ctx]
    nb = sum(pages) + 3
    order = torch.randperm(nb, generator=g)
    width = max(pages) + 1
    bt = torch.full((len(lens_ctx), width), -1, dtype=torch.int32)
    kc = torch.zeros(nb, BS, KVH, D, dtype=dtype)
    vc = torch.zeros_like(kc)
    k_full, v_full, q_new, at = [], [], [], 0
    for s, (c, n) in enumerate(zip(lens_ctx, lens_new)):
        bt[s, :pages[s]] = order[at:at + pages[s]].to(torch.int32)
        at += pages[s]
        k = torch.randn(c, KVH, D, generator=g).to(dtype)
        v = torch.randn(c, KVH, D, generator=g).to(dtype)
        pos = torch.arange(c)
        slots = bt[s, pos // BS].long() * BS + pos % BS
        kc.view(-1, KVH, D)[slots] = k
        vc.view(-1, KVH, D)[slots] = v
        k_full.append(k), v_full.append(v)
        q_new.append(torch.randn(n, H, D, generator=g).to(dtype))
    cu_q = torch.tensor([0] + list(np.cumsum(lens_new)), dtype=torch.int32)
    return dict(q=torch.cat(q_new), kc=kc, vc=vc, bt=bt, ctx=torch.tensor(lens_ctx, dtype=torch.int32), cu_q=cu_q,
                k_full=k_full, v_full=v_full, q_new=q_new)


@pytest.mark.parametrize("BS", [16, 256])
def test_paged_prefill_oracle_equals_dense_tail(BS):
    """causal_paged_attention (no reference implementation exists: the reference's prefill ignores the
    pages, A-8) is anchored on causal_varlen_attention, which IS pinned to the reference: queries at the
    last n positions of a sequence give the last n rows of the dense causal attention over all ctx
    tokens, whatever the queries at earlier positions are.  With no prefix the two coincide."""
    H, KVH, D = 4, 2, 32
    lens_ctx, lens_new = [40, 300, 17, 129], [40, 45, 1, 128]     # first sequence: no cached prefix
    c = _paged_case(7, lens_ctx, lens_new, H, KVH, D, BS)
    got = prefill_oracle.causal_paged_attention(c["q"], c["kc"], c["vc"], c["cu_q"], c["bt"], c["ctx"], 0.2, H, KVH,
                                                D, BS)
    for s, (ctx, n) in enumerate(zip(lens_ctx, lens_new)):
        q_full = torch.zeros(ctx, H, D)
        q_full[ctx - n:] = c["q_new"][s]
        dense = prefill_oracle.causal_varlen_attention(q_full, c["k_full"][s], c["v_full"][s],
                                                       torch.tensor([0, ctx]), 0.2, H, KVH, D)
        b = int(c["cu_q"][s])
        assert (got[b:b + n] - dense[ctx - n:]).abs().max() < 1e-12
