"""GPU (-m gpu): our kernels against the reference's OWN GPU kernels (Triton) on the same inputs, when
the unmodified reference package is present under baseline/_ref (git-ignored install, see DESIGN.md
section 5; skipped otherwise).  The CPU goldens pin the oracle; this pins the product against the
reference's GPU implementation directly, at north_star's tolerances."""
import os
import sys

import pytest
import torch

from cases import make_inputs

pytestmark = pytest.mark.gpu
_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")


@pytest.fixture(scope="module")
def ref_attention():
    if not os.path.isdir(os.path.join(_REF, "myvllm")):
        pytest.skip("reference package not installed under baseline/_ref")
    sys.path.insert(0, _REF)
    try:
        from myvllm.layers import attention as ref
    except Exception as exc:  # triton / transformers import problems on this box
        pytest.skip(f"reference package does not import here: {exc!r}")
    finally:
        sys.path.remove(_REF)
    return ref


@pytest.mark.parametrize("name", ["c1_bench_shape", "ragged_bf16", "ragged_f16", "long_split", "g8_ragged", "g5_d64", "bs256_bf16", "aliased_bf16"])
def test_decode_and_store_match_reference_triton(cuda_device, ref_attention, name):
    from minivllm_b200.layers.attention import paged_attention_decode, store_kvcache
    try:
        inp = make_inputs(name)
    except KeyError:
        pytest.skip(f"no golden case {name}")
    dev = cuda_device
    args = (inp["q"].to(dev), inp["k_cache"].to(dev), inp["v_cache"].to(dev), inp["block_tables"].to(dev),
            inp["context_lens"].to(dev), inp["scale"], inp["H"], inp["KVH"], inp["D"], inp["BS"])
    want = ref_attention.paged_attention_decode(*args)
    got = paged_attention_decode(*args)
    tol = 1e-4 if inp["q"].dtype == torch.float32 else 2e-2
    assert (got.float() - want.float()).abs().max().item() <= tol
    # store: bit-exact against the reference's Triton store kernel
    B, KVH, D, BS = inp["q"].shape[0], inp["KVH"], inp["D"], inp["BS"]
    g = torch.Generator().manual_seed(1)
    k = torch.randn(B, KVH, D, generator=g).to(inp["q"].dtype).to(dev)
    v = torch.randn(B, KVH, D, generator=g).to(inp["q"].dtype).to(dev)
    last = inp["context_lens"].long() - 1
    slots = (inp["block_tables"][torch.arange(B), last // BS].long() * BS + last % BS).to(dev)
    ka, va = inp["k_cache"].to(dev).clone(), inp["v_cache"].to(dev).clone()
    kb, vb = ka.clone(), va.clone()
    ref_attention.store_kvcache(k, v, ka, va, slots, BS)
    store_kvcache(k, v, kb, vb, slots, BS)
    assert torch.equal(ka.view(torch.uint8), kb.view(torch.uint8)) and torch.equal(va.view(torch.uint8), vb.view(torch.uint8))


def test_prefill_matches_reference_triton(cuda_device, ref_attention):
    from minivllm_b200.layers.attention import flash_attention_prefill
    H, KVH, D = 8, 2, 128
    g = torch.Generator().manual_seed(2)
    lens = [300, 77, 513]
    T = sum(lens)
    q = torch.randn(T, H, D, generator=g).to(torch.bfloat16).to(cuda_device)
    k = torch.randn(T, KVH, D, generator=g).to(torch.bfloat16).to(cuda_device)
    v = torch.randn(T, KVH, D, generator=g).to(torch.bfloat16).to(cuda_device)
    cu = torch.tensor([0, 300, 377, 890], dtype=torch.int32, device=cuda_device)
    want = ref_attention.flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D)
    for flags in (0, 1, 2, 4):
        got = flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, flags=flags)
        assert (got.float() - want.view_as(got).float()).abs().max().item() <= 2e-2
