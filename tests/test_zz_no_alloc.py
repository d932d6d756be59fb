"""GPU (-m gpu), collected last: mvl_paged_decode_static, the entry point that allocates nothing.

Who it is for: a caller that cannot give the library a workspace — an engine whose KV cache already
owns all free HBM, a binding that holds device pointers but no allocator, a loop that must never
hit cudaMalloc's device-wide synchronisation.  Its split-KV partials live in a fixed 16 MiB array
inside the library instead.

What is checked, through the C ABI (ctypes, raw pointers):
  * two calls in a row (different shapes and dtypes, so the second one finds the first one's
    leftovers in the shared array) each match the ordinary entry point, mvl_paged_decode with a
    caller-owned workspace, within the tolerance the parity tests use, and the reference's own
    golden outputs for the same two cases;
  * between the two calls neither the device's free memory nor torch's allocator moves;
  * a problem whose partials exceed the array is refused with MVL_ERR_LIMIT and a message that
    states both sizes, and nothing is written.
The file is self-contained so it passes alone in a fresh process as well as after the others.
"""
import os

import numpy as np
import pytest
import torch

from cases import make_inputs
from minivllm_b200 import _ext
from minivllm_b200.layers.attention import paged_attention_decode

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")

# same bound as tests/test_gpu_parity.py: |err| <= A + R * |want|
TOL = {torch.float32: (1e-4, 0.0), torch.bfloat16: (1e-3, 2.0 ** -8), torch.float16: (2e-4, 2.0 ** -11)}
CODE = {torch.float32: _ext.F32, torch.float16: _ext.F16, torch.bfloat16: _ext.BF16}


def _on(inp, dev):
    return {k: (v.to(dev).contiguous() if isinstance(v, torch.Tensor) else v) for k, v in inp.items()}


def _static_call(inp, out, chunk=0):
    """mvl_paged_decode_static with raw pointers; the caller owns inputs and output."""
    bt, cl = inp["block_tables"], inp["context_lens"]
    return _ext.lib().mvl_paged_decode_static(
        out.data_ptr(), inp["q"].data_ptr(), inp["k_cache"].data_ptr(), inp["v_cache"].data_ptr(),
        bt.data_ptr(), cl.data_ptr(), int(cl.dtype == torch.int64), inp["q"].shape[0], bt.shape[1],
        bt.stride(0), inp["H"], inp["KVH"], inp["D"], inp["BS"], inp["k_cache"].shape[0],
        float(inp["scale"]), CODE[inp["q"].dtype], chunk,
        torch.cuda.current_stream().cuda_stream)


def _workspace_bytes(inp, chunk=0):
    return _ext.lib().mvl_paged_decode_workspace_bytes(
        inp["q"].shape[0], inp["H"], inp["KVH"], inp["D"], inp["block_tables"].shape[1], inp["BS"],
        CODE[inp["q"].dtype], chunk)


def _close(got, want, dtype):
    got = got.float().cpu().numpy().astype(np.float64)
    want = want.float().cpu().numpy().astype(np.float64)
    a, r = TOL[dtype]
    return float((np.abs(got - want) - (a + r * np.abs(want))).max()) <= 0.0


def test_static_entry_matches_workspace_entry_twice_in_a_row(cuda_device):
    lib = _ext.lib()
    assert lib.mvl_static_scratch_bytes() == 16 << 20
    first = _on(make_inputs("long_split"), cuda_device)    # fp32, 5000 / 3333 tokens, many chunks
    second = _on(make_inputs("ragged_bf16"), cuda_device)  # bf16, 8 ragged sequences
    for inp in (first, second):                            # both really use the scratch
        assert 0 < _workspace_bytes(inp) <= lib.mvl_static_scratch_bytes()

    want_first = paged_attention_decode(first["q"], first["k_cache"], first["v_cache"],
                                        first["block_tables"], first["context_lens"], first["scale"],
                                        first["H"], first["KVH"], first["D"], first["BS"])
    want_second = paged_attention_decode(second["q"], second["k_cache"], second["v_cache"],
                                         second["block_tables"], second["context_lens"],
                                         second["scale"], second["H"], second["KVH"], second["D"],
                                         second["BS"])
    out_first = torch.full_like(first["q"], float("nan"))
    out_second = torch.full_like(second["q"], float("nan"))
    torch.cuda.synchronize()

    # call 1 (also the one-time address look-up on this device, outside any capture)
    assert _static_call(first, out_first) == 0, lib.mvl_static_last_error().decode()
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(cuda_device)
    held_before = torch.cuda.memory_allocated(cuda_device)
    # call 2, straight after, other shape and dtype: the array still holds call 1's partials
    assert _static_call(second, out_second) == 0, lib.mvl_static_last_error().decode()
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(cuda_device)
    assert free_after == free_before and torch.cuda.memory_allocated(cuda_device) == held_before

    assert _close(out_first, want_first, torch.float32)
    assert _close(out_second, want_second, torch.bfloat16)
    # and against the reference's own outputs for these two cases (tests/golden/make_golden.py):
    # 1e-4 abs with both sides fp32, 2e-2 abs against the reference run in bf16
    ref_first = np.load(os.path.join(GOLD, "long_split.npz"))["ref_out"]
    ref_second = np.load(os.path.join(GOLD, "ragged_bf16.npz"))["ref_out"]
    assert np.abs(out_first.float().cpu().numpy() - ref_first).max() <= 1e-4
    assert np.abs(out_second.float().cpu().numpy() - ref_second).max() <= 2e-2
    # and once more with the first problem, now over call 2's leftovers
    out_again = torch.full_like(first["q"], float("nan"))
    assert _static_call(first, out_again) == 0
    torch.cuda.synchronize()
    assert _close(out_again, want_first, torch.float32)


def test_static_entry_refuses_what_does_not_fit(cuda_device):
    lib = _ext.lib()
    # 256 sequences x 32 heads x 16 chunks of partials (64k-token tables) need ~68 MB > 16 MiB
    B, H, KVH, D, BS, W = 256, 32, 8, 128, 16, 4096
    q = torch.zeros(B, H, D, dtype=torch.bfloat16, device=cuda_device)
    kc = torch.zeros(4, BS, KVH, D, dtype=torch.bfloat16, device=cuda_device)
    inp = dict(q=q, k_cache=kc, v_cache=kc, H=H, KVH=KVH, D=D, BS=BS, scale=1.0,
               block_tables=torch.zeros(B, W, dtype=torch.int32, device=cuda_device),
               context_lens=torch.ones(B, dtype=torch.int64, device=cuda_device))
    need = _workspace_bytes(inp)
    assert need > lib.mvl_static_scratch_bytes()
    out = torch.full_like(q, 3.0)
    assert _static_call(inp, out) == -7  # MVL_ERR_LIMIT
    msg = lib.mvl_static_last_error().decode()
    assert str(need) in msg and str(16 << 20) in msg
    torch.cuda.synchronize()
    assert bool((out == 3.0).all())  # nothing was launched
    with pytest.raises(ValueError):
        _ext.check(-7)
    # the same batch fits once a chunk covers every sequence (one chunk -> no partials at all)
    assert _workspace_bytes(inp, chunk=W * BS) == 0
    assert _static_call(inp, out, chunk=W * BS) == 0
    torch.cuda.synchronize()
