"""Golden vectors for prefill attention from the REFERENCE's own PyTorch implementation
(pytorch_standard_attention, /root/reference/benchmark_prefilling.py:9-52), run on CPU in the build
container.  Writes tests/golden/prefill_<case>.npz: ref_out_bits (the reference run in the case
dtype, raw 16-bit patterns), ref_out_f32 (reference fed fp32-upcast inputs, every third token),
digest of the seeded inputs.
    python tests/golden/make_golden_prefill.py"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.environ.get("MVL_REFERENCE", "/root/reference"))
import benchmark_prefilling as ref  # noqa: E402
from cases_prefill import PREFILL_CASES, make_prefill_inputs, prefill_digest  # noqa: E402

for name in PREFILL_CASES:
    inp = make_prefill_inputs(name)
    args = (inp["cu_seqlens"], inp["scale"], inp["H"], inp["KVH"], inp["D"])
    native = ref.pytorch_standard_attention(inp["q"], inp["k"], inp["v"], *args)
    f32 = ref.pytorch_standard_attention(inp["q"].float(), inp["k"].float(), inp["v"].float(), *args)
    # ref_out: the reference's 16-bit result, stored as its raw bits (exact, half the size);
    # ref_out_f32: every third token of the fp32 run
    np.savez_compressed(os.path.join(HERE, f"prefill_{name}.npz"),
                        ref_out_bits=native.contiguous().view(torch.int16).numpy(),
                        ref_out_f32=f32[::3].numpy(), digest=np.array(prefill_digest(inp)))
    print(f"{name:12s} {tuple(native.shape)} {native.dtype}")
