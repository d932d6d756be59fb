"""Generate tests/golden/*.npz from the REFERENCE's own PyTorch paged attention.

Run in the build container only (needs /root/reference, which is absent on the GPU box):
    python tests/golden/make_golden.py
Imports decode_torch_optimized and naive_decode_attention from
/root/reference/benchmark_decoding.py (:136-189, :192-269), runs them on CPU on the seeded inputs
of tests/golden/cases.py and stores, per case:
    ref_out        decode_torch_optimized in the case dtype (as float32 values)
    ref_out_f32    the same function fed the inputs upcast to float32 ("both sides fp32")
    naive_maxdiff  max |decode_torch_optimized - naive_decode_attention| (cross-check, fp32)
    digest         SHA-256 of the regenerated input bytes
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.environ.get("MVL_REFERENCE", "/root/reference"))

import benchmark_decoding as ref  # noqa: E402  (the reference; Triton decorators are lazy)
from cases import CASES, input_digest, make_inputs  # noqa: E402


def main():
    torch.manual_seed(0)
    for name in CASES:
        inp = make_inputs(name)
        args = (inp["block_tables"], inp["context_lens"], inp["scale"], inp["H"], inp["KVH"], inp["D"],
                inp["BS"])
        native = ref.decode_torch_optimized(inp["q"], inp["k_cache"], inp["v_cache"], *args)
        up = [inp[k].float() for k in ("q", "k_cache", "v_cache")]
        f32 = ref.decode_torch_optimized(*up, *args)
        naive = ref.naive_decode_attention(*up, *args)
        maxdiff = float((f32 - naive).abs().max())
        np.savez_compressed(os.path.join(HERE, name + ".npz"), ref_out=native.float().numpy(),
                            ref_out_f32=f32.numpy(), naive_maxdiff=np.float64(maxdiff),
                            digest=np.array(input_digest(inp)))
        print(f"{name:18s} out {tuple(native.shape)} {native.dtype}  optimized-vs-naive {maxdiff:.2e}")


if __name__ == "__main__":
    main()
