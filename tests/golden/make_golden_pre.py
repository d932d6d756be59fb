"""Golden vectors for the pre-attention pass from the REFERENCE's own modules (build container only):
LayerNorm.rms_forward (reference src/myvllm/layers/layernorm.py:9-17) and RotaryEmbedding.forward
(src/myvllm/layers/rotary_embedding.py:74-83), run eagerly on CPU (TORCHDYNAMO_DISABLE=1 turns the
reference's @torch.compile decorators into plain calls).  Writes tests/golden/pre_<case>.npz:
q_ref, k_ref (float32 values of the reference's outputs) and the digest of the seeded inputs.
    TORCHDYNAMO_DISABLE=1 python tests/golden/make_golden_pre.py
"""
import hashlib
import os
import sys

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
import numpy as np  # noqa: E402
import torch  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(os.environ.get("MVL_REFERENCE", "/root/reference"), "src"))

from myvllm.layers.layernorm import LayerNorm  # noqa: E402
from myvllm.layers.rotary_embedding import RotaryEmbedding  # noqa: E402
from cases_pre import PRE_CASES, make_pre_inputs, pre_digest  # noqa: E402


def main():
    for name, c in PRE_CASES.items():
        inp = make_pre_inputs(name)
        norm = LayerNorm(inp["gamma"].clone(), eps=c["eps"])
        rope = RotaryEmbedding(base=c["base"], rotary_embedding=c["D"], max_position=c["max_position"])
        q, k = inp["q"], inp["k"]
        if c["norm"]:
            q, k = norm(q), norm(k)
        q, k = rope(inp["positions"], q, k)
        table = rope.cos_sin_cache.numpy()
        np.savez_compressed(os.path.join(HERE, f"pre_{name}.npz"), q_ref=q.float().numpy(),
                            k_ref=k.float().numpy(), digest=np.array(pre_digest(inp)),
                            table_digest=np.array(hashlib.sha256(table.tobytes()).hexdigest()),
                            table_shape=np.array(table.shape))
        print(f"{name:14s} q {tuple(q.shape)} {q.dtype}")


if __name__ == "__main__":
    main()
