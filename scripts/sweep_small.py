"""Decode-attention latency at the reference's own sweep sizes (reference benchmark_models.py:
batch {1..16} x context {512..16384}, Qwen3-8B shape 32Q/8KV/D128, block 16), bf16.
GPU time per call measured under CUDA-graph replay (20 calls per graph) so host launch cost does
not hide the kernel; prints per-size microseconds, algorithmic GB/s and the split size used."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from minivllm_b200 import _ext  # noqa: E402
from minivllm_b200.layers.attention import decode_step  # noqa: E402

dev = torch.device("cuda", 0)
H, KVH, D, BS = 32, 8, 128, 16
# arguments: split sizes (0 = the library's choice), or "grid" / "stream" / "pdl" to force the chunk grid, the
# persistent stream kernel, or to add MVL_DECODE_OVERLAP_PREV, with the library's split size
chunks = [c if c in ("grid", "stream", "pdl") else int(c) for c in sys.argv[1:]] or [0]
FL = {"grid": _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SPLIT_GRID, "stream": _ext.DECODE_FUSED_COMBINE | _ext.DECODE_STREAM,
      "pdl": _ext.DECODE_FUSED_COMBINE | _ext.DECODE_OVERLAP_PREV}
for B in (1, 4, 16, 64):
    for ctx in (512, 4096, 16384, 65536):
        if B * ctx > 2 ** 21:
            continue
        nblk = (ctx + BS - 1) // BS
        kc = torch.randn(B * nblk, BS, KVH, D, device=dev).to(torch.bfloat16)
        vc = torch.randn_like(kc)
        q = torch.randn(B, H, D, device=dev).to(torch.bfloat16)
        k = torch.randn(B, KVH, D, device=dev).to(torch.bfloat16)
        v = torch.randn_like(k)
        tables = torch.randperm(B * nblk, device=dev).to(torch.int32).view(B, nblk)
        cl = torch.full((B,), ctx, dtype=torch.int64, device=dev)
        slots = tables[:, -1].long() * BS + (ctx - 1) % BS
        out = torch.empty_like(q)
        nbytes = B * ctx * KVH * D * 2 * 2 + 2 * B * H * D * 2
        row = {"B": B, "ctx": ctx, "MB": round(nbytes / 1e6, 1)}
        for chunk in chunks:
            named = isinstance(chunk, str)
            call = lambda: decode_step(q, k, v, kc, vc, slots, tables, cl, D ** -0.5, H, KVH, D, BS, out=out,
                                       chunk_tokens=0 if named else chunk, flags=FL[chunk] if named else None)
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(20):
                    call()
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1000 / 200
            used = chunk if named else (chunk or "auto")
            row[f"{used}"] = f"{us:.1f}us {nbytes / us / 1e3:.0f}GB/s"
        print(json.dumps(row))
