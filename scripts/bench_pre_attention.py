"""Micro-benchmark of the fused pre-attention pass (q/k RMSNorm + RoPE + KV store) on one GPU:
prefill-sized (bandwidth-bound) and decode-sized (launch-bound) inputs, CUDA-event timing.
Prints one JSON line per size; algorithmic bytes = read q,k,v once + write q', k', v' once
(+ positions, slots, and the rotation-table rows)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from minivllm_b200.layers.pre_attention import fused_qk_norm_rope_store, rope_cache  # noqa: E402

dev = torch.device("cuda", 0)
H, KVH, D, BS = 32, 8, 128, 16
peak = 6555.8
if os.path.exists("MEASURED_PEAKS.json"):
    peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
table = rope_cache(1e6, D, 32768, device=dev)
gamma = torch.ones(D, device=dev)
for n in (256, 4096, 65536, 262144):
    qkv = torch.randn(n, (H + 2 * KVH) * D, device=dev).to(torch.bfloat16)
    q, k, v = (t.view(n, -1, D) for t in qkv.split([H * D, KVH * D, KVH * D], dim=-1))
    nb = (n + BS - 1) // BS
    kc = torch.zeros(nb, BS, KVH, D, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    slots = torch.randperm(nb * BS, device=dev)[:n]
    pos = torch.randint(0, 32768, (n,), device=dev)
    q_out = torch.empty(n, H, D, dtype=torch.bfloat16, device=dev)
    run = lambda: fused_qk_norm_rope_store(q, k, v, pos, gamma, 1e-5, table, kc, vc, slots, q_out=q_out)
    for _ in range(5):
        run()
    reps = 50
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    nbytes = 2 * n * (H + 2 * KVH) * D * 2 + n * (8 + 8) + n * D * 4
    print(json.dumps({"kernel": "qk_norm_rope_store_token_kernel" if n >= 8192 else "qk_norm_rope_store_kernel", "tokens": n, "shape": f"{H}Q/{KVH}KV/D{D} bf16",
                      "ms": ms, "algorithmic_bytes": nbytes, "GBps": nbytes / ms / 1e6,
                      "frac_of_measured_peak": nbytes / ms / 1e6 / peak}))
