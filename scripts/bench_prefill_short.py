"""Prefill on short prompts: where per-CTA setup and wave quantisation show (one to eight key tiles)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200.layers.attention import flash_attention_prefill
dev = torch.device("cuda", 0)
FLAGS = 1 if "--mma-sync" in sys.argv else 2 if "--single" in sys.argv else 4 if "--pair" in sys.argv else 0
H, KVH, D = 32, 8, 128
for nseq, n in ((64, 128), (32, 256), (16, 512), (8, 1024), (32, 1024), (16, 2048), (4, 4096), (2, 16384)):
    T = nseq * n
    q = torch.randn(T, H, D, device=dev).to(torch.bfloat16)
    k = torch.randn(T, KVH, D, device=dev).to(torch.bfloat16)
    v = torch.randn_like(k)
    cu = torch.arange(0, T + 1, n, dtype=torch.int32, device=dev)
    run = lambda: flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n, flags=FLAGS)
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    flops = 4.0 * D * H * nseq * n * (n + 1) / 2
    print(json.dumps({"flags": FLAGS, "seqs": nseq, "len": n, "us": ms * 1e3, "TFLOPs": flops / ms / 1e9}))
