"""store_kvcache at prefill sizes: bytes = read k, v once + write them once (+ slots)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200.layers.attention import store_kvcache
dev = torch.device("cuda", 0)
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6555.8
KVH, D, BS = 8, 128, 16
for n in (256, 4096, 32768, 262144, 1048576):
    k = torch.randn(n, KVH, D, device=dev).to(torch.bfloat16)
    v = torch.randn_like(k)
    kc = torch.zeros(n // BS + 1, BS, KVH, D, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    slots = torch.randperm(n, device=dev)
    run = lambda: store_kvcache(k, v, kc, vc, slots, BS)
    for _ in range(5): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    nbytes = 4 * n * KVH * D * 2 + 8 * n
    print(json.dumps({"kernel": "store_kvcache_kernel", "tokens": n, "us": ms * 1e3, "GBps": nbytes / ms / 1e6,
                      "frac_of_measured_peak": nbytes / ms / 1e6 / peak}))
