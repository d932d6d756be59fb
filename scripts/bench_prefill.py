"""Throughput of the prefill kernel (causal, varlen) at a few sizes; FLOPs = 4 * D * H * sum(n_i^2)/2."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200 import _ext
from minivllm_b200.layers.attention import flash_attention_prefill
FLAGS = _ext.PREFILL_MMA_SYNC if "--mma-sync" in sys.argv else 0
KERNEL = "flash_prefill_kernel (mma.sync)" if FLAGS else "flash_prefill_tcgen05_kernel"
dev = torch.device("cuda", 0)
peak = json.load(open("MEASURED_PEAKS.json"))["bf16_tflops"] if os.path.exists("MEASURED_PEAKS.json") else 1590.0
H, KVH, D = 32, 8, 128
for nseq, n in ((8, 1024), (4, 4096), (2, 16384)):
    T = nseq * n
    q = torch.randn(T, H, D, device=dev).to(torch.bfloat16)
    k = torch.randn(T, KVH, D, device=dev).to(torch.bfloat16)
    v = torch.randn_like(k)
    cu = torch.arange(0, T + 1, n, dtype=torch.int32, device=dev)
    run = lambda: flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n, flags=FLAGS)
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flops = 4.0 * D * H * nseq * n * (n + 1) / 2
    print(json.dumps({"kernel": KERNEL, "seqs": nseq, "len": n, "shape": f"{H}Q/{KVH}KV/D{D} bf16",
                      "ms": ms, "TFLOPs": flops / ms / 1e9, "frac_of_measured_bf16_peak": flops / ms / 1e9 / peak}))
