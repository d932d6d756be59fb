"""Throughput of prefill over cached pages (mvl_flash_prefill_paged): chunked prefill of long prompts
and prompts with a cached prefix.  FLOPs = 4 * D * H * sum over queries of the keys each one sees."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200.layers.attention import flash_attention_prefill_paged
dev = torch.device("cuda", 0)
peak = json.load(open("MEASURED_PEAKS.json"))["bf16_tflops"] if os.path.exists("MEASURED_PEAKS.json") else 1590.0
H, KVH, D = 32, 8, 128
# (sequences, context after the step, new tokens per sequence, block size)
for nseq, ctx, new, BS in ((2, 16384, 16384, 256), (2, 16384, 16384, 16), (4, 16384, 2048, 256), (4, 16384, 2048, 16),
                           (8, 8192, 1024, 16), (16, 4096 + 512, 512, 16)):
    pages = (ctx + BS - 1) // BS
    nb = nseq * pages
    kc = torch.randn(nb, BS, KVH, D, device=dev).to(torch.bfloat16)
    vc = torch.randn_like(kc)
    bt = torch.randperm(nb, device=dev).to(torch.int32).view(nseq, pages)
    q = torch.randn(nseq * new, H, D, device=dev).to(torch.bfloat16)
    cu = torch.arange(0, nseq * new + 1, new, dtype=torch.int32, device=dev)
    cl = torch.full((nseq,), ctx, dtype=torch.int32, device=dev)
    FLAGS = 4 if '--pair' in sys.argv else 2 if '--single' in sys.argv else 0
    run = lambda: flash_attention_prefill_paged(q, kc, vc, cu, bt, cl, D ** -0.5, H, KVH, D, BS, max_seqlen_q=new, flags=FLAGS)
    for _ in range(3): run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    keys = sum(ctx - new + i + 1 for i in range(new))
    flops = 4.0 * D * H * nseq * keys
    print(json.dumps({"kernel": "flash_prefill_tcgen05 (paged, %s)" % {0: "auto", 2: "one tile", 4: "pair"}[FLAGS], "seqs": nseq, "context": ctx, "new_tokens": new,
                      "block_size": BS, "shape": f"{H}Q/{KVH}KV/D{D} bf16", "ms": ms, "TFLOPs": flops / ms / 1e9,
                      "frac_of_measured_bf16_peak": flops / ms / 1e9 / peak}))
