"""Throughput of the prefill kernel (causal, varlen) at a few sizes; FLOPs = 4 * D * H * sum(n_i^2)/2."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200 import _ext
for a in sys.argv:   # --lib=NAME: use minivllm_b200/csrc/NAME (experiment builds)
    if a.startswith("--lib="):
        _ext.LIB_PATH = os.path.join(os.path.dirname(_ext.LIB_PATH), a[6:])
SIZES = ((2, 16384),) if "--long" in sys.argv else ((8, 1024), (4, 4096), (2, 16384))
from minivllm_b200.layers.attention import flash_attention_prefill
FLAGS = 1 if "--mma-sync" in sys.argv else 2 if "--single" in sys.argv else 4 if "--pair" in sys.argv else 0
KERNEL = "flash_prefill_f32_kernel" if "--fp32" in sys.argv else {0: "auto", 1: "flash_prefill_kernel (mma.sync)", 2: "flash_prefill_tcgen05_kernel", 4: "flash_prefill_tcgen05_pair_kernel"}[FLAGS]
dev = torch.device("cuda", 0)
peak = json.load(open("MEASURED_PEAKS.json"))["bf16_tflops"] if os.path.exists("MEASURED_PEAKS.json") else 1590.0
H, KVH = 32, 8
DT = torch.float16 if "--fp16" in sys.argv else torch.float32 if "--fp32" in sys.argv else torch.bfloat16
D = 64 if "--d64" in sys.argv else 128
for nseq, n in SIZES:
    T = nseq * n
    q = torch.randn(T, H, D, device=dev).to(DT)
    k = torch.randn(T, KVH, D, device=dev).to(DT)
    v = torch.randn_like(k)
    cu = torch.arange(0, T + 1, n, dtype=torch.int32, device=dev)
    run = lambda: flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n, flags=FLAGS)
    for _ in range(3): run()
    if "--clocks" in sys.argv:   # sustained run with clock / power samples
        import subprocess, threading, time
        stop, samples = [False], []
        def poll():
            while not stop[0]:
                o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True).stdout.strip()
                samples.append(o); time.sleep(0.02)
        th = threading.Thread(target=poll); th.start()
        t0 = time.time()
        while time.time() - t0 < 1.5: run()
        torch.cuda.synchronize(); stop[0] = True; th.join()
        print("clock/power samples:", samples[len(samples)//2:len(samples)//2+4])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flops = 4.0 * D * H * nseq * n * (n + 1) / 2
    print(json.dumps({"kernel": KERNEL, "seqs": nseq, "len": n, "shape": f"{H}Q/{KVH}KV/D{D} " + {torch.float16: "fp16", torch.float32: "fp32", torch.bfloat16: "bf16"}[DT],
                      "ms": ms, "TFLOPs": flops / ms / 1e9, "frac_of_measured_bf16_peak": flops / ms / 1e9 / peak}))
