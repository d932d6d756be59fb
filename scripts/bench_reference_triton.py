"""Secondary GPU baseline: the reference's own Triton kernels — the only
pre-existing GPU implementation of the path — timed on the same B200 next to ours, on the same
synthetic batches.  The reference is the unmodified package installed into baseline/_ref (git-ignored;
`pip install --no-index --no-deps --ignore-requires-python --target baseline/_ref <copy of /root/reference>`,
its pyproject pins python < 3.12).  Nothing here is on the product path."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import torch
import workloads
from minivllm_b200.layers import attention as ours

try:
    from myvllm.layers import attention as ref
except Exception as exc:   # not installed on this box
    print(json.dumps({"unavailable": repr(exc)}))
    sys.exit(0)
dev = torch.device("cuda", 0)


def timed(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


# c4 is left out by default: the reference kernel faults there (illegal memory access; its offsets are
# 32-bit and C4's caches hold 6.5e9 elements each), which poisons the CUDA context of this process
names = sys.argv[1:] or ["c1", "c2", "c3", "prefill"]
for name in [n for n in names if n != "prefill"]:
    sp = workloads.SPECS[name]
    t = workloads.build_tables(name)
    w = workloads.materialize(t, dev)
    args = (w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], sp.H, sp.KVH, sp.D,
            sp.BS)
    alg = workloads.algorithmic_bytes(t, kvh_local=sp.KVH, h_local=sp.H, unique=True)
    try:
        o_ref = ref.paged_attention_decode(*args)
        ms_ref = timed(lambda: ref.paged_attention_decode(*args), reps=5)
    except Exception as exc:
        print(json.dumps({"workload": name, "reference_triton": "failed: " + repr(exc)[:200]}))
        continue
    o = ours.paged_attention_decode(*args)
    ms = timed(lambda: ours.paged_attention_decode(*args))
    err = (o.float() - o_ref.float()).abs().max().item()
    print(json.dumps({"workload": name, "batch": int(w["q"].shape[0]), "dtype": sp.dtype,
                      "reference_triton_ms": ms_ref, "reference_triton_GBps": alg / ms_ref / 1e6,
                      "ours_ms": ms, "ours_GBps": alg / ms / 1e6, "speedup": ms_ref / ms,
                      "max_abs_diff_vs_reference_triton": err}))

# ---- prefill: the reference's flash_attention_prefill (Triton) vs ours
H, KVH, D = 32, 8, 128
for nseq, n in (((8, 1024), (4, 4096)) if "prefill" in names else ()):
    T = nseq * n
    g = torch.Generator(device=dev).manual_seed(n)
    q = torch.randn(T, H, D, device=dev, generator=g).to(torch.bfloat16)
    k = torch.randn(T, KVH, D, device=dev, generator=g).to(torch.bfloat16)
    v = torch.randn(T, KVH, D, device=dev, generator=g).to(torch.bfloat16)
    cu = torch.arange(0, T + 1, n, dtype=torch.int32, device=dev)
    flops = 4.0 * D * H * nseq * n * (n + 1) / 2
    try:
        o_ref = ref.flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D)
        ms_ref = timed(lambda: ref.flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D), reps=5)
    except Exception as exc:
        print(json.dumps({"workload": f"prefill {nseq}x{n}", "reference_triton": "failed: " + repr(exc)[:200]}))
        continue
    o = ours.flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n)
    ms = timed(lambda: ours.flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n))
    print(json.dumps({"workload": f"prefill {nseq} x {n} tokens, 32Q/8KV/D128 bf16", "reference_triton_ms": ms_ref,
                      "reference_triton_TFLOPs": flops / ms_ref / 1e9, "ours_ms": ms, "ours_TFLOPs": flops / ms / 1e9,
                      "speedup": ms_ref / ms,
                      "max_abs_diff_vs_reference_triton": (o.float() - o_ref.view_as(o).float()).abs().max().item()}))
