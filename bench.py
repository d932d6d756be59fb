#!/usr/bin/env python
"""bench.py — throughput of the paged-KV decode-attention hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1)

A "step" is ONE decode-attention call of one layer over one synthetic batch:
store_kvcache of the batch's new tokens, then paged decode attention (split kernel + combine)
[+ NCCL all-gather of the per-rank heads when N > 1].  tokens/s = batch / step time, the
definition of reference benchmark_models.py:161.  Default workload: BASELINE.json configs[2]
("c3": Qwen3-8B shape, batch 256, contexts 1k-8k, bf16) — the config the north star's roofline
target is quoted on.  N > 1 runs the SAME batch sharded by KV head (strong scaling).

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      dominant kernel (the split-KV decode kernel): algorithmic bytes per launch /
                CUDA-event duration of that launch, against MEASURED_PEAKS.json
  cpu_baseline  the oracle port of the reference's vectorised-PyTorch CPU path, bounded sample
  e2e           same metric through the host-buffer C-ABI entry (pinned H2D + D2H every step)
--impl reference times that CPU path alone (rank 0; other ranks exit 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import workloads  # noqa: E402

METRIC = "decode attn tokens/s (HBM GB/s and % of B200 peak in roofline)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--workload", default="c3", choices=sorted(workloads.SPECS) + ["c2e"],
                    help="c1..c5: one attention call per step; c2e: 28-layer engine-shaped decode "
                         "loop of config 2 under CUDA-graph replay")
    ap.add_argument("--eager", action="store_true", help="c2e: no CUDA graphs")
    ap.add_argument("--share-prefix", choices=["auto", "on", "off", "narrow", "mma"], default="auto",
                    help="score sequences with aliased block-table prefixes together "
                         "(MVL_DECODE_SHARE_PREFIX); auto = on for the shared-prefix workload c5")
    ap.add_argument("--cuda-cores", action="store_true",
                    help="16-bit caches: packed-fp32 FFMA2 kernel instead of the tensor-core one "
                         "(MVL_DECODE_CUDA_CORES)")
    ap.add_argument("--tma", action="store_true",
                    help="fetch K/V tiles with tensor-map TMA instead of cp.async (MVL_DECODE_LOAD_TMA)")
    ap.add_argument("--unfused", action="store_true",
                    help="three launches per step (store, split, combine) instead of the fused one")
    ap.add_argument("--shard-of", type=int, default=1,
                    help="tuning aid: run rank 0's KV-head shard of an N-way split on ONE GPU, no gather")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--chunk", type=int, default=0, help="split size override (tokens)")
    ap.add_argument("--cpu-seqs", type=int, default=8, help="sequences in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-neighbours", action="store_true",
                    help="skip the prefill / pre-attention kernels measured after the headline (N=1, c3 only)")
    ap.add_argument("--kernel", default="auto", choices=["auto", "grid", "stream"],
                    help="tensor-core decode kernel: one CTA per chunk (grid), the persistent evenly cut "
                         "stream (stream), or the library's own choice (auto)")
    ap.add_argument("--no-overlap", action="store_true",
                    help="plain stream order between steps (default: MVL_DECODE_OVERLAP_PREV, a step's kernel "
                         "starts up while the previous step's finishes, as between the layers of a decode step)")
    ap.add_argument("--sustain-s", type=float, default=2.0,
                    help="seconds of back-to-back steps for the 'sustained' sub-record (0 = skip)")
    ap.add_argument("--parity-seqs", type=int, default=4, help="sequences checked against the oracle after timing")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N>1: how the per-rank heads are gathered (fused peer stores, or NCCL)")
    return ap.parse_args()


def workload_config(args, t, world):
    """The `config` object of the JSON line: what was measured, identical for both arms (--impl ours /
    reference).  How each arm ran it goes under `implementation`."""
    sp = t.spec
    unique = workloads.algorithmic_bytes(t, unique=True)
    return {"workload": args.workload, "shape": f"{sp.H}Q/{sp.KVH}KV/D{sp.D}", "batch": len(t.ctx),
            "block_size": sp.BS, "context_tokens": int(t.ctx.sum()),
            "context_range": [int(t.ctx.min()), int(t.ctx.max())],
            "parallelism": f"kv-head-shard x{world}" if args.shard_of == 1 else
            f"rank 0's shard of a {args.shard_of}-way head split, alone on one GPU",
            "l2": "inputs larger than L2 (KV bytes per step >> 126 MB); no flush needed"
                  if unique / max(world, args.shard_of) > 4 * 126e6 else "KV fits in L2: number is L2-assisted"}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def reference_triton_decode(w, alg_bytes, our_ms):
    """Secondary GPU baseline: the reference's own Triton decode kernel — the
    only pre-existing GPU implementation of the path — on this step's tensors, when the unmodified
    reference package is installed under baseline/_ref (git-ignored).  Not the reference arm (that is
    the CPU path, --impl reference); reported beside the headline, 2 warm-up + 5 timed launches."""
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_root, "myvllm")):
        return {"unavailable": "baseline/_ref/myvllm not installed"}
    sys.path.insert(0, ref_root)
    try:
        from myvllm.layers import attention as ref
    finally:
        sys.path.remove(ref_root)
    call = lambda: ref.paged_attention_decode(w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"],
                                              w["scale"], w["H"], w["KVH"], w["D"], w["BS"])
    for _ in range(2):
        o_ref = call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    from minivllm_b200.layers.attention import paged_attention_decode
    ours = paged_attention_decode(w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"],
                                  w["H"], w["KVH"], w["D"], w["BS"])
    return {"kernel": "myvllm.layers.attention.paged_attention_decode (Triton, JIT for this GPU)", "ms_per_step": ms,
            "GBps": alg_bytes / ms / 1e6, "ours_faster_by": ms / our_ms,
            "max_abs_diff_ours_vs_reference_triton": float((ours.float() - o_ref.float()).abs().max())}


def neighbour_kernels(dev):
    """The kernels either side of the hot path, each timed alone
    after the headline measurement: 3 warm-up + 10 timed launches, CUDA events, inputs larger than L2.
    Reported next to the headline line, not part of it."""
    from minivllm_b200.layers.attention import flash_attention_prefill, flash_attention_prefill_paged
    from minivllm_b200.layers.pre_attention import fused_qk_norm_rope_store, rope_cache
    tf_burst, tf_sust = 1590.0, 1400.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            mp = json.load(f)
        tf_burst, tf_sust = float(mp.get("bf16_tflops", tf_burst)), float(mp.get("bf16_tflops_sustained", tf_sust))
    hbm, _ = measured_peaks()
    H, KVH, D = 32, 8, 128

    def timed(run):
        for _ in range(3):
            run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        e0.record()
        for _ in range(10):
            run()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / 10

    out = {"shape": f"{H}Q/{KVH}KV/D{D} bf16", "timing": "3 warm-up + 10 launches, CUDA events",
           "tensor_peak_tflops": {"burst": tf_burst, "sustained": tf_sust, "source": "MEASURED_PEAKS.json (cuBLAS bf16)"}}
    # prefill, dense: 2 sequences x 16384 tokens, causal
    nseq, n = 2, 16384
    q = torch.randn(nseq * n, H, D, device=dev).to(torch.bfloat16)
    k = torch.randn(nseq * n, KVH, D, device=dev).to(torch.bfloat16)
    v = torch.randn_like(k)
    cu = torch.arange(0, nseq * n + 1, n, dtype=torch.int32, device=dev)
    ms = timed(lambda: flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n))
    fl = 4.0 * D * H * nseq * n * (n + 1) / 2
    out["prefill"] = {"kernel": "flash_prefill_tcgen05_pair_kernel", "workload": "2 x 16384 tokens, causal",
                      "ms": ms, "tflops": fl / ms / 1e9, "frac_of_burst_peak": fl / ms / 1e9 / tf_burst}
    # prefill over cached pages: 4 sequences, 2048 new tokens each on top of a 14336-token cached prefix
    nseq, ctx, new, BS = 4, 16384, 2048, 16
    pages = ctx // BS
    kc = torch.randn(nseq * pages, BS, KVH, D, device=dev).to(torch.bfloat16)
    vc = torch.randn_like(kc)
    bt = torch.randperm(nseq * pages, device=dev).to(torch.int32).view(nseq, pages)
    qn = q[:nseq * new]
    cuq = torch.arange(0, nseq * new + 1, new, dtype=torch.int32, device=dev)
    cl = torch.full((nseq,), ctx, dtype=torch.int32, device=dev)
    ms = timed(lambda: flash_attention_prefill_paged(qn, kc, vc, cuq, bt, cl, D ** -0.5, H, KVH, D, BS, max_seqlen_q=new))
    fl = 4.0 * D * H * nseq * sum(ctx - new + i + 1 for i in range(new))
    out["prefill_cached_prefix"] = {"kernel": "flash_prefill_tcgen05_pair_kernel (paged K/V)",
                                    "workload": "4 x 2048 new tokens over 14336 cached, 16-slot pages", "ms": ms,
                                    "tflops": fl / ms / 1e9, "frac_of_burst_peak": fl / ms / 1e9 / tf_burst}
    del kc, vc, bt, k, v
    # pre-attention pass: q/k RMSNorm + RoPE + KV store over 262144 tokens
    nt = 262144
    qkv = torch.randn(nt, (H + 2 * KVH) * D, device=dev).to(torch.bfloat16)
    qq, kk, vv = (x.view(nt, -1, D) for x in qkv.split([H * D, KVH * D, KVH * D], dim=-1))
    kc = torch.zeros(nt // 16, 16, KVH, D, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    slots = torch.randperm(nt, device=dev)
    pos = torch.randint(0, 32768, (nt,), device=dev)
    table = rope_cache(1e6, D, 32768, device=dev)
    gamma = torch.ones(D, device=dev)
    q_out = torch.empty(nt, H, D, dtype=torch.bfloat16, device=dev)
    ms = timed(lambda: fused_qk_norm_rope_store(qq, kk, vv, pos, gamma, 1e-5, table, kc, vc, slots, q_out=q_out))
    nbytes = 2 * nt * (H + 2 * KVH) * D * 2 + nt * 16 + nt * D * 4   # in + out + (pos, slot) + table rows
    out["pre_attention"] = {"kernel": "qk_norm_rope_store_token_kernel", "workload": "262144 tokens", "ms": ms,
                            "GBps": nbytes / ms / 1e6, "frac_of_hbm_peak": nbytes / ms / 1e6 / hbm}
    return out


# ------------------------------------------------------------------ clocks during the timed region
class ClockSampler:
    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.01)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons - {"gpu_idle"}), "samples": len(self.samples)}


# ------------------------------------------------------------------ CPU baseline (oracle port)
def cpu_sample(t: workloads.Tables, nseq: int):
    """A bounded sample of the workload on CPU: ``nseq`` evenly strided sequences with a compacted
    block pool (only the blocks those sequences reference)."""
    sp = t.spec
    B = len(t.ctx)
    pick = np.unique(np.linspace(0, B - 1, min(nseq, B)).astype(np.int64))
    ctx = t.ctx[pick]
    nblk = (ctx + sp.BS - 1) // sp.BS
    used = np.unique(np.concatenate([t.tables[b, :n] for b, n in zip(pick, nblk)]))
    remap = np.full(t.num_blocks, -1, dtype=np.int32)
    remap[used] = np.arange(len(used), dtype=np.int32)
    width = int(nblk.max())
    tables = np.full((len(pick), width), -1, dtype=np.int32)
    for i, (b, n) in enumerate(zip(pick, nblk)):
        tables[i, :n] = remap[t.tables[b, :n]]
    dt = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}[sp.dtype]
    g = torch.Generator().manual_seed(sp.seed)
    q = torch.randn(len(pick), sp.H, sp.D, generator=g).to(dt)
    kc = torch.randn(len(used), sp.BS, sp.KVH, sp.D, generator=g).to(dt)
    vc = torch.randn(len(used), sp.BS, sp.KVH, sp.D, generator=g).to(dt)
    return dict(q=q, k_cache=kc, v_cache=vc, block_tables=torch.from_numpy(tables),
                context_lens=torch.from_numpy(ctx), scale=1.0 / sp.D ** 0.5, H=sp.H, KVH=sp.KVH,
                D=sp.D, BS=sp.BS, n=len(pick), tokens=int(ctx.sum()))


def cpu_step(s):
    from oracle import paged_attention_oracle as po  # the checker, used here only as CPU baseline
    return po.decode_attention_padded(s["q"], s["k_cache"], s["v_cache"], s["block_tables"],
                                      s["context_lens"], s["scale"], s["H"], s["KVH"], s["D"], s["BS"])


def time_cpu(s, steps: int, warmup: int, budget_s: float = 1e9, warm_budget_s: float = 1e9):
    torch.set_num_threads(os.cpu_count() or 1)
    t0 = time.perf_counter()
    for _ in range(warmup):
        cpu_step(s)
        if time.perf_counter() - t0 > warm_budget_s:
            break
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        cpu_step(s)
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    return (time.perf_counter() - t0) / done, done


def run_reference(args, rank):
    if rank != 0:
        return
    t = workloads.build_tables(args.workload)
    s = cpu_sample(t, args.cpu_seqs)
    warm = max(args.warmup, 0)
    per_step, done = time_cpu(s, args.steps, warm, budget_s=240.0, warm_budget_s=120.0)
    value = s["n"] / per_step
    sample = (f"{s['n']} of {len(t.ctx)} sequences of workload {args.workload} (evenly strided, "
              f"{s['tokens']} context tokens) per step; {done} steps")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": done, "warmup": warm, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": t.spec.dtype, "data": "synthetic",
        "config": workload_config(args, t, args.gpus),
        "implementation": {"path": "oracle port of decode_torch_optimized (vectorised PyTorch, CPU), unsharded, "
                                   "rank 0's host cores"},
        "cpu_baseline": {"value": value, "unit": "tokens/s", "cores": torch.get_num_threads(),
                         "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ the CUDA arm
def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist

    from minivllm_b200 import _ext
    from minivllm_b200.head_shard import gather_heads, shard_plan
    from minivllm_b200.layers.attention import (HostDecodeStep, decode_step, paged_attention_decode,
                                                store_kvcache)

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    t = workloads.build_tables(args.workload)
    sp = t.spec
    sh = shard_plan(sp.H, sp.KVH, world, rank)
    if world == 1 and args.shard_of > 1:
        sh = shard_plan(sp.H, sp.KVH, args.shard_of, 0)
    w = workloads.materialize(t, dev, kv_heads=sh.kv_heads, seed_offset=rank)
    B = w["B"]
    out = torch.empty_like(w["q"])
    staging = torch.empty((world, B, w["H"], w["D"]), dtype=out.dtype, device=dev) if world > 1 else None
    fused, gather_kind = None, "none"
    if args.unfused and world > 1:
        args.gather = "nccl"   # the peer-store gather signals once per step: one launch per step only
    if world > 1 and args.gather == "peer":
        try:
            from minivllm_b200.head_shard import PeerGatheredDecode
            fused = PeerGatheredDecode(sh, B, sp.H, sp.D, out.dtype, dev)
            gather_kind = ("peer stores from the decode kernel, step announced from its epilogue, one-warp wait "
                           "kernel (mvl_decode_step_peers_sync + mvl_peer_wait); two result buffers in turns")
        except Exception as exc:  # symmetric memory unavailable on this box: say so, use NCCL
            gather_kind = f"NCCL all-gather (peer-memory setup failed: {type(exc).__name__})"
        ok = torch.tensor([1 if fused is not None else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok) == 0:
            fused = None
    elif world > 1:
        gather_kind = "NCCL all-gather + permute copy"
    dec_args = (w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"],
                w["KVH"], w["D"], w["BS"])
    lib = _ext.lib()
    # auto: on for the shared-prefix workload (also on head shards: with the tcgen05 group kernel a one-kv-head
    # shard of C5 runs 0.104 ms shared against 0.127 ms unshared; the mma.sync group kernel lost there, 0.181)
    share = args.share_prefix in ("on", "narrow", "mma") or (args.share_prefix == "auto" and args.workload == "c5")
    tma = ((_ext.DECODE_LOAD_TMA if args.tma else 0) | (_ext.DECODE_CUDA_CORES if args.cuda_cores else 0) |
           (_ext.DECODE_SHARE_PREFIX if share else 0) |
           (_ext.DECODE_SHARE_NARROW if args.share_prefix == "narrow" else 0) |
           (_ext.DECODE_GROUP_MMA if args.share_prefix == "mma" else 0) |
           {"auto": 0, "grid": _ext.DECODE_SPLIT_GRID, "stream": _ext.DECODE_STREAM}[args.kernel] |
           (0 if args.no_overlap else _ext.DECODE_OVERLAP_PREV))
    if tma:
        from minivllm_b200.layers import attention as att_mod
        att_mod.DEFAULT_DECODE_FLAGS |= tma
    code = {"f32": 0, "f16": 1, "bf16": 2}[sp.dtype]
    chunk = args.chunk or lib.mvl_paged_decode_default_chunk(B, w["H"], w["KVH"], w["D"],
                                                             t.tables.shape[1], w["BS"], code)
    multi_split = t.tables.shape[1] * sp.BS > chunk

    fargs = (w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["BS"])
    new_kv = (w["k"], w["v"], w["slot_mapping"])

    def step(ev=None):
        """One decode step of a layer.  Default: ONE launch (store + split + combine fused,
        mvl_decode_step).  --unfused: the three separate launches, for profiling comparisons."""
        if args.unfused:
            store_kvcache(w["k"], w["v"], w["k_cache"], w["v_cache"], w["slot_mapping"], w["BS"])
            if ev is not None:
                ev[0].record()
            if fused is not None:
                fused(*fargs, chunk_tokens=args.chunk, flags=_ext.DECODE_SKIP_COMBINE | tma, barrier=False)
            else:
                paged_attention_decode(w["q"], *dec_args, out=out, chunk_tokens=args.chunk,
                                       flags=_ext.DECODE_SKIP_COMBINE | tma)
            if ev is not None:
                ev[1].record()
            if multi_split:
                if fused is not None:
                    fused(*fargs, chunk_tokens=args.chunk, flags=_ext.DECODE_ONLY_COMBINE, barrier=False)
                else:
                    paged_attention_decode(w["q"], *dec_args, out=out, chunk_tokens=args.chunk,
                                           flags=_ext.DECODE_ONLY_COMBINE)
        else:
            if ev is not None:
                ev[0].record()
            if fused is not None:
                fused(*fargs, chunk_tokens=args.chunk, barrier=False, store=new_kv)
            else:
                decode_step(w["q"], w["k"], w["v"], w["k_cache"], w["v_cache"], w["slot_mapping"],
                            w["block_tables"], w["context_lens"], w["scale"], w["H"], w["KVH"], w["D"],
                            w["BS"], out=out, chunk_tokens=args.chunk)
            if ev is not None:
                ev[1].record()
        if fused is not None:
            fused.barrier()
            return fused.out
        if world > 1:
            return gather_heads(out, staging=staging)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    barrier()
    K = args.steps
    # The dominant kernel's own duration comes from an event pair around its launch inside the timed
    # region.  At N = 1 every step carries one; at N > 1, where a step is ~100 us and two extra stream
    # operations per step are not free (N = 8: ~10 us of a 130 us step), every 8th step does.
    stride = 1 if world == 1 else min(8, max(1, K // 3))
    kev = {i: (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for i in range(0, K, stride)}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        e0.record()
        for i in range(K):
            step(kev.get(i))
        e1.record()
        barrier()
    ms_total = e0.elapsed_time(e1)
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kev.values()]))
    if world > 1:
        red = torch.tensor([ms_total, kern_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        ms_total, kern_ms = red.tolist()
    ms_step = ms_total / K
    value = B / (ms_step * 1e-3)

    # ---- sustained sub-measurement: back-to-back steps for >= sustain_s seconds, its own clocks
    sustained = None
    if args.sustain_s > 0:
        n_sus = max(int(args.sustain_s * 1e3 / ms_step) + 1, K)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(local_rank) as sclocks:
            barrier()
            s0.record()
            for _ in range(n_sus):
                step()
            s1.record()
            barrier()
        sus_ms = s0.elapsed_time(s1) / n_sus
        if world > 1:
            red = torch.tensor([sus_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(red, op=dist.ReduceOp.MAX)
            sus_ms = float(red[0])
        sustained = {"steps": n_sus, "seconds": sus_ms * n_sus / 1e3, "ms_per_step": sus_ms,
                     "value": B / (sus_ms * 1e-3), "unit": "tokens/s", "vs_burst": ms_step / sus_ms,
                     "clocks": sclocks.summary()}

    # ---- output check: the last step's result against the oracle on sampled sequences (every rank
    # checks the heads it computed; worst error over ranks), and at N > 1 the gathered buffer
    # against an NCCL all-gather of the local results, bit for bit.
    parity = None
    if args.parity_seqs > 0:
        from oracle import c_oracle  # the checker; nothing timed goes through it
        res = step()
        torch.cuda.synchronize()
        if world > 1:
            barrier()
        local = res[:, sh.q_heads] if world > 1 else res
        pick = sorted(set(np.linspace(0, B - 1, args.parity_seqs).astype(int).tolist() +
                          workloads.highest_page_sequences(t, 1)))
        smp = workloads.oracle_sample(t, w, pick)
        want = c_oracle.paged_decode(smp["q"], smp["k_cache"], smp["v_cache"], smp["block_tables"],
                                     smp["context_lens"], smp["scale"], smp["H"], smp["KVH"], smp["D"], smp["BS"])
        err = float((local[pick].float().cpu() - want).abs().max())
        tol = 1e-4 if sp.dtype == "f32" else 2e-2
        gather_ok = None
        if world > 1:
            mine = local.contiguous()
            ref_gather = gather_heads(mine)
            gather_ok = bool(torch.equal(ref_gather, res))
            red = torch.tensor([err, 0.0 if gather_ok else 1.0], device=dev, dtype=torch.float64)
            dist.all_reduce(red, op=dist.ReduceOp.MAX)
            err, gather_ok = float(red[0]), bool(red[1] == 0.0)
        parity = {"max_abs_err": err, "tol": tol, "ok": bool(err <= tol), "sequences": pick,
                  "heads": "every rank checks its own heads; max over ranks" if world > 1 else "all",
                  "checked_against": "oracle/paged_oracle.c (float64 accumulation) on the rows' own pages",
                  "gather_verified": gather_ok}

    # ---- end-to-end through the host-buffer C-ABI entry (pinned H2D + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        hs = HostDecodeStep(w["k_cache"], w["v_cache"], B, t.tables.shape[1], w["H"], w["scale"],
                            index_dtype=torch.int64, chunk_tokens=args.chunk, flags=tma)
        hs.q_h.copy_(w["q"].cpu()), hs.k_h.copy_(w["k"].cpu()), hs.v_h.copy_(w["v"].cpu())
        hs.slot_h.copy_(w["slot_mapping"].cpu()), hs.ctx_h.copy_(w["context_lens"].cpu())
        hs.bt_h.copy_(w["block_tables"].cpu())
        for _ in range(3):
            hs()
        ke = max(K // 4, min(K, 10))
        barrier()
        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        checksum = 0.0
        h0.record()
        for _ in range(ke):
            res = hs()                       # synchronises: the host reads the result every step
            checksum += float(res[0, 0, 0])
        h1.record()
        barrier()
        e2e_ms = h0.elapsed_time(h1) / ke
        if world > 1:
            red = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(red, op=dist.ReduceOp.MAX)
            e2e_ms = float(red[0])
        e2e = {"value": B / (e2e_ms * 1e-3), "unit": "tokens/s", "ms_per_step": e2e_ms, "steps": ke,
               "h2d_bytes_per_step": hs.h2d_bytes, "d2h_bytes_per_step": hs.d2h_bytes,
               "api": "HostDecodeStep -> mvl_decode_step_host (per rank: its head shard); k, v, slots, lengths and "
                      "tables in one pinned H2D copy, " + ("q read and result rows written" if hs.direct_q else
                      "q in that copy too (prefix sharing: no PCIe round trip per work item), result rows written") +
                      " by the kernel across PCIe (device-mapped pinned memory), stream-synchronised every step"}

    # a peer wait that gave up (a lost announcement) would show as a multi-second step: say so explicitly
    wait_timeouts = None
    if fused is not None:
        red = torch.tensor([float(fused.timed_out())], device=dev, dtype=torch.float64)
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        wait_timeouts = int(red[0])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    alg = workloads.algorithmic_bytes(t, kvh_local=w["KVH"], h_local=w["H"])
    alg_unique = workloads.algorithmic_bytes(t, kvh_local=w["KVH"], h_local=w["H"], unique=True)
    achieved = alg_unique / (kern_ms * 1e-3) / 1e9
    # a prefix-sharing step is four launches: the pre-pass (lcp_kernel, group_kernel), the kernel of the shared
    # chunks and the split kernel; its roofline figure is the step's (unique bytes over all of them)
    wide_share = share and args.share_prefix != "narrow"
    roof = {"bound": "hbm", "kernel": ("lcp_kernel + group_kernel + " + (
                "paged_decode_group_kernel (mma.sync)" if args.share_prefix == "mma" else "paged_decode_group_tc5_kernel (tcgen05)")
                + " + " if wide_share else "lcp_kernel + group_kernel + " if share else "") + "paged_decode_split_kernel" + (
                "" if args.unfused else " (store and combine fused in)"), "achieved": achieved, "peak": peak,
            "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
            "frac_note": ("frac > 1: the measured peak is a read+write copy (MEASURED_PEAKS.json), this kernel almost only "
                          "reads; frac_of_nominal_8TBs and read_only_stream_gbs are the sanity figures"
                          if achieved > peak else None),
            "algorithmic_bytes_per_launch": alg_unique, "logical_bytes_per_launch": alg,
            "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms / ms_step,
            "kernel_ms_samples": len(kev),
            "frac_of_nominal_8TBs": achieved / 8000.0}
    # measured DRAM bytes per launch exist only for the exact (workload, shard, kernel) ncu captured
    shard_n = world if world > 1 else args.shard_of
    math_key = "ffma2" if (args.cuda_cores or args.tma or sp.dtype == "f32") else "mma"
    traffic_key = f"{args.workload}/shard{shard_n}/{math_key}/{args.kernel}" + (("/share" if args.share_prefix in ("auto", "on") else "/share-" + args.share_prefix) if share else "")
    roof["traffic_key"] = traffic_key
    # a read-only stream through a library reduction over the same K cache, timed the same way: what a
    # kernel that only reads can reach on this box (the copy-measured peak is not a ceiling for it)
    try:
        flat = w["k_cache"].view(-1)
        flat = flat[:(flat.numel() // 8) * 8].view(torch.int64)
        for _ in range(2):
            flat.sum()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        r0.record()
        for _ in range(5):
            flat.sum()
        r1.record()
        torch.cuda.synchronize()
        roof["read_only_stream_gbs"] = flat.numel() * 8 / (r0.elapsed_time(r1) / 5 * 1e-3) / 1e9
        roof["read_only_stream_how"] = "torch.sum over the K cache viewed as int64 (5 launches, CUDA events)"
    except Exception as exc:
        roof["read_only_stream_gbs"] = None
        roof["read_only_stream_how"] = repr(exc)[:120]
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            with open(prof) as f:
                roof["traffic"] = json.load(f).get(traffic_key)
        except Exception:
            pass

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        s = cpu_sample(t, args.cpu_seqs)
        per_step, done = time_cpu(s, 50, 1, budget_s=15.0)
        cpu = {"value": s["n"] / per_step, "unit": "tokens/s", "cores": torch.get_num_threads(),
               "kind": "port", "ms_per_sample": per_step * 1e3,
               "sample": f"{s['n']} of {B} sequences (evenly strided, {s['tokens']} context tokens), "
                         f"{done} repetitions, oracle port of decode_torch_optimized in {sp.dtype}"}

    line = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": K,
        "warmup": warm, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": sp.dtype, "data": "synthetic",
        "config": workload_config(args, t, world),
        "implementation": {"chunk_tokens": int(chunk), "gather": gather_kind, "prefix_sharing": bool(share),
                           "kv_load": "tensor-map TMA" if args.tma else "cp.async", "kernel": args.kernel,
                           "peer_wait_gave_up_on_rank_plus_1": wait_timeouts,
                           "launch": "plain stream order" if args.no_overlap else
                           "programmatic dependent launch between steps (MVL_DECODE_OVERLAP_PREV)",
                           "math": "FFMA2 (CUDA cores)" if (args.cuda_cores or args.tma or sp.dtype == "f32")
                           else "mma.sync m16n8k16 (tensor cores)"},
        "hbm_gbs_step": alg_unique / (ms_step * 1e-3) / 1e9,
        "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "parity": parity, "sustained": sustained,
        "gpu_launches": K * (((3 if multi_split else 2) if args.unfused else 1) + (3 if wide_share else 2 if share else 0)),
        "clocks": clocks.summary(),
    }
    if world == 1 and args.workload == "c3" and args.shard_of == 1 and not args.no_neighbours:
        try:
            line["reference_triton"] = reference_triton_decode(w, alg_unique, ms_step)
        except Exception as exc:
            line["reference_triton"] = {"unavailable": repr(exc)[:200]}
        try:
            line["neighbours"] = neighbour_kernels(dev)
        except Exception as exc:   # the headline line must not depend on them
            line["neighbours"] = {"error": repr(exc)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_engine(args):
    """BASELINE.json configs[1] at the attention-path level: Qwen3-0.6B shape (16Q/8KV/D128, 28
    layers, bf16), batch 64, context 2k growing by one token per step, block_size 256 (main.py:19),
    driven like the reference engine: pool.append -> prepare_decode (host prep + H2D) -> graph replay
    of 28 x (store, split, combine).  tokens/s = batch / step time (llm_engine.py:58,79-83)."""
    from minivllm_b200.engine import BlockPool, DecodeRunner, SeqState

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    Lyr, H, KVH, D, BS, B, CTX = 28, 16, 8, 128, 256, 64, 2048
    warm = max(args.warmup, 3)
    K = args.steps
    max_len = ((CTX + warm + K + 8 + BS - 1) // BS) * BS
    nb = B * (max_len // BS)
    runner = DecodeRunner(Lyr, H, KVH, D, BS, nb, B, max_len, torch.bfloat16, dev, enforce_eager=args.eager)
    runner.kv_cache.normal_()
    runner.q.normal_(), runner.k.normal_(), runner.v.normal_()
    if not args.eager:
        runner.capture_cudagraph([B])
    pool = BlockPool(nb, BS, seed=2)
    seqs = [SeqState() for _ in range(B)]
    for s in seqs:
        pool.allocate(s, CTX)

    def step():
        for s in seqs:
            pool.append_token(s)
        runner.prepare_decode(seqs)
        return runner.run()

    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx_start = seqs[0].num_tokens
    with ClockSampler(0) as clocks:
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(K):
            step()
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    ms_step = e0.elapsed_time(e1) / K
    ctx_mean = (ctx_start + seqs[0].num_tokens) / 2 + 0.5
    es = 2
    alg = Lyr * (B * ctx_mean * KVH * D * 2 * es + 2 * B * H * D * es + 4 * B * np.ceil(ctx_mean / BS) + 8 * B
                 + 4 * B * KVH * D * es + 8 * B)
    peak, peak_src = measured_peaks()
    achieved = alg / (ms_step * 1e-3) / 1e9
    value = B / (ms_step * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": 1, "steps": K, "warmup": warm,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "c2e: 28-layer engine-shaped decode loop, attention path only "
                               "(projections/MLP/sampler out of scope)", "shape": "16Q/8KV/D128",
                   "layers": Lyr, "batch": B, "block_size": BS, "context": [ctx_start, seqs[0].num_tokens],
                   "cuda_graph": not args.eager,
                   "l2": "KV bytes per step (15 GB) >> L2; no flush needed"},
        "roofline": {"bound": "hbm", "kernel": "whole step: CUDA graph of 28 x (store_kvcache, "
                     "paged_decode_split, paged_decode_combine) + host prep", "achieved": achieved,
                     "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": alg},
        "cpu_baseline": None,
        "e2e": {"value": B / (wall / K), "unit": "tokens/s", "h2d_bytes_per_step": int(runner.h2d_bytes_per_step),
                "d2h_bytes_per_step": 0, "api": "DecodeRunner.prepare_decode + run (wall clock, "
                "metadata H2D every step; activations stay on the device as in the reference engine)"},
        "gpu_launches": K * Lyr * 3, "clocks": clocks.summary(),
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "c2e":
        if args.impl == "reference":
            args.workload = "c2"   # the CPU arm times one layer of the same shape
        elif world == 1:
            run_engine(args)
            return
        else:
            raise SystemExit("c2e is a single-GPU workload")
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if args.gpus != 1 or world != 1:
            raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={world})")
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
