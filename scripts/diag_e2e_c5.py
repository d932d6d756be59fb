"""End-to-end host step of C5 (prefix sharing): with and without the zero-copy q / result rows / new k, v rows,
and with sharing off, next to the device-resident step (CUDA events, 30 repetitions each)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import workloads  # noqa: E402
from minivllm_b200 import _ext  # noqa: E402
from minivllm_b200.layers.attention import HostDecodeStep  # noqa: E402

dev = torch.device("cuda", 0)
t = workloads.build_tables("c5")
w = workloads.materialize(t, dev)
B = w["B"]


def timed(fn, n=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


res = {}
share = _ext.DECODE_SHARE_PREFIX
for name, kw in (("share_direct_all", dict(flags=share)),
                 ("share_direct_none", dict(flags=share, direct_out=False, direct_q=False, direct_kv=False)),
                 ("share_direct_out_only", dict(flags=share, direct_q=False, direct_kv=False)),
                 ("share_direct_q_only", dict(flags=share, direct_out=False, direct_kv=False)),
                 ("share_int64_index", dict(flags=share, index_dtype=torch.int64)),
                 ("share_overlap_prev", dict(flags=share | _ext.DECODE_OVERLAP_PREV)),
                 ("share_overlap_prev_int64", dict(flags=share | _ext.DECODE_OVERLAP_PREV, index_dtype=torch.int64)),
                 ("unshared_direct_all", dict(flags=0))):
    hs = HostDecodeStep(w["k_cache"], w["v_cache"], B, t.tables.shape[1], w["H"], w["scale"], pieces=1, **kw)
    hs.q_h.copy_(w["q"].cpu()), hs.k_h.copy_(w["k"].cpu()), hs.v_h.copy_(w["v"].cpu())
    hs.slot_h.copy_(w["slot_mapping"].cpu()), hs.ctx_h.copy_(w["context_lens"].cpu())
    hs.bt_h.copy_(w["block_tables"].cpu())
    res[name + "_ms"] = round(timed(lambda: hs()), 4)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(30):
        hs(synchronize=False)
    res[name + "_host_issue_us"] = round((time.perf_counter() - t0) / 30 * 1e6, 1)
    torch.cuda.synchronize()
    if name == "share_direct_all":
        res["h2d_ms"] = round(timed(lambda: (hs._d_arena.copy_(hs._h_arena, non_blocking=True), torch.cuda.synchronize())), 4)
        res["h2d_bytes"] = hs.h2d_bytes
print(json.dumps(res))
