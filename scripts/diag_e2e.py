"""Where the end-to-end step's time goes: copies alone, kernel alone, the serial host step, and the
overlapped host step at several piece counts (CUDA events, 20 repetitions each)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import workloads  # noqa: E402
from minivllm_b200.layers.attention import HostDecodeStep  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
dev = torch.device("cuda", 0)
t = workloads.build_tables(name)
w = workloads.materialize(t, dev)
B = w["B"]


def timed(fn, n=20):
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


res = {"workload": name}
for pieces in (1, 2, 4, 8):
    hs = HostDecodeStep(w["k_cache"], w["v_cache"], B, t.tables.shape[1], w["H"], w["scale"], pieces=pieces)
    hs.q_h.copy_(w["q"].cpu()), hs.k_h.copy_(w["k"].cpu()), hs.v_h.copy_(w["v"].cpu())
    hs.slot_h.copy_(w["slot_mapping"].cpu()), hs.ctx_h.copy_(w["context_lens"].cpu())
    hs.bt_h.copy_(w["block_tables"].cpu())
    res[f"host_step_pieces{pieces}_ms"] = timed(lambda: hs())
    if pieces == 1:
        res["h2d_ms"] = timed(lambda: (hs._d_arena.copy_(hs._h_arena, non_blocking=True), torch.cuda.synchronize()))
        res["d2h_ms"] = timed(lambda: (hs.out_h.copy_(hs.out_d, non_blocking=True), torch.cuda.synchronize()))
        res["h2d_bytes"], res["d2h_bytes"] = hs.h2d_bytes, hs.d2h_bytes
        import time
        t0 = time.perf_counter()
        for _ in range(50):
            hs(synchronize=False)
        res["host_issue_us_pieces1"] = (time.perf_counter() - t0) / 50 * 1e6
        torch.cuda.synchronize()
    else:
        import time
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(50):
            hs(synchronize=False)
        res[f"host_issue_us_pieces{pieces}"] = (time.perf_counter() - t0) / 50 * 1e6
        torch.cuda.synchronize()
for name_, kw in (("direct_out", dict(direct_out=True)), ("direct_q", dict(direct_q=True)),
                  ("direct_out_q", dict(direct_out=True, direct_q=True))):
    hs = HostDecodeStep(w["k_cache"], w["v_cache"], B, t.tables.shape[1], w["H"], w["scale"], pieces=1, **kw)
    hs.q_h.copy_(w["q"].cpu()), hs.k_h.copy_(w["k"].cpu()), hs.v_h.copy_(w["v"].cpu())
    hs.slot_h.copy_(w["slot_mapping"].cpu()), hs.ctx_h.copy_(w["context_lens"].cpu())
    hs.bt_h.copy_(w["block_tables"].cpu())
    res[f"host_step_{name_}_ms"] = timed(lambda: hs())
    ref = HostDecodeStep(w["k_cache"], w["v_cache"], B, t.tables.shape[1], w["H"], w["scale"], pieces=1)
    for a in ("q_h", "k_h", "v_h", "slot_h", "ctx_h", "bt_h"):
        getattr(ref, a).copy_(getattr(hs, a))
    res[f"{name_}_bitwise_equal"] = bool(torch.equal(ref().clone(), hs().clone()))
print(json.dumps(res))
