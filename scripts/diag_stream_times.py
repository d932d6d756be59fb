"""Per-CTA timeline of the persistent stream kernel on one head shard (mvl_debug_set_times)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import workloads  # noqa: E402
from minivllm_b200 import _ext  # noqa: E402
from minivllm_b200.layers.attention import paged_attention_decode  # noqa: E402

name, shard = sys.argv[1], int(sys.argv[2])
dev = torch.device("cuda", 0)
t = workloads.build_tables(name)
from minivllm_b200.head_shard import shard_plan  # noqa: E402
sh = shard_plan(t.spec.H, t.spec.KVH, shard, 0)
w = workloads.materialize(t, dev, kv_heads=sh.kv_heads)
args = (w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"], w["KVH"],
        w["D"], w["BS"])
flags = _ext.DECODE_FUSED_COMBINE | _ext.DECODE_STREAM
for _ in range(5):
    paged_attention_decode(*args, flags=flags)
buf = torch.zeros(592 * 5, dtype=torch.int64, device=dev)
_ext.lib().mvl_debug_set_times(buf.data_ptr())
torch.cuda.synchronize()
paged_attention_decode(*args, flags=flags)
torch.cuda.synchronize()
_ext.lib().mvl_debug_set_times(None)
raw = buf.cpu().numpy().reshape(592, 5)
raw = raw[raw[:, 0] > 0]
sm = raw[:, 4]
ts = raw[:, :4].astype(np.float64)
t0 = ts[:, 0].min()
rel = (ts - t0) / 1e3  # us
out = {"workload": name, "shard": shard, "ctas": int(len(ts))}
for i, lab in enumerate(["entry", "piece_known", "first_tile", "exit"]):
    col = rel[:, i]
    out[lab] = {"min": round(float(col.min()), 2), "p10": round(float(np.percentile(col, 10)), 2),
                "median": round(float(np.median(col)), 2), "p90": round(float(np.percentile(col, 90)), 2),
                "max": round(float(col.max()), 2)}
out["cta_busy_us_median"] = round(float(np.median(rel[:, 3] - rel[:, 0])), 2)
# does a CTA's speed depend on where it ran?  busy time per SM (mean of its CTAs), by SM id range
busy = rel[:, 3] - rel[:, 2]
per_sm = {int(i): float(busy[sm == i].mean()) for i in np.unique(sm)}
ids = sorted(per_sm)
out["busy_by_sm_octile"] = [round(float(np.mean([per_sm[i] for i in ids[k * len(ids) // 8:(k + 1) * len(ids) // 8]])), 1)
                            for k in range(8)]
out["busy_sm_min_max"] = [round(min(per_sm.values()), 1), round(max(per_sm.values()), 1)]
within = [float(busy[sm == i].max() - busy[sm == i].min()) for i in ids if (sm == i).sum() > 1]
out["spread_within_sm_median"] = round(float(np.median(within)), 2)
out["busy_per_sm"] = [round(per_sm[i], 1) for i in ids]
print(json.dumps(out))
