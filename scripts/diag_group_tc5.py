"""Role timeline of the tcgen05 group kernel (CTA 0, first item): per 128-token tile, when the producer got
the stage / finished issuing, when the MMA thread issued QK and PV, when the softmax warp started / ended."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import workloads  # noqa: E402
from minivllm_b200 import _ext  # noqa: E402
from minivllm_b200.layers.attention import paged_attention_decode  # noqa: E402

dev = torch.device("cuda", 0)
t = workloads.build_tables("c5")
w = workloads.materialize(t, dev)
args = (w["q"], w["k_cache"], w["v_cache"], w["block_tables"], w["context_lens"], w["scale"], w["H"], w["KVH"],
        w["D"], w["BS"])
flags = _ext.DECODE_FUSED_COMBINE | _ext.DECODE_SHARE_PREFIX
for _ in range(3):
    paged_attention_decode(*args, flags=flags)
buf = torch.zeros(32 * 6 + 2960, dtype=torch.int64, device=dev)
_ext.lib().mvl_debug_set_times(buf.data_ptr())
torch.cuda.synchronize()
paged_attention_decode(*args, flags=flags)
torch.cuda.synchronize()
_ext.lib().mvl_debug_set_times(None)
ts = buf.cpu().numpy()[:192].reshape(32, 6).astype(np.float64)
t0 = ts[ts > 0].min()
rel = np.where(ts > 0, (ts - t0) / 1e3, -1)
print("tile  prod_got_stage prod_issued  mma_qk_issue mma_pv_issue  softmax_start softmax_end   (us)")
for j in range(32):
    print(j, " ".join(f"{v:9.2f}" for v in rel[j]))

it = buf.cpu().numpy()[192:192 + 2960].reshape(148, 4, 5)
print("\nper item (CTA, item): start, streamed, counted, merged (us from first stamp), last_mask")
for c in (0, 1, 7, 8, 60, 119, 120, 147):
    for k in range(4):
        if it[c, k, 0] == 0:
            continue
        r = [(it[c, k, q] - t0) / 1e3 for q in range(4)]
        print(c, k, " ".join(f"{v:8.1f}" for v in r), hex(int(it[c, k, 4])))
ends = [(it[c, k, 3] - t0) / 1e3 for c in range(148) for k in range(4) if it[c, k, 3] > 0]
starts = [(it[c, k, 0] - t0) / 1e3 for c in range(148) for k in range(4) if it[c, k, 0] > 0]
merge = [(it[c, k, 3] - it[c, k, 2]) / 1e3 for c in range(148) for k in range(4) if it[c, k, 3] > 0]
stream = [(it[c, k, 1] - it[c, k, 0]) / 1e3 for c in range(148) for k in range(4) if it[c, k, 3] > 0]
print("items", len(ends), "first start", min(starts), "last end", max(ends), "stream us: median", np.median(stream), "max", max(stream),
      "merge us: median", np.median(merge), "max", max(merge), "sum of merges > 5us:", sum(1 for m in merge if m > 5))
