"""Host-side cost of one decode_step call at a tiny shape (C1): where the Python wrapper spends its time."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import workloads
from minivllm_b200.layers import attention as att
dev = torch.device("cuda", 0)
t = workloads.build_tables("c1")
w = workloads.materialize(t, dev)
sp = t.spec
args = (w["q"], w["k"], w["v"], w["k_cache"], w["v_cache"], w["slot_mapping"], w["block_tables"], w["context_lens"],
        w["scale"], sp.H, sp.KVH, sp.D, sp.BS)
out = torch.empty_like(w["q"])
for _ in range(50):
    att.decode_step(*args, out=out)
torch.cuda.synchronize()
n = 5000
t0 = time.perf_counter()
for _ in range(n):
    att.decode_step(*args, out=out)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host time per call {1e6 * (t1 - t0) / n:.1f} us; with drain {1e6 * (t2 - t0) / n:.1f} us")
pr = cProfile.Profile()
pr.enable()
for _ in range(2000):
    att.decode_step(*args, out=out)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
