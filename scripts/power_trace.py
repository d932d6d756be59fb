"""Step-time, SM-clock and power trajectory of a long C3 run (is the sustained regime power-capped?)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, pynvml
import workloads
from minivllm_b200.layers.attention import decode_step
dev = torch.device("cuda", 0)
t = workloads.build_tables(sys.argv[1] if len(sys.argv) > 1 else "c3")
w = workloads.materialize(t, dev)
out = torch.empty_like(w["q"])
step = lambda: decode_step(w["q"], w["k"], w["v"], w["k_cache"], w["v_cache"], w["slot_mapping"], w["block_tables"],
                           w["context_lens"], w["scale"], w["H"], w["KVH"], w["D"], w["BS"], out=out)
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
samples, stop = [], threading.Event()
def loop():
    while not stop.is_set():
        samples.append((time.perf_counter(), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                        pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM),
                        pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
                        pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)))
        stop.wait(0.02)
for _ in range(10): step()
torch.cuda.synchronize()
K = 1200
ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
th = threading.Thread(target=loop); th.start()
t0 = time.perf_counter()
ev[0].record()
for i in range(K):
    step(); ev[i + 1].record()
torch.cuda.synchronize()
stop.set(); th.join()
ms = np.array([ev[i].elapsed_time(ev[i + 1]) for i in range(K)])
for a in range(0, K, 100):
    print(f"steps {a:4d}-{a+99:4d}: median {np.median(ms[a:a+100]):.3f} ms  min {ms[a:a+100].min():.3f}")
for s in samples[::4]:
    print(f"t={s[0]-t0:6.3f}s sm {s[1]} MHz mem {s[2]} MHz power {s[3]:.0f} W reasons {s[4]:#x}")
