#!/bin/bash
# Does a TMA load path help the start-up-bound shapes?  The packed-fp32 kernel has both load paths in one
# template (cp.async / cp.async.bulk.tensor): same kernel, same shapes, only the copy instruction differs.
out=gpurun_out/r02_sweep_tma_startup.jsonl
: > $out
run() { python bench.py --no-cpu-baseline --no-neighbours --no-e2e --steps 100 --warmup 5 --sustain-s 0 --parity-seqs 0 "$@" >> $out 2>> gpurun_out/r02_sweep_tma_startup.err; }
for f in "--cuda-cores" "--tma"; do
  run --workload c3 --shard-of 8 $f
  run --workload c2 $f
done
python - <<'PY'
import json
for l in open("gpurun_out/r02_sweep_tma_startup.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"], d["config"]["parallelism"][:24], d["implementation"]["kv_load"], "ms", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3))
PY
