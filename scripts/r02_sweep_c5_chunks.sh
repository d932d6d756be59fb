#!/bin/bash
# Split size of the C5 step with prefix sharing (tcgen05 group kernel + split kernel), N = 1 and one shard of 8.
out=gpurun_out/r02_sweep_c5_chunks.jsonl
: > $out
run() { python bench.py --no-cpu-baseline --no-neighbours --no-e2e --steps 100 --warmup 10 --sustain-s 0 --parity-seqs 0 --workload c5 "$@" >> $out 2>> gpurun_out/r02_sweep_c5_chunks.err; }
for c in 512 1024 2048 4096 8192; do run --chunk $c; run --shard-of 8 --chunk $c; done
run; run --shard-of 8
python - <<'PY'
import json
for l in open("gpurun_out/r02_sweep_c5_chunks.jsonl"):
    d = json.loads(l)
    print(d["config"]["parallelism"], d["implementation"]["chunk_tokens"], round(d["ms_per_step"], 4), round(d["roofline"]["frac"], 3))
PY
