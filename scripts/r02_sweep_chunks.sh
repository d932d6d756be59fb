#!/bin/bash
# Split size (tokens per chunk CTA) of the chunk-grid kernel on head shards and small launches.
out=gpurun_out/r02_sweep_chunks.jsonl
: > $out
run() { python bench.py --no-cpu-baseline --no-neighbours --no-e2e --steps 100 --warmup 5 --sustain-s 0 --parity-seqs 0 --kernel grid "$@" >> $out 2>> gpurun_out/r02_sweep_chunks.err; }
for c in 128 256 512 1024 2048 4096; do
  for n in 8 4 2; do run --workload c3 --shard-of $n --chunk $c; done
  run --workload c2 --chunk $c
  run --workload c4 --shard-of 8 --chunk $c
  run --workload c5 --shard-of 8 --chunk $c
  run --workload c3 --chunk $c
done
python - <<'PY'
import json, collections
rows = [json.loads(l) for l in open("gpurun_out/r02_sweep_chunks.jsonl")]
tab = collections.OrderedDict()
for d in rows:
    key = (d["config"]["workload"], d["config"]["parallelism"][:24])
    tab.setdefault(key, []).append((d["implementation"]["chunk_tokens"], round(d["ms_per_step"], 4)))
for k, v in tab.items():
    print(k, v)
PY
