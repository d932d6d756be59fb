#!/bin/bash
# Chunk-grid kernel vs persistent stream kernel on every workload and on the head shards (one GPU).
out=gpurun_out/r02_sweep_stream.jsonl
: > $out
run() { python bench.py --no-cpu-baseline --no-neighbours --steps 50 --warmup 5 --sustain-s 0 "$@" >> $out 2>> gpurun_out/r02_sweep_stream.err; }
for k in grid stream; do
  for w in c1 c2 c3 c4 c5; do run --workload $w --kernel $k; done
  for n in 2 4 8; do run --workload c3 --shard-of $n --kernel $k; done
  run --workload c4 --shard-of 8 --kernel $k
  run --workload c5 --shard-of 8 --kernel $k --share-prefix off
done
python - <<'PY'
import json
for l in open("gpurun_out/r02_sweep_stream.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"], d["config"]["parallelism"][:22], d["implementation"]["kernel"], "ms", round(d["ms_per_step"], 4),
          "frac", round(d["roofline"]["frac"], 3), "e2e_ms", round(d["e2e"]["ms_per_step"], 4) if d.get("e2e") else None,
          "parity", d["parity"]["max_abs_err"] if d.get("parity") else None)
PY
