#!/bin/bash
# Effect of MVL_DECODE_OVERLAP_PREV (programmatic dependent launch between steps) per workload / shard.
out=gpurun_out/r02_sweep_pdl.jsonl
: > $out
run() { python bench.py --no-cpu-baseline --no-neighbours --no-e2e --steps 100 --warmup 5 --sustain-s 0 "$@" >> $out 2>> gpurun_out/r02_sweep_pdl.err; }
for o in "" "--no-overlap"; do
  for k in grid stream; do
    for w in c2 c3 c5; do run --workload $w --kernel $k $o; done
    for n in 4 8; do run --workload c3 --shard-of $n --kernel $k $o; done
    run --workload c4 --shard-of 8 --kernel $k $o
  done
done
python - <<'PY'
import json
for l in open("gpurun_out/r02_sweep_pdl.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"], d["config"]["parallelism"][:22], d["implementation"]["kernel"], d["implementation"]["launch"][:5], "ms", round(d["ms_per_step"], 4),
          "frac_step", round(d["hbm_gbs_step"] / d["roofline"]["peak"], 3), "parity", d["parity"]["max_abs_err"] if d.get("parity") else None)
PY
