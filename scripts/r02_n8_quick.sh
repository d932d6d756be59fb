#!/bin/bash
N=${1:-8}
run() { name=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N "$@" > gpurun_out/r02_bench_n${N}_$name.json 2>> gpurun_out/r02_bench_n${N}.err; }
run c3_grid --workload c3 --steps 200 --warmup 10
run c3_short --workload c3 --steps 20 --warmup 5
run c3_nopdl --workload c3 --steps 200 --warmup 10 --no-overlap --no-e2e
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_bench_n${N}_c3_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f.split("/")[-1], "ms", round(d["ms_per_step"], 4), "tok/s", round(d["value"]), "kernel_ms", round(d["roofline"]["kernel_ms"], 4),
          "e2e_ms", round(d["e2e"]["ms_per_step"], 4) if d.get("e2e") else None, "parity", d["parity"]["max_abs_err"], d["parity"]["gather_verified"],
          "sust", round(d["sustained"]["ms_per_step"], 4), "gave_up", d["implementation"].get("peer_wait_gave_up_on_rank_plus_1"))
PY
tail -3 gpurun_out/r02_bench_n${N}.err
