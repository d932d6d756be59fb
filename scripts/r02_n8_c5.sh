#!/bin/bash
# Eight GPUs, C5: prefix sharing on every rank (one kv head each: tcgen05 group kernel + split kernel), result rows
# stored into all ranks' buffers, completion protocol.  One run, bounded.
N=${1:-8}
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus $N --workload c5 --steps 100 --warmup 10 --no-cpu-baseline --sustain-s 0 > gpurun_out/r02_bench_n${N}_c5_grid.json 2> gpurun_out/r02_bench_n${N}_c5.err
python - <<PY
import json
d = json.loads(open("gpurun_out/r02_bench_n${N}_c5_grid.json").read().strip().splitlines()[-1])
print("n", d["n_gpus"], "ms", round(d["ms_per_step"], 4), "tok/s", round(d["value"]), "kernel_ms", round(d["roofline"]["kernel_ms"], 4), "e2e", d["e2e"]["ms_per_step"] if d.get("e2e") else None,
      "parity", d["parity"]["max_abs_err"], d["parity"]["gather_verified"], "gave_up", d["implementation"].get("peer_wait_gave_up_on_rank_plus_1"))
PY
tail -2 gpurun_out/r02_bench_n${N}_c5.err
