#!/bin/bash
# 8 ranks: real-rank tests, then the head-sharded bench lines.
N=${1:-8}
python -m pytest tests/test_gpu_multi.py -m gpu -q -x -k "real_ranks" 2>&1 | tail -3
run() { name=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 200 --warmup 10 "$@" > gpurun_out/r02_bench_n${N}_$name.json 2>> gpurun_out/r02_bench_n${N}.err; }
run c3_grid --workload c3
if [ "$N" -le 2 ]; then
  run c3_stream --workload c3 --kernel stream
  run c3_grid_nopdl --workload c3 --no-overlap
fi
run c3_nccl --workload c3 --gather nccl
run c4_grid --workload c4
run c5_grid --workload c5
# the driver's own settings
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_n${N}_c3_short.json 2>> gpurun_out/r02_bench_n${N}.err
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_bench_n${N}_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f.split("/")[-1], "ms", round(d["ms_per_step"], 4), "tok/s", round(d["value"]), "kernel_ms", round(d["roofline"]["kernel_ms"], 4),
          "frac", round(d["roofline"]["frac"], 3), "e2e_ms", round(d["e2e"]["ms_per_step"], 4), "parity", d["parity"]["max_abs_err"], d["parity"]["gather_verified"],
          "sust", round(d["sustained"]["ms_per_step"], 4))
PY
tail -5 gpurun_out/r02_bench_n${N}.err
