#!/bin/bash
# Two GPUs: the real-rank tests, then C5 (prefix sharing on every rank, result rows stored into both ranks'
# buffers, completion protocol) and C3 with the driver's arguments.
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3 > $O/r02_pytest_gpu_multi_n2.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus 2 --workload c5 --steps 100 --warmup 10 --no-cpu-baseline 2> $O/r02_bench_n2_c5.err | tail -1 > $O/r02_bench_n2_c5_grid.json
timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 20 --warmup 5 2>> $O/r02_bench_n2_c5.err | tail -1 > $O/r02_bench_n2_c3_short.json
tail -2 $O/r02_pytest_gpu_multi_n2.txt
python - <<'PY'
import json
for f in ("r02_bench_n2_c5_grid", "r02_bench_n2_c3_short"):
    d = json.loads(open(f"gpurun_out/{f}.json").read())
    print(f, d["n_gpus"], round(d["ms_per_step"], 4), round(d["value"]), d["parity"], d["implementation"].get("peer_wait_gave_up_on_rank_plus_1"))
PY
